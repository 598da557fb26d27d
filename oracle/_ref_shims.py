"""Import the UNMODIFIED reference hot-path files from /root/reference with shims.

TEST INFRASTRUCTURE ONLY.  Used by ``oracle/gen_golden.py`` (run in the dev
container, where /root/reference is mounted) to pin the CPU restatement in
``oracle/scflow_oracle.py`` against the reference's own code.  Nothing under
``scflow_b200/`` may import this; nothing on the GPU box can (the mount is
absent there).

The reference is pure Python but depends on wheels that are not installed
(mmcv, mmengine, kornia, pytorch3d, tkinter).  Only the seven third-party names
the hot path touches are shimmed; the reference files
themselves are executed as they are from the read-only mount:

  models/decoder/raft_decoder.py, models/decoder/scflow_decoder.py,
  models/utils/corr_lookup.py, models/utils/pose.py, models/head/pose_head.py,
  models/encoder/raft_encoder.py, models/backbone/resnet.py
"""
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn

_STAGED = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
# the read-only mount in the dev container; on the GPU box the byte-identical copies staged by
# oracle/stage_ref.py under baseline/_ref/ (git-ignored, travels with the snapshot)
REF_ROOT = os.environ.get("SCFLOW_REFERENCE_ROOT") or (
    "/root/reference" if os.path.isfile("/root/reference/models/decoder/scflow_decoder.py") else _STAGED)


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "models/decoder/scflow_decoder.py"))


# --------------------------------------------------------------------------- shims
class _BaseModule(nn.Module):
    """mmengine.model.BaseModule: nn.Module that swallows ``init_cfg``."""

    def __init__(self, init_cfg=None):
        super().__init__()
        self.init_cfg = init_cfg

    def init_weights(self):
        pass


class _Sequential(_BaseModule, nn.Sequential):
    def __init__(self, *args, init_cfg=None):
        _BaseModule.__init__(self, init_cfg)
        nn.Sequential.__init__(self, *args)


_ACTS = {"ReLU": lambda: nn.ReLU(inplace=True), "Sigmoid": nn.Sigmoid, "Tanh": nn.Tanh}


def _build_norm_layer(cfg, num_features, postfix=""):
    cfg = dict(cfg)
    t = cfg.pop("type")
    cfg.pop("requires_grad", None)
    if t == "BN":
        return "bn" + str(postfix), nn.BatchNorm2d(num_features, **cfg)
    if t == "IN":
        return "in" + str(postfix), nn.InstanceNorm2d(num_features, **cfg)
    if t == "GN":
        return "gn" + str(postfix), nn.GroupNorm(num_channels=num_features, **cfg)
    raise KeyError(t)


def _build_conv_layer(cfg, *args, **kwargs):
    assert cfg is None or cfg.get("type", "Conv2d") in ("Conv2d", "Conv")
    return nn.Conv2d(*args, **kwargs)


class _ConvModule(nn.Module):
    """mmcv.cnn.ConvModule restricted to what the hot path uses:
    conv -> optional norm -> optional activation, bias iff no norm,
    attribute names ``conv`` / ``gn|bn|in`` / ``activate`` (infer.ipynb cell 3)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0,
                 dilation=1, groups=1, bias="auto", conv_cfg=None, norm_cfg=None,
                 act_cfg=dict(type="ReLU"), inplace=True, **_):
        super().__init__()
        self.with_norm = norm_cfg is not None
        self.with_activation = act_cfg is not None
        if bias == "auto":
            bias = not self.with_norm
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size, stride=stride,
                              padding=padding, dilation=dilation, groups=groups, bias=bias)
        self.norm_name = None
        if self.with_norm:
            self.norm_name, norm = _build_norm_layer(norm_cfg, out_channels)
            self.add_module(self.norm_name, norm)
        if self.with_activation:
            self.activate = _ACTS[act_cfg["type"]]()

    def forward(self, x):
        x = self.conv(x)
        if self.with_norm:
            x = getattr(self, self.norm_name)(x)
        if self.with_activation:
            x = self.activate(x)
        return x


class _Registry:
    def __init__(self):
        self._m = {}

    def register_module(self, name=None, force=False, module=None):
        def deco(cls):
            self._m[name or cls.__name__] = cls
            return cls
        return deco if module is None else deco(module)

    def build(self, cfg):
        cfg = dict(cfg)
        t = cfg.pop("type")
        cls = self._m[t] if isinstance(t, str) else t
        return cls(**cfg)


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _pkg(name, path):
    m = types.ModuleType(name)
    m.__path__ = [path]
    sys.modules[name] = m
    return m


def _load(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF_ROOT, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_LOADED = None


def load_reference():
    """Returns a namespace with the reference's own classes/functions."""
    global _LOADED
    if _LOADED is not None:
        return _LOADED
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")

    # (1) turtle.forward  (raft_decoder.py:3 imports it; tkinter is absent)
    _mod("turtle", forward=lambda *a, **k: None)
    # (2) mmengine.model, (3) mmcv.cnn
    _mod("mmengine")
    _mod("mmengine.model", BaseModule=_BaseModule, Sequential=_Sequential, BaseModel=_BaseModule)
    _mod("mmcv")
    _mod("mmcv.cnn", ConvModule=_ConvModule, build_conv_layer=_build_conv_layer,
         build_norm_layer=_build_norm_layer, build_plugin_layer=None)
    _mod("mmcv.utils")
    _mod("mmcv.utils.parrots_wrapper", _BatchNorm=nn.modules.batchnorm._BatchNorm)
    _mod("mmengine.utils")
    _mod("mmengine.utils.dl_utils")
    _mod("mmengine.utils.dl_utils.parrots_wrapper", _BatchNorm=nn.modules.batchnorm._BatchNorm)
    # (4) registry.MODELS
    reg = _Registry()
    _mod("registry", MODELS=reg)
    # (5) kornia (quaternion mode only), (6) datasets.pose.remap_pose (unused on the path)
    _mod("kornia")
    _mod("datasets")
    _mod("datasets.pose", remap_pose=None)
    if "cv2" not in sys.modules:
        try:
            import cv2  # noqa: F401
        except Exception:
            _mod("cv2")
    # (7) package stubs that bypass models/__init__.py -> pytorch3d
    for p in ("models", "models.utils", "models.decoder", "models.head",
              "models.encoder", "models.backbone"):
        _pkg(p, os.path.join(REF_ROOT, p.replace(".", "/")))

    corr_lookup = _load("models.utils.corr_lookup", "models/utils/corr_lookup.py")
    pose = _load("models.utils.pose", "models/utils/pose.py")
    mu = sys.modules["models.utils"]
    for n in ("get_flow_from_delta_pose_and_points", "get_pose_from_delta_pose",
              "cal_3d_2d_corr", "get_flow_from_delta_pose_and_depth", "lift_2d_to_3d"):
        setattr(mu, n, getattr(pose, n))
    mu.CorrLookup = corr_lookup.CorrLookup
    raft_decoder = _load("models.decoder.raft_decoder", "models/decoder/raft_decoder.py")
    pose_head = _load("models.head.pose_head", "models/head/pose_head.py")
    scflow_decoder = _load("models.decoder.scflow_decoder", "models/decoder/scflow_decoder.py")
    resnet = _load("models.backbone.resnet", "models/backbone/resnet.py")
    raft_encoder = _load("models.encoder.raft_encoder", "models/encoder/raft_encoder.py")

    _LOADED = types.SimpleNamespace(
        corr_lookup=corr_lookup, pose=pose, raft_decoder=raft_decoder, pose_head=pose_head,
        scflow_decoder=scflow_decoder, resnet=resnet, raft_encoder=raft_encoder, registry=reg,
        CorrelationPyramid=raft_decoder.CorrelationPyramid, CorrLookup=corr_lookup.CorrLookup,
        ConvGRU=raft_decoder.ConvGRU, MotionEncoder=raft_decoder.MotionEncoder,
        XHead=raft_decoder.XHead, SCFlowDecoder=scflow_decoder.SCFlowDecoder,
        MultiClassPoseHead=pose_head.MultiClassPoseHead, RAFTEncoder=raft_encoder.RAFTEncoder,
    )
    return _LOADED


DECODER_CFG = dict(
    net_type="Basic", num_levels=4, radius=4, iters=8,
    detach_flow=True, detach_mask=True, detach_pose=True, detach_depth_for_xy=True,
    mask_flow=False, mask_corr=False,
    pose_head_cfg=dict(type="MultiClassPoseHead", num_class=21, in_channels=224,
                       net_type="Basic", rotation_mode="ortho6d",
                       norm_cfg=dict(type="GN", num_groups=32, requires_grad=True),
                       act_cfg=dict(type="ReLU")),
    corr_lookup_cfg=dict(align_corners=True), gru_type="SeqConv", act_cfg=dict(type="ReLU"))


class ReferenceRefiner(nn.Module):
    """The part of SCFlowRefiner.get_pose the metric covers (models/refiner/scflow_refiner.py:84-138), assembled
    from the UNMODIFIED reference classes: shared IN RAFTEncoder on real and rendered images, BN context
    RAFTEncoder on the rendered images, tanh / relu split, SCFlowDecoder with a zero init flow and
    invalid_flow_num = 0.  (scflow_refiner.py itself needs the pytorch3d renderer to import.)"""

    def __init__(self, image_size: int = 256, state=None):
        super().__init__()
        import copy
        R = load_reference()
        self.encoder = R.RAFTEncoder(in_channels=3, out_channels=256, net_type="Basic", norm_cfg=dict(type="IN"))
        self.context = R.RAFTEncoder(in_channels=3, out_channels=256, net_type="Basic", norm_cfg=dict(type="BN"))
        cfg = copy.deepcopy(DECODER_CFG)
        cfg["pose_head_cfg"]["feat_size"] = (image_size // 8, image_size // 8)
        self.decoder = R.SCFlowDecoder(**cfg)
        if state is not None:
            self.encoder.load_state_dict(state[0])
            self.context.load_state_dict(state[1])
            self.decoder.load_state_dict(state[2])
        self.eval()

    @torch.no_grad()
    def forward(self, batch):
        real_feat = self.encoder(batch["real_images"])
        render_feat = self.encoder(batch["rendered_images"])
        cxt = self.context(batch["rendered_images"])
        h, c = torch.split(cxt, [128, 128], dim=1)
        n, _, hh, ww = batch["real_images"].shape
        init_flow = torch.zeros((n, 2, hh, ww), dtype=torch.float32, device=real_feat.device)
        out = self.decoder(render_feat, real_feat, torch.tanh(h), torch.relu(c), batch["R0"], batch["t0"],
                           batch["depth"], batch["K"], batch["labels"], init_flow, 0.)
        return out[2][-1], out[3][-1]
