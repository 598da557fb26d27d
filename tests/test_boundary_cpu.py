"""CPU tier: the C-ABI library loads and exports every symbol include/scflow_b200.h declares,
the module surface mirrors the reference (constructor kwargs, state_dict keys/shapes), and
unsupported configurations / CPU tensors are rejected instead of falling back."""
import copy
import json
import os
import re

import pytest
import torch

import scflow_b200 as sb
from scflow_b200 import _lib, build
from scflow_b200.refiner import DECODER_CFG, SCFlowRefinerCore, decoder_cfg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "scflow_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(scflow_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 12
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name


def test_abi_version_and_strerror(lib):
    assert lib.scflow_abi_version() == _lib.ABI_VERSION == 7
    assert lib.scflow_strerror(0) == b"ok"
    assert b"not implemented" in lib.scflow_strerror(-2)


def test_null_and_shape_rejected_without_gpu(lib):
    # argument validation happens before any CUDA call, so it is checkable on the CPU box
    assert lib.scflow_corr_lookup_fwd(None, 4, None, None, None, 1, 32, 32, 4, 0, 1, None) == -4
    assert lib.scflow_norm_apply_split(*([None] * 3), 0, 0, None, None, None, 0, 0, None, 1, 16, 64, 1e-5, 1, None) == -4
    assert lib.scflow_nhwc_to_nchw(None, None, 1, 16, 64, 1, None) == -4
    assert lib.scflow_corr_pyramid_fwd(None, None, None, 4, 1, 256, 32, 32, 1, None, 0, None) == -4
    assert lib.scflow_upsample_fwd(None, None, None, 1, 1, 4, 4, 8, 1.0, None) == -4
    # entry points added in round 2
    assert lib.scflow_corr_pyramid_bwd(None, None, None, None, None, 1, 256, 32, 32, 1, None) == -4
    assert lib.scflow_tap_sum_fwd(None, 4, 20, None, None, 1, 32, 32, 3, 2, None) == -4
    assert lib.scflow_pose_compose_fwd(None, None, None, None, None, None, 4, 10.0, None) == -4
    assert lib.scflow_pose_compose_bwd(*([None] * 10), 4, 10.0, 0, None) == -4
    assert lib.scflow_gru_reset_fwd(None, None, None, 16, None) == -4
    assert lib.scflow_gru_blend_bwd(*([None] * 7), 16, None) == -4
    assert lib.scflow_expand_taps_fwd(None, None, 1, 2, 7, 32, 32, None) == -4


def test_decoder_state_dict_matches_reference(golden_dir):
    dec = sb.SCFlowDecoder(**copy.deepcopy(DECODER_CFG))
    ref = {k: tuple(v) for k, v in json.load(open(os.path.join(golden_dir, "decoder_state_shapes.json"))).items()}
    mine = {k: tuple(v.shape) for k, v in dec.state_dict().items()}
    assert mine == ref
    # reference attributes read from outside (scflow_refiner.py:51-53,150-158)
    assert (dec.iters, dec.h_channels, dec.cxt_channels) == (8, 128, 128)
    dec.iters = 4
    # zero-initialised pose head (pose_head.py:187-198)
    assert float(dec.pose_pred.rotation_pred.weight.abs().max()) == 0.0
    assert dec.pose_pred.rotation_pred.bias[:6].tolist() == [1, 0, 0, 0, 1, 0]


def test_encoder_state_dict_matches_reference(golden_dir):
    core = SCFlowRefinerCore()
    es = json.load(open(os.path.join(golden_dir, "encoder_state_shapes.json")))
    for key, mod in (("IN", core.encoder), ("BN", core.context)):
        assert {n: list(v.shape) for n, v in mod.state_dict().items()} == es[key]


def test_registry_builds_from_config_dict():
    dec = sb.MODELS.build(dict(type="SCFlowDecoder", **decoder_cfg()))
    assert isinstance(dec, sb.SCFlowDecoder)
    dec2 = sb.MODELS.build(dict(type=sb.SCFlowDecoder, **decoder_cfg(feat_size=(64, 64))))
    assert dec2.pose_pred.fc_layers[0][0].in_features == 128 * 8 * 8


def test_install_is_opt_in_and_overrides_a_foreign_registry():
    """Importing the package registers nothing outside it; `install(registry)` replaces the three reference class
    names in the registry it is given (the reference's `registry.MODELS`), force=True."""
    class Foreign:
        def __init__(self):
            self.mods = {"SCFlowDecoder": object, "Other": int}
        def register_module(self, name=None, force=False, module=None):
            key = name or module.__name__
            if key in self.mods and not force:
                raise KeyError(key)
            self.mods[key] = module
            return module
    reg = Foreign()
    names = sb.install(reg)
    assert set(names) >= {"SCFlowDecoder", "MultiClassPoseHead", "RAFTEncoder"}
    assert reg.mods["SCFlowDecoder"] is sb.SCFlowDecoder and reg.mods["Other"] is int


@pytest.mark.parametrize("over", [dict(net_type="Small"), dict(mask_corr=True), dict(mask_flow=True),
                                  dict(depth_transform="linear"), dict(gru_type="Conv"),
                                  dict(detach_flow=False)])
def test_unsupported_configs_rejected(over):
    with pytest.raises((NotImplementedError, AssertionError)):
        sb.SCFlowDecoder(**decoder_cfg(**over))


def test_lookup_modes_rejected():
    with pytest.raises(NotImplementedError):
        sb.CorrLookup(radius=4, align_corners=False)
    with pytest.raises(NotImplementedError):
        sb.CorrLookup(radius=4, padding_mode="border")


def test_cpu_tensors_rejected_no_fallback(lib):
    pyr = sb.CorrelationPyramid(4)
    with pytest.raises(_lib.ScflowError):
        pyr(torch.zeros(1, 32, 8, 32), torch.zeros(1, 32, 8, 32))
    with pytest.raises(_lib.ScflowError):
        sb.CorrLookup(4)([torch.zeros(256, 1, 8, 32)], torch.zeros(1, 2, 8, 32))
    with pytest.raises(_lib.ScflowError):
        sb.ConvGRU(128, 256)(torch.zeros(1, 128, 8, 8), torch.zeros(1, 256, 8, 8))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.ScflowError):
        _lib.load()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "scflow_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
