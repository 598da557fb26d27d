"""mmflow-style module surface of the hot path (same names, constructor arguments, forward
signatures and state_dict keys as the reference), backed by the sm_100a kernels.

  CorrelationPyramid  models/decoder/raft_decoder.py:19-58
  CorrLookup          models/utils/corr_lookup.py:71-136
  ConvGRU             models/decoder/raft_decoder.py:168-253
  MotionEncoder       models/decoder/raft_decoder.py:61-166   (cuDNN; ranked "next")
  XHead               models/decoder/raft_decoder.py:256-294  (cuDNN; "next")
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import convtc as C
from . import ops


def _needs_grad(*tensors) -> bool:
    """True when autograd must record the call (training, BASELINE.json configs[4])."""
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


_ACT = {"ReLU": lambda: nn.ReLU(inplace=True), "Sigmoid": nn.Sigmoid, "Tanh": nn.Tanh}


def _norm(cfg, ch):
    cfg = dict(cfg)
    t = cfg.pop("type")
    cfg.pop("requires_grad", None)
    if t == "GN":
        return "gn", nn.GroupNorm(num_channels=ch, **cfg)
    if t == "BN":
        return "bn", nn.BatchNorm2d(ch, **cfg)
    if t == "IN":
        return "in", nn.InstanceNorm2d(ch, **cfg)
    raise KeyError(f"norm type {t}")


class ConvModule(nn.Module):
    """conv -> [norm] -> [activation] with mmcv.cnn.ConvModule's attribute names (``conv``,
    ``gn``/``bn``/``in``, ``activate``) so checkpoints load with strict=True."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0,
                 conv_cfg=None, norm_cfg=None, act_cfg=dict(type="ReLU")):
        super().__init__()
        if conv_cfg is not None and conv_cfg.get("type", "Conv2d") not in ("Conv2d", "Conv"):
            raise NotImplementedError(f"conv_cfg {conv_cfg}")
        self.with_norm = norm_cfg is not None
        self.with_activation = act_cfg is not None
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size, stride=stride,
                              padding=padding, bias=not self.with_norm)
        self.norm_name = None
        if self.with_norm:
            self.norm_name, norm = _norm(norm_cfg, out_channels)
            self.add_module(self.norm_name, norm)
        if self.with_activation:
            self.activate = _ACT[act_cfg["type"]]()

    def _small_cin_ok(self, x) -> bool:
        c = self.conv
        return (not self.with_norm and c.in_channels <= 2 and c.kernel_size in ((3, 3), (7, 7))
                and c.stride == (1, 1) and c.padding == (c.kernel_size[0] // 2,) * 2 and c.dilation == (1, 1)
                and c.groups == 1 and c.out_channels % 32 == 0 and c.out_channels <= 256
                and x.shape[-1] % 8 == 0
                and (not self.with_activation or isinstance(self.activate, nn.ReLU)))

    def forward_split(self, x: "C.Split", **kw):
        """Tensor-core form on split-plane activations (convtc.Split in; kwargs as convtc.conv: out=, out_f32=,
        want=, res=, f32_tf32=).  conv -> bias -> ReLU in one launch; a GroupNorm layer is the raw convolution
        followed by the one-launch GN+ReLU kernel, returned as the fp32 channels-last tensor."""
        relu = self.with_activation and isinstance(self.activate, nn.ReLU)
        if self.with_activation and not relu:
            raise NotImplementedError(f"ConvModule activation {type(self.activate).__name__} on the tensor-core path")
        if not self.with_norm:
            return C.conv(x, C.packed_conv(self.conv), relu=relu, **kw)
        if self.norm_name != "gn" or not (self.gn.affine and self.gn.num_channels == 4 * self.gn.num_groups):
            raise NotImplementedError(f"ConvModule norm {self.norm_name} on the tensor-core path")
        raw = C.conv(x, C.packed_conv(self.conv, with_bias=False), want="f32")
        if kw.get("want") == "split":           # the next layer is a tensor-core convolution: split planes only
            n, h, w, c = raw.shape
            return ops.group_norm4_act(raw, self.gn, relu=relu, split_out=C.alloc_split(n, h, w, c, raw.device))
        return ops.group_norm4_act(raw, self.gn, relu=relu)

    def forward_small_cin_split(self, x: torch.Tensor) -> "C.Split":
        """1-2 input channels (FMA kernel, csrc/small_conv.cu) with the result written as split-bf16 planes: the
        operand of the tensor-core convolution that follows it in flow_net / delta_flow_encoder / mask_encoder."""
        if not self._small_cin_ok(x):
            return C.to_split(self.forward(x).permute(0, 2, 3, 1))
        c = self.conv
        if x.shape[-2] % 16 == 0 and c.in_channels * c.kernel_size[1] <= 16:
            # tensor-core form: the kw x-neighbours of the 1-2 channels expanded to 16 channels, then a kh x 1
            # convolution (kh taps of one k-step) instead of kh * kw * c FMAs per output on the FP32 pipe
            return C.conv(C.expand_taps(x, c.kernel_size[1]), C.packed_small_cin(c), relu=self.with_activation)
        return ops.conv_small_cin(x, c, relu=self.with_activation, split_out=True)

    def forward(self, x):
        if _needs_grad(x, *self.parameters()):
            # training (BASELINE.json configs[4]): the nn.Modules under autograd supply the backward
            x = self.conv(x)
            if self.with_norm:
                x = getattr(self, self.norm_name)(x)
            if self.with_activation:
                x = self.activate(x)
            return x
        if not x.is_cuda or x.dtype != torch.float32 or x.dim() != 4:
            raise ops._lib.ScflowError("ConvModule: expected a CUDA fp32 NCHW tensor (scflow_b200 has no CPU path)")
        if self._small_cin_ok(x):
            # 1-2 input channels: FP32-FMA kernel with bias + ReLU fused (csrc/small_conv.cu)
            return ops.conv_small_cin(x, self.conv, relu=self.with_activation)
        # stand-alone call on an NCHW-shaped tensor: one layout/split pass in, channels-last fp32 out
        xs = C.to_split(x.permute(0, 2, 3, 1))
        y = self.forward_split(xs, want="f32") if not self.with_norm else self.forward_split(xs)
        return y[..., :self.conv.out_channels].permute(0, 3, 1, 2)


class CorrelationPyramid(nn.Module):
    """All-pairs correlation + avg-pool pyramid in one tcgen05 kernel (a1)."""

    def __init__(self, num_levels: int = 4, precision: str = "tf32") -> None:
        super().__init__()
        self.num_levels = num_levels
        self.precision = {"fp32": ops.PREC_FP32, "tf32": ops.PREC_TF32}[precision]

    def forward(self, feat1: torch.Tensor, feat2: torch.Tensor) -> Sequence[torch.Tensor]:
        if _needs_grad(feat1, feat2):
            from . import autograd
            return autograd.corr_pyramid(feat1, feat2, self.num_levels, self.precision)
        return ops.corr_pyramid(feat1, feat2, self.num_levels, self.precision)


class CorrLookup(nn.Module):
    """Radius-r bilinear window lookup (a2).  Only the configuration every shipped config uses
    is implemented (bilinear / zeros / align_corners=True / radius 4); others are rejected."""

    def __init__(self, radius: int = 4, mode: str = "bilinear", padding_mode: str = "zeros",
                 align_corners: bool = True, div_mode: str = "ieee") -> None:
        super().__init__()
        if mode != "bilinear" or padding_mode != "zeros" or not align_corners:
            raise NotImplementedError(
                f"CorrLookup(mode={mode}, padding_mode={padding_mode}, align_corners={align_corners})")
        self.r = radius
        self.mode, self.padding_mode, self.align_corners = mode, padding_mode, align_corners
        self.div_mode = {"ieee": ops.DIV_IEEE, "rcp": ops.DIV_RCP}[div_mode]

    def forward(self, corr_pyramid: Sequence[torch.Tensor], flow: torch.Tensor) -> torch.Tensor:
        if _needs_grad(flow):
            # every shipped config detaches the flow before the lookup (scflow_decoder.py:193-194,
            # detach_flow=True); the coordinate gradient of grid_sample is not implemented
            raise NotImplementedError("CorrLookup: gradient wrt the flow is not implemented; detach it")
        if _needs_grad(*corr_pyramid):
            from . import autograd
            return autograd.corr_lookup(corr_pyramid, flow, self.r, self.div_mode)
        return ops.corr_lookup(corr_pyramid, flow, self.r, self.div_mode)

    def forward_split(self, corr_pyramid: Sequence[torch.Tensor], flow: torch.Tensor) -> "C.Split":
        """Inference: the lookup written as split-bf16 planes, the operand format of MotionEncoder.corr_net[0]."""
        return ops.corr_lookup(corr_pyramid, flow, self.r, self.div_mode, split=True)


class ConvGRU(nn.Module):
    """SeqConv GRU (a3).  Parameters keep the reference names conv_{z,r,q}.{0,1}.conv.*"""
    _kernel = {"Conv": 3, "SeqConv": ((1, 5), (5, 1))}
    _padding = {"Conv": 1, "SeqConv": ((0, 2), (2, 0))}

    def __init__(self, h_channels: int, x_channels: int, net_type: str = "SeqConv") -> None:
        super().__init__()
        if net_type != "SeqConv":
            raise NotImplementedError("gru_type='Conv' is not implemented (no config uses it)")
        self.h_channels, self.x_channels = h_channels, x_channels
        mk = lambda k, p, act: ConvModule(h_channels + x_channels, h_channels, k, padding=p,
                                          act_cfg=dict(type=act))
        ks, ps = self._kernel[net_type], self._padding[net_type]
        self.conv_z = nn.ModuleList([mk(k, p, "Sigmoid") for k, p in zip(ks, ps)])
        self.conv_r = nn.ModuleList([mk(k, p, "Sigmoid") for k, p in zip(ks, ps)])
        self.conv_q = nn.ModuleList([mk(k, p, "Tanh") for k, p in zip(ks, ps)])

    def init_weights(self) -> None:
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.orthogonal_(m.weight)

    def forward(self, h: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        from . import gru
        return gru.conv_gru_forward(self, h, x)


class MotionEncoder(nn.Module):
    def __init__(self, num_levels: int = 4, radius: int = 4, net_type: str = "Basic",
                 conv_cfg=None, norm_cfg=None, act_cfg=None) -> None:
        super().__init__()
        if net_type != "Basic":
            raise NotImplementedError("net_type != 'Basic'")
        kw = dict(conv_cfg=conv_cfg, norm_cfg=norm_cfg, act_cfg=act_cfg)
        cin = num_levels * (2 * radius + 1) ** 2
        self.corr_net = nn.Sequential(ConvModule(cin, 256, 1, padding=0, **kw),
                                      ConvModule(256, 192, 3, padding=1, **kw))
        self.flow_net = nn.Sequential(ConvModule(2, 128, 7, padding=3, **kw),
                                      ConvModule(128, 64, 3, padding=1, **kw))
        self.out_channels = [126]
        self.out_net = nn.Sequential(ConvModule(192 + 64, 126, 3, padding=1, **kw))

    def features(self, corr: torch.Tensor, flow: torch.Tensor) -> torch.Tensor:
        """The 126 learned motion channels (raft_decoder.py:162-165), before `flow` is appended."""
        return self.out_net(torch.cat([self.corr_net(corr), self.flow_net(flow)], dim=1))

    def features_into(self, corr: torch.Tensor, flow: torch.Tensor, out: "C.Split") -> None:
        """Inference: motion features + flow (raft_decoder.py:162-166) written as split planes into `out`, a
        128-channel slice (126 motion channels, then the flow pair) of the ConvGRU's operand buffer.
        corr: (N,324,h,w)-shaped channels-last lookup result; flow: planar (N,2,h,w).
        Tensor-core / FMA launches only; what the reference concatenates is written in place: corr_net and
        flow_net store into the two channel slices of one split buffer that out_net reads."""
        if isinstance(corr, C.Split):          # CorrLookup.forward_split: channels in the kernel's lane-major order
            cs = corr
            (n, h, w), dev = cs.shape, cs.buf.device
            c0 = self.corr_net[0]
            if c0.with_norm or not (c0.with_activation and isinstance(c0.activate, nn.ReLU)):
                raise NotImplementedError("MotionEncoder.corr_net[0] must be conv + ReLU on the tensor-core path")
            a = C.conv(cs, C.packed_conv(c0.conv, tag="lookup_order", in_order=ops.lookup_split_order()), relu=True)
        else:
            n, _, h, w = corr.shape
            dev = corr.device
            a = self.corr_net[0].forward_split(C.to_split(corr.permute(0, 2, 3, 1)))
        cat = C.alloc_split(n, h, w, 256, dev)
        c1 = self.corr_net[1].conv.out_channels
        self.corr_net[1].forward_split(a, out=cat.slice(0, c1))
        f0 = self.flow_net[0].forward_small_cin_split(flow)             # 7x7 on 2 channels: FMA kernel, split planes out
        self.flow_net[1].forward_split(f0, out=cat.slice(c1, 256 - c1))
        co = self.out_net[0].conv.out_channels
        self.out_net[0].forward_split(cat, out=out.slice(0, co))        # writes co rounded up to 8 (zeros)
        C.split_planar(flow, C.Split(out.buf, 2, out.off + co))        # the flow pair over those zeros

    def forward(self, corr: torch.Tensor, flow: torch.Tensor) -> torch.Tensor:
        if _needs_grad(corr, flow, *self.parameters()):
            return torch.cat([self.features(corr, flow), flow], dim=1)
        n, _, h, w = corr.shape
        co = self.out_net[0].conv.out_channels
        buf = C.alloc_split(n, h, w, co + 2, corr.device)
        self.features_into(corr.contiguous(memory_format=torch.channels_last), flow.contiguous(), buf)
        return buf.float().permute(0, 3, 1, 2)


class XHead(nn.Module):
    def __init__(self, in_channels: int, feat_channels: Sequence[int], x_channels: int, x: str) -> None:
        super().__init__()
        layers = []
        for ch in feat_channels:
            layers.append(ConvModule(in_channels, ch, 3, padding=1))
            in_channels = ch
        self.layers = nn.Sequential(*layers)
        if x in ("flow", "tradeoff"):
            self.predict_layer = nn.Conv2d(feat_channels[-1], x_channels, kernel_size=3, padding=1)
        elif x == "mask":
            self.predict_layer = nn.Conv2d(feat_channels[-1], x_channels, kernel_size=1, padding=0)
        else:
            raise ValueError(f"x must be 'flow' or 'mask', but got {x}")

    def forward_split(self, hs: "C.Split") -> torch.Tensor:
        """Inference on the split-plane hidden state: 3x3 convs on the tensor cores, the 1-2 channel predictor
        folded into the last layer's epilogue (convtc.head_predict); planar (N, x_channels, h, w)."""
        y = hs
        for layer in self.layers[:-1]:
            y = layer.forward_split(y, want="split")
        last = self.layers[-1]
        if last.with_norm or not (last.with_activation and isinstance(last.activate, nn.ReLU)):
            raise NotImplementedError("XHead: the last feature layer must be conv + ReLU on the tensor-core path")
        return C.head_predict(y, C.packed_conv(last.conv), self.predict_layer, relu=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if _needs_grad(x, *self.parameters()):
            return self.predict_layer(self.layers(x))
        if not x.is_cuda or x.dtype != torch.float32:
            raise ops._lib.ScflowError("XHead: expected a CUDA fp32 tensor (scflow_b200 has no CPU path)")
        return self.forward_split(C.to_split(x.permute(0, 2, 3, 1)))
