"""RAFTEncoder 'Basic' (models/encoder/raft_encoder.py:13-314 with BasicBlock/ResLayer from
models/backbone/resnet.py:12-92,676-771).  Upstream of the hot path; stays PyTorch/cuDNN
(ranked "next" 8f-3) but is part of the timed end-to-end metric, so it lives
here with the reference's parameter names (conv1, in1|bn1, res_layer{1,2,3}.{0,1}.*, conv2)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .registry import register


# ------------------------------------------------------------------ fused inference path
# Convolutions stay on cuDNN (SURVEY 8f-3 "next") but run channels-last end to end, and every
# norm + ReLU (+ residual add) is two launches of our HBM-bound kernels (csrc/norm_act.cu)
# instead of the reference's bias add_ / batch_norm / relu_ / add / relu_ chain.
def _conv_cl(conv: nn.Conv2d, x: torch.Tensor, with_bias: bool = False) -> torch.Tensor:
    """cuDNN convolution on a channels-last (N,C,H,W)-shaped tensor; the filter is re-laid once."""
    w = conv.weight
    if not w.is_contiguous(memory_format=torch.channels_last):
        conv.weight.data = w.data.contiguous(memory_format=torch.channels_last)
        w = conv.weight
    y = F.conv2d(x, w, conv.bias if with_bias else None, conv.stride, conv.padding)
    return y.contiguous(memory_format=torch.channels_last)      # no-op when cuDNN kept the layout


def _s2d_enabled(conv: nn.Conv2d, x: torch.Tensor) -> bool:
    """The 7x7 stride-2 stem on 3 channels is the one layer cuDNN has no tensor-core kernel for
    (it runs a generic 3-channel implicit GEMM: 416 us per 64 crops, 7 % of the step).  Rewritten
    as the equivalent 4x4 stride-1 convolution on the 2x2 space-to-depth input (12 -> 16 channels)
    it takes cuDNN's regular TF32 path.  Only where TF32 convolutions are allowed (the mode the
    reference itself runs in on a GPU): the strict-fp32 parity mode keeps the filter as shipped."""
    import os
    return (os.environ.get("SCFLOW_STEM_S2D", "1") != "0" and torch.backends.cudnn.allow_tf32
            and conv.kernel_size == (7, 7) and conv.stride == (2, 2) and conv.padding == (3, 3)
            and x.shape[1] <= 4 and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0)


def _s2d_filter(w: torch.Tensor) -> torch.Tensor:
    """(O,C,7,7) stride-2 filter -> (O,16,4,4) stride-1 filter over pixel_unshuffle(x, 2): tap
    k' = k + 1 = 2a + b lands on phase b at offset a - 2 (taps padded to 8 with a leading zero)."""
    o, c = w.shape[:2]
    wp = F.pad(w, (1, 0, 1, 0)).reshape(o, c, 4, 2, 4, 2).permute(0, 1, 3, 5, 2, 4).reshape(o, 4 * c, 4, 4)
    return F.pad(wp, (0, 0, 0, 0, 0, 16 - 4 * c)).contiguous(memory_format=torch.channels_last)


_S2D_LAST = [None, None, None]     # (weakref to the image tensor, its version, transformed tensor)


def _s2d_input(x: torch.Tensor) -> torch.Tensor:
    """(N,C,H,W) -> channels-last (N,16,H/2+3,W/2+3): phases as channels (c*4 + 2*by + bx), zero
    channels up to 16, and the stride-1 filter's (2,1) zero border written in -- one launch
    (scflow_stem_s2d_fwd).  The feature and the context encoder are fed the same rendered images
    (scflow_refiner.py:84-106): the last transform is kept and reused for the same, unmodified tensor."""
    import weakref
    ref, ver, val = _S2D_LAST
    if ref is not None and ref() is x and ver == x._version and val.device == x.device:
        return val
    val = ops.stem_s2d(x)
    _S2D_LAST[:] = [weakref.ref(x), x._version, val]
    return val


def _stem_cached(conv: nn.Conv2d, w: torch.Tensor) -> torch.Tensor:
    key = (w.data_ptr(), w._version)
    hit = conv.__dict__.get("_scflow_s2d")
    if hit is not None and hit[0] == key:
        return hit[1]
    with torch.no_grad():
        val = _s2d_filter(w.detach().float())
    conv.__dict__["_scflow_s2d"] = (key, val)
    return val


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    return x.permute(0, 2, 3, 1)


def _bn_affine(bn: nn.BatchNorm2d, bias):
    """Eval BatchNorm2d folded with the preceding convolution bias: y = conv*scale + shift."""
    key = tuple((t.data_ptr(), t._version) for t in (bn.weight, bn.bias, bn.running_mean, bn.running_var, bias)
                if t is not None)
    hit = bn.__dict__.get("_scflow_affine")
    if hit is not None and hit[0] == key:
        return hit[1]
    with torch.no_grad():
        scale = bn.weight.float() * torch.rsqrt(bn.running_var.float() + bn.eps)
        b = bias.float() if bias is not None else 0.0
        shift = bn.bias.float() + (b - bn.running_mean.float()) * scale
        val = (scale.contiguous(), shift.contiguous())
    bn.__dict__["_scflow_affine"] = (key, val)
    return val


def _bn_folded(conv: nn.Conv2d, bn: nn.BatchNorm2d):
    """Eval BatchNorm2d folded INTO the preceding convolution: (w * scale[cout], shift) as a
    channels-last filter + bias, so cuDNN's fused conv+bias(+residual)+ReLU kernels produce the
    block's output in one launch.  Cached on the conv, keyed on every tensor involved."""
    key = tuple((t.data_ptr(), t._version) for t in (conv.weight, conv.bias, bn.weight, bn.bias,
                                                     bn.running_mean, bn.running_var) if t is not None)
    hit = conv.__dict__.get("_scflow_folded")
    if hit is not None and hit[0] == key:
        return hit[1]
    scale, shift = _bn_affine(bn, conv.bias)
    with torch.no_grad():
        w = (conv.weight.float() * scale.view(-1, 1, 1, 1)).contiguous(memory_format=torch.channels_last)
    conv.__dict__["_scflow_folded"] = (key, (w, shift))
    return w, shift


def _side(norm: nn.Module, y: torch.Tensor, bias):
    """Normalisation coefficients of conv output y (channels-last view): kwargs for ops.norm_act."""
    if isinstance(norm, nn.InstanceNorm2d):
        # the convolution bias cancels in (v - mean): it is never added
        if norm.affine or norm.track_running_stats:
            raise NotImplementedError("InstanceNorm2d(affine/track_running_stats) on the fused path")
        return dict(sums=ops.chan_stats(y)), norm.eps
    if isinstance(norm, nn.BatchNorm2d):
        return dict(affine=_bn_affine(norm, bias)), norm.eps
    raise NotImplementedError(type(norm))


def _in_fused(norm: nn.Module) -> bool:
    """Plain InstanceNorm2d layers take the one-launch cluster kernel (SCFLOW_IN_FUSED=0: the
    statistics + normalise launch pair, kept as the cross-check)."""
    import os
    return (isinstance(norm, nn.InstanceNorm2d) and not norm.affine and not norm.track_running_stats
            and os.environ.get("SCFLOW_IN_FUSED", "1") != "0")


def _norm(kind: str, ch: int, postfix):
    if kind == "IN":
        return f"in{postfix}", nn.InstanceNorm2d(ch)
    if kind == "BN":
        return f"bn{postfix}", nn.BatchNorm2d(ch)
    raise NotImplementedError(f"norm {kind}")


class BasicBlock(nn.Module):
    def __init__(self, inplanes, planes, stride, norm, downsample=None):
        super().__init__()
        self.norm1_name, n1 = _norm(norm, planes, 1)
        self.norm2_name, n2 = _norm(norm, planes, 2)
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride=stride, padding=1, bias=True)
        self.add_module(self.norm1_name, n1)
        self.conv2 = nn.Conv2d(planes, planes, 3, padding=1, bias=True)
        self.add_module(self.norm2_name, n2)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        identity = x
        out = self.relu(getattr(self, self.norm1_name)(self.conv1(x)))
        out = getattr(self, self.norm2_name)(self.conv2(out))
        if self.downsample is not None:
            identity = self.downsample(x)
        return self.relu(out + identity)

    def forward_fused(self, x: torch.Tensor) -> torch.Tensor:
        """x, result: channels-last (N,C,H,W)-shaped tensors."""
        n1, n2 = getattr(self, self.norm1_name), getattr(self, self.norm2_name)
        if isinstance(n1, nn.BatchNorm2d):
            # eval BN folded into the filters: three cuDNN launches per block, no separate norm pass
            w1, b1 = _bn_folded(self.conv1, n1)
            w2, b2 = _bn_folded(self.conv2, n2)
            o = torch.cudnn_convolution_relu(x, w1, b1, self.conv1.stride, self.conv1.padding, (1, 1), 1)
            idt = x
            if self.downsample is not None:
                wd, bd = _bn_folded(self.downsample[0], self.downsample[1])
                idt = F.conv2d(x, wd, bd, self.downsample[0].stride, self.downsample[0].padding)
                idt = idt.contiguous(memory_format=torch.channels_last)
            return torch.cudnn_convolution_add_relu(o, w2, idt, 1.0, b2, self.conv2.stride, self.conv2.padding,
                                                    (1, 1), 1)
        o = _conv_cl(self.conv1, x)
        if _in_fused(n1):
            ops.instnorm_act(_nhwc(o), relu=True, eps=n1.eps)
        else:
            kw, eps = _side(n1, _nhwc(o), self.conv1.bias)
            ops.norm_act(_nhwc(o), relu=True, eps=eps, **kw)
        o2 = _conv_cl(self.conv2, o)
        if self.downsample is None and _in_fused(n2):
            ops.instnorm_act(_nhwc(o2), res=_nhwc(x), relu=True, eps=n2.eps)
            return o2
        kw, eps = _side(n2, _nhwc(o2), self.conv2.bias)
        if self.downsample is not None:
            d = _conv_cl(self.downsample[0], x)
            kr, _ = _side(self.downsample[1], _nhwc(d), self.downsample[0].bias)
            kw.update({"res_" + k: v for k, v in kr.items()})
            ops.norm_act(_nhwc(o2), res=_nhwc(d), relu=True, eps=eps, **kw)
        else:
            ops.norm_act(_nhwc(o2), res=_nhwc(x), relu=True, eps=eps, **kw)
        return o2


def _res_layer(inplanes, planes, stride, norm):
    down = None
    if stride != 1 or inplanes != planes:
        down = nn.Sequential(nn.Conv2d(inplanes, planes, 1, stride=stride, bias=True),
                             _norm(norm, planes, "")[1])
    return nn.Sequential(BasicBlock(inplanes, planes, stride, norm, down),
                         BasicBlock(planes, planes, 1, norm))


@register
class RAFTEncoder(nn.Module):
    def __init__(self, in_channels: int, out_channels: int, scale: float = 1 / 8,
                 net_type: str = "Basic", norm_cfg: dict = dict(type="BN", requires_grad=True),
                 init_cfg=None, **unsupported):
        super().__init__()
        if net_type != "Basic" or scale != 1 / 8 or any(v for v in unsupported.values()):
            raise NotImplementedError(f"RAFTEncoder({net_type}, scale={scale}, {unsupported})")
        norm = norm_cfg["type"]
        self.in_channels, self.out_channels = in_channels, out_channels
        self.conv1 = nn.Conv2d(in_channels, 64, 7, stride=2, padding=3, bias=True)
        self.norm1_name, n1 = _norm(norm, 64, 1)
        self.add_module(self.norm1_name, n1)
        self.relu = nn.ReLU(inplace=True)
        self.res_layers = []
        inpl = 64
        for i, (planes, stride) in enumerate(((64, 1), (96, 2), (128, 2))):
            name = f"res_layer{i + 1}"
            self.add_module(name, _res_layer(inpl, planes, stride, norm))
            self.res_layers.append(name)
            inpl = planes
        self.conv2 = nn.Conv2d(128, out_channels, 1)

    def _fused_ok(self, x: torch.Tensor) -> bool:
        n1 = getattr(self, self.norm1_name)
        if torch.is_grad_enabled() or not x.is_cuda or x.dtype != torch.float32:
            return False
        return isinstance(n1, nn.InstanceNorm2d) or not self.training

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._fused_ok(x):
            return self.forward_fused(x)
        x = self.relu(getattr(self, self.norm1_name)(self.conv1(x)))
        for name in self.res_layers:
            x = getattr(self, name)(x)
        return self.conv2(x)

    def forward_fused(self, x: torch.Tensor) -> torch.Tensor:
        """Inference path: same values as forward() (the IN-cancelled conv bias aside, a rounding-
        level difference), channels-last activations, norms through csrc/norm_act.cu."""
        n1 = getattr(self, self.norm1_name)
        s2d = _s2d_enabled(self.conv1, x)
        x = _s2d_input(x) if s2d else x.contiguous(memory_format=torch.channels_last)
        if isinstance(n1, nn.BatchNorm2d):
            w1, b1 = _bn_folded(self.conv1, n1)
            if s2d:
                y = torch.cudnn_convolution_relu(x, _stem_cached(self.conv1, w1), b1, (1, 1), (0, 0), (1, 1), 1)
            else:
                y = torch.cudnn_convolution_relu(x, w1, b1, self.conv1.stride, self.conv1.padding, (1, 1), 1)
        else:
            if s2d:
                y = F.conv2d(x, _stem_cached(self.conv1, self.conv1.weight))
                y = y.contiguous(memory_format=torch.channels_last)
            else:
                y = _conv_cl(self.conv1, x)
            if _in_fused(n1):
                ops.instnorm_act(_nhwc(y), relu=True, eps=n1.eps)
            else:
                kw, eps = _side(n1, _nhwc(y), self.conv1.bias)
                ops.norm_act(_nhwc(y), relu=True, eps=eps, **kw)
        for name in self.res_layers:
            for block in getattr(self, name):
                y = block.forward_fused(y)
        return _conv_cl(self.conv2, y, with_bias=True)
