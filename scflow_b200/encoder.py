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


# ------------------------------------------------------------------ tensor-core inference path
# Every convolution of the encoders runs on scflow_conv_tc_fwd (split-bf16, three tcgen05 MMAs per product:
# fp32-grade results, csrc/conv_tc.cu) with channels-last split-plane activations end to end:
#   * eval BatchNorm is folded into the packed filters, so a BN block is three launches (conv+bias+ReLU,
#     [1x1 down-sample], conv+bias+residual+ReLU) and nothing else;
#   * an InstanceNorm layer is the convolution (raw fp32 out, per-(image, channel) sum / sum of squares
#     accumulated by its epilogue) + ONE normalise(+residual)+ReLU pass that writes the next operand;
#   * the 7x7 stride-2 stem on 3 channels is the equivalent 4x4 stride-1 convolution on the 2x2 space-to-depth
#     image (12 -> 16 channels), read as 4 filter rows of 64 contiguous channels (four x-neighbours).
from . import convtc as C


def _s2d_filter(w: torch.Tensor) -> torch.Tensor:
    """(O,C,7,7) stride-2 filter -> (O,16,4,4) stride-1 filter over pixel_unshuffle(x, 2): tap
    k' = k + 1 = 2a + b lands on phase b at offset a - 2 (taps padded to 8 with a leading zero)."""
    o, c = w.shape[:2]
    wp = F.pad(w, (1, 0, 1, 0)).reshape(o, c, 4, 2, 4, 2).permute(0, 1, 3, 5, 2, 4).reshape(o, 4 * c, 4, 4)
    return F.pad(wp, (0, 0, 0, 0, 0, 16 - 4 * c))


class StemInput:
    """2x2 space-to-depth image of a planar (N,C<=4,H,W) batch as two bf16 planes (scflow_stem_s2d_split_fwd);
    computed once per image batch and shared by the encoders that read it (scflow_refiner.py:84-106 feeds the
    rendered images to both the feature and the context encoder)."""

    def __init__(self, x: torch.Tensor):
        n, c, h, w = x.shape
        if not x.is_cuda or x.dtype != torch.float32 or c > 4 or h % 2 or w % 2:
            raise ops._lib.ScflowError(f"encoder stem: expected a CUDA fp32 (N,<=4,even,even) batch, got "
                                       f"{tuple(x.shape)} {x.dtype} {x.device}")
        self.n, self.h2, self.w2 = n, h // 2, w // 2
        self.planes = ops.stem_s2d_split(x)                    # (2, n, h/2+3, w/2+3, 16) bf16

    def split(self) -> "C.Split":
        """The planes as a Split the conv launcher can address, plus the overlapped-window view."""
        _, n, ho, wo, _ = self.planes.shape
        buf = self.planes.view(2 * n, ho, wo, 16)               # only data_ptr / device are used with `view`
        view = (n, ho, wo - 3, 64, n * ho * wo * 16, 16, wo * 16, ho * wo * 16)
        return C.Split(buf, 64), view


def _stem_pack(conv: nn.Conv2d, bn=None) -> "C.Packed":
    ts = [conv.weight, conv.bias] + ([bn.weight, bn.bias, bn.running_mean, bn.running_var] if bn is not None else [])
    key = C._key(*ts)
    hit = conv.__dict__.get("_scflow_stem")
    if hit is not None and hit[0] == key:
        return hit[1]
    if conv.kernel_size != (7, 7) or conv.stride != (2, 2) or conv.padding != (3, 3):
        raise ops._lib.ScflowError("encoder stem: expected the 7x7 stride-2 pad-3 convolution")
    with torch.no_grad():
        w = _s2d_filter(conv.weight.detach().float())                       # (O, 16, dy, dx)
        w = w.permute(0, 3, 1, 2).reshape(w.shape[0], 64, 4, 1).contiguous()   # cin = dx*16 + ch, kh = dy
        if bn is not None:
            scale = bn.weight.float() * torch.rsqrt(bn.running_var.float() + bn.eps)
            b0 = conv.bias.float() if conv.bias is not None else 0.0
            shift = bn.bias.float() + (b0 - bn.running_mean.float()) * scale
            pk = C.pack_weights(w, shift, 1, 0, scale=scale)
        else:
            pk = C.pack_weights(w, None, 1, 0)                 # InstanceNorm cancels the bias
    conv.__dict__["_scflow_stem"] = (key, pk)
    return pk


class _Stats:
    """One zeroed fp64 buffer per encoder pass for the InstanceNorm statistics of all its layers."""

    def __init__(self, n: int, device, floats: int):
        self.buf = torch.zeros(floats, device=device, dtype=torch.float64)
        self.n, self.used = n, 0

    def take(self, c: int) -> torch.Tensor:
        k = self.n * C.STATS_REPLICAS * c * 2
        out = self.buf[self.used:self.used + k].view(self.n, C.STATS_REPLICAS, c, 2)
        self.used += k
        assert self.used <= self.buf.numel()
        return out


def _plain_in(norm: nn.Module) -> None:
    if not isinstance(norm, nn.InstanceNorm2d) or norm.affine or norm.track_running_stats:
        raise NotImplementedError(f"{type(norm).__name__}(affine/track_running_stats) on the tensor-core path")


def _conv_raw(x, conv: nn.Conv2d, st: _Stats, view=None, pk=None):
    """Convolution without bias (InstanceNorm cancels it) -> raw fp32 output + its statistics."""
    pk = pk or C.packed_conv(conv, with_bias=False, tag="in")
    sums = st.take(pk.cout)
    raw = C.conv(x, pk, want="f32", stats=sums, view=view)
    return raw, sums


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

    def forward_tc(self, x, st):
        """x, result: split-plane activations (convtc.Split); st: the pass's statistics pool."""
        n1, n2 = getattr(self, self.norm1_name), getattr(self, self.norm2_name)
        if isinstance(n1, nn.BatchNorm2d):
            o = C.conv(x, C.packed_conv(self.conv1, bn=n1, tag="bn"), relu=True)
            idt = x
            if self.downsample is not None:
                idt = C.conv(x, C.packed_conv(self.downsample[0], bn=self.downsample[1], tag="bn"))
            return C.conv(o, C.packed_conv(self.conv2, bn=n2, tag="bn"), relu=True, res=idt)
        _plain_in(n1), _plain_in(n2)
        raw, sums = _conv_raw(x, self.conv1, st)
        o = C.norm_apply(raw, sums, relu=True, eps=n1.eps)
        raw2, sums2 = _conv_raw(o, self.conv2, st)
        if self.downsample is None:
            return C.norm_apply(raw2, sums2, relu=True, eps=n2.eps, res=x)
        _plain_in(self.downsample[1])
        rawd, sumsd = _conv_raw(x, self.downsample[0], st)
        return C.norm_apply(raw2, sums2, relu=True, eps=n2.eps, res_raw=rawd, res_sums=sumsd)


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

    def forward(self, x: torch.Tensor, stem: "StemInput" = None, planar_tf32: bool = False,
                taps: list = None) -> torch.Tensor:
        """Inference (no autograd): the tensor-core path, CUDA only.  With autograd recording (training,
        BASELINE.json configs[4]) the layers run as the nn.Modules they are, so that autograd supplies the
        backward.  `stem`: a StemInput of x computed by the caller (shared between encoders);
        `planar_tf32`: return the features NCHW-contiguous and rounded to the nearest TF32 -- the correlation
        kernel's operand format -- instead of a channels-last tensor."""
        n1 = getattr(self, self.norm1_name)
        if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters())):
            x = self.relu(n1(self.conv1(x)))
            for name in self.res_layers:
                x = getattr(self, name)(x)
            return self.conv2(x)
        if isinstance(n1, nn.BatchNorm2d) and self.training:
            raise NotImplementedError("RAFTEncoder(BN) in train mode without autograd: no batch-statistics kernel")
        return self.forward_tc(x, stem, planar_tf32, taps)

    def forward_tc(self, x: torch.Tensor, stem=None, planar_tf32: bool = False, taps: list = None) -> torch.Tensor:
        """`taps` (tests): collects the stem output and every residual block's output as fp32 (n,h,w,c)."""
        n1 = getattr(self, self.norm1_name)
        stem = stem or StemInput(x)
        xs, view = stem.split()
        n = stem.n
        st = None
        if isinstance(n1, nn.BatchNorm2d):
            y = C.conv(xs, _stem_pack(self.conv1, n1), relu=True, view=view)
        else:
            _plain_in(n1)
            st = _Stats(n, x.device, n * C.STATS_REPLICAS * 2 * (64 * 5 + 96 * 5 + 128 * 5))
            raw, sums = _conv_raw(xs, self.conv1, st, view=view, pk=_stem_pack(self.conv1))
            y = C.norm_apply(raw, sums, relu=True, eps=n1.eps)
        if taps is not None:
            taps.append(y.float())
        for name in self.res_layers:
            for block in getattr(self, name):
                y = block.forward_tc(y, st)
                if taps is not None:
                    taps.append(y.float())
        pk = C.packed_conv(self.conv2)
        if planar_tf32:
            out = C.conv(y, pk, want="planar", planar_tf32=True)
            out._scflow_planar_tf32 = True          # ops.corr_pyramid takes it as is
            return out
        return C.conv(y, pk, want="f32").permute(0, 3, 1, 2)     # NCHW-shaped view of channels-last memory
