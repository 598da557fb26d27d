"""Host side of the tensor-core convolution family (csrc/conv_tc.cu, `scflow_conv_tc_fwd`).

Activations travel between the convolutions as SPLIT planes: a channels-last buffer whose pixel holds
[hi C | lo C] bf16 with value = hi + lo (16 significand bits).  `Split` describes such a buffer or a
channel slice of one (so that producers write straight into the slices of what the reference builds with
`torch.cat`).  Weights stay fp32 `nn.Parameter`s under the reference names; `pack()` derives the
[tap][k-block][hi|lo][cout][bk] bf16 copy (optionally with an eval BatchNorm folded in), cached on the
module and keyed on the parameters' version counters and storage, never saved.

There is no other backend behind these functions: CPU tensors and unsupported shapes raise.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .ops import _ptr, _stream, _timed


class _Desc(ctypes.Structure):
    """Mirror of scflow_conv_tc_desc (include/scflow_b200.h)."""
    _fields_ = [(k, ctypes.c_int) for k in ("n", "h", "w", "cin", "cout", "kh", "kw", "stride", "pad_h", "pad_w",
                                            "bk")] + [
        ("x", ctypes.c_void_p), ("x_lo_off", ctypes.c_longlong), ("x_stride_w", ctypes.c_longlong),
        ("x_stride_h", ctypes.c_longlong), ("x_stride_n", ctypes.c_longlong),
        ("wpack", ctypes.c_void_p), ("bias", ctypes.c_void_p), ("relu", ctypes.c_int),
        ("res_split", ctypes.c_void_p), ("res_pitch", ctypes.c_longlong), ("res_lo", ctypes.c_int),
        ("res_f32", ctypes.c_void_p), ("resf_pitch", ctypes.c_longlong),
        ("out_split", ctypes.c_void_p), ("out_pitch", ctypes.c_longlong), ("out_lo", ctypes.c_int),
        ("out_f32", ctypes.c_void_p), ("outf_pitch", ctypes.c_longlong), ("f32_tf32", ctypes.c_int),
        ("out_planar", ctypes.c_void_p), ("planar_tf32", ctypes.c_int),
        ("stats", ctypes.c_void_p), ("bn", ctypes.c_int), ("epi", ctypes.c_int), ("gru_c", ctypes.c_int),
        ("gru_h", ctypes.c_void_p), ("gru_z", ctypes.c_void_p), ("algo", ctypes.c_int),
        ("x2", ctypes.c_void_p), ("cin2", ctypes.c_int), ("x2_lo_off", ctypes.c_longlong),
        ("x2_stride_w", ctypes.c_longlong), ("x2_stride_h", ctypes.c_longlong), ("x2_stride_n", ctypes.c_longlong),
        ("dot_w", ctypes.c_void_p), ("dot_out", ctypes.c_void_p), ("dot_n", ctypes.c_int), ("dot_pitch", ctypes.c_int),
        ("dot_slots", ctypes.c_int), ("mma_terms", ctypes.c_int)]


EPI_PLAIN, EPI_GRU_ZR, EPI_GRU_GATE = 0, 1, 2
# scflow_conv_tc_desc.algo: operand movement (include/scflow_b200.h)
MMA_TERMS = 3               # bf16 MMAs per product of every conv() launch: 3 = fp32-grade (the refiner's one parity
                            # mode); 2 / 1 = the opt-in reduced modes of scflow_conv_tc_desc.mma_terms (set_precision)
DOT_PITCH = 20              # SCFLOW_CONV_DOT_PITCH
STATS_REPLICAS = 8          # SCFLOW_STATS_REPLICAS: accumulators per (image, channel) of the conv epilogue statistics
ALGO_AUTO, ALGO_TAP_BOXES, ALGO_PATCH, ALGO_PATCH_PAIR = 0, 1, 2, 3


def _rup(v: int, m: int) -> int:
    return (v + m - 1) // m * m


@dataclass
class Split:
    """Channels [off, off + c) of a split-plane buffer `buf` (n, h, w, 2*cp) bf16."""
    buf: torch.Tensor
    c: int
    off: int = 0

    @property
    def cp(self) -> int:
        return self.buf.shape[-1] // 2

    @property
    def shape(self):
        return self.buf.shape[:3]

    @property
    def pitch(self) -> int:
        return self.buf.shape[-1]

    def hi_ptr(self) -> int:
        return self.buf.data_ptr() + 2 * self.off

    def slice(self, off: int, c: int) -> "Split":
        assert off % 8 == 0 and self.off + off + c <= self.cp
        return Split(self.buf, c, self.off + off)

    def float(self) -> torch.Tensor:
        """(n, h, w, c) fp32 = hi + lo (test / debugging helper)."""
        cp = self.cp
        hi = self.buf[..., self.off:self.off + self.c].float()
        lo = self.buf[..., cp + self.off:cp + self.off + self.c].float()
        return hi + lo


def alloc_split(n: int, h: int, w: int, c: int, device, zero: bool = False) -> Split:
    cp = _rup(c, 8)
    mk = torch.zeros if zero else torch.empty
    return Split(mk((n, h, w, 2 * cp), device=device, dtype=torch.bfloat16), c)


def to_split(x: torch.Tensor, out: Optional[Split] = None) -> Split:
    """fp32 channels-last (n, h, w, c) tensor (any pixel pitch: a channel slice of a wider buffer is fine)
    -> split planes (one launch)."""
    if not x.is_cuda or x.dtype != torch.float32 or x.dim() != 4:
        raise _lib.ScflowError("to_split: expected a CUDA fp32 (n,h,w,c) tensor")
    n, h, w, c = x.shape
    if x.stride(3) != 1 or x.stride(1) != w * x.stride(2) or x.stride(0) != h * x.stride(1):
        x = x.contiguous()
    if out is None:
        out = alloc_split(n, h, w, c, x.device)
    assert tuple(out.shape) == (n, h, w) and out.c == c
    lib = _lib.load()
    with _timed("scflow_split_bf16", 1, 8 * x.numel()):
        rc = lib.scflow_split_bf16(_ptr(x), x.stride(2), ctypes.c_void_p(out.hi_ptr()), out.pitch, out.cp, c,
                                   n * h * w, _stream())
    _lib.check(rc, "scflow_split_bf16")
    return out


@dataclass
class Packed:
    w: torch.Tensor          # packed bf16 weights
    bias: Optional[torch.Tensor]     # fp32, padded
    cout: int
    cin: int
    kh: int
    kw: int
    stride: int
    pad_h: int
    pad_w: int
    bk: int


def choose_bk(cout: int, stride: int = 2) -> int:
    """Stride-1 filters: 64-channel k-blocks (128-byte swizzle rows), what the patch-resident kernel reads.
    Others (per-tap kernel): 64 while three ring stages fit next to the weight tile, 32-channel k-blocks
    (64-byte rows) for the wide tiles."""
    if stride == 1:
        return 64
    bn = _rup(cout, 16)
    if bn > 256:
        nt = (bn + 255) // 256
        bn = _rup((bn + nt - 1) // nt, 16)
    stage64 = 2 * (128 * 64 * 2 + bn * 64 * 2)
    return 64 if (227 * 1024 - 2048) // stage64 >= 3 else 32


def pack_weights(weight: torch.Tensor, bias: Optional[torch.Tensor], stride=1, padding=0, scale=None,
                 bk: Optional[int] = None) -> Packed:
    """weight (cout, cin, kh, kw) fp32 [* scale[cout]] -> Packed."""
    lib = _lib.load()
    if not weight.is_cuda:
        raise _lib.ScflowError("pack_weights: expected CUDA parameters (scflow_b200 has no CPU path)")
    w = weight.detach().float().contiguous()
    cout, cin, kh, kw = w.shape
    s = (stride, stride) if isinstance(stride, int) else tuple(stride)
    bk = bk or choose_bk(cout, s[0])
    nbytes = lib.scflow_conv_tc_packed_bytes(cout, cin, kh, kw, bk)
    if nbytes == 0:
        raise _lib.ScflowError(f"conv_tc: unsupported filter {tuple(w.shape)}")
    out = torch.empty(nbytes // 2, device=w.device, dtype=torch.bfloat16)
    sc = None if scale is None else scale.detach().float().contiguous()
    with _timed("scflow_conv_tc_pack_weights"):
        rc = lib.scflow_conv_tc_pack_weights(_ptr(w), _ptr(sc), _ptr(out), cout, cin, kh, kw, bk, _stream())
    _lib.check(rc, "scflow_conv_tc_pack_weights")
    cpad = _rup(cout, 32) + 32
    b = torch.zeros(cpad, device=w.device, dtype=torch.float32)
    if bias is not None:
        b[:cout] = bias.detach().float()
    p = (padding, padding) if isinstance(padding, int) else tuple(padding)
    if s[0] != s[1]:
        raise _lib.ScflowError("conv_tc: anisotropic stride")
    return Packed(out, b if bias is not None else None, cout, cin, kh, kw, s[0], p[0], p[1], bk)


def _key(*tensors):
    return tuple((t.data_ptr(), t._version) for t in tensors if t is not None)


def packed_conv(conv: nn.Conv2d, bn: Optional[nn.BatchNorm2d] = None, with_bias: bool = True,
                tag: str = "", in_order: Optional[torch.Tensor] = None) -> Packed:
    """Cached pack of an nn.Conv2d, optionally with an eval BatchNorm folded in (y = conv*scale + shift).
    `in_order` (needs its own `tag`): input channel c' of the packed filter is the module's channel in_order[c']
    (an operand whose producer wrote its channels in another order, e.g. the lookup's split planes)."""
    ts = [conv.weight, conv.bias]
    if bn is not None:
        ts += [bn.weight, bn.bias, bn.running_mean, bn.running_var]
    key = (_key(*ts), with_bias)
    cache = conv.__dict__.setdefault("_scflow_ctc", {})
    hit = cache.get(tag)
    if hit is not None and hit[0] == key:
        return hit[1]
    if conv.groups != 1 or conv.dilation != (1, 1):
        raise _lib.ScflowError("conv_tc: groups / dilation are not supported")
    with torch.no_grad():
        if bn is not None:
            scale = bn.weight.float() * torch.rsqrt(bn.running_var.float() + bn.eps)
            b0 = conv.bias.float() if conv.bias is not None else 0.0
            shift = bn.bias.float() + (b0 - bn.running_mean.float()) * scale
            pk = pack_weights(conv.weight, shift, conv.stride, conv.padding, scale=scale)
        else:
            wt = conv.weight if in_order is None else conv.weight[:, in_order.to(conv.weight.device)]
            pk = pack_weights(wt, conv.bias if with_bias else None, conv.stride, conv.padding)
    cache[tag] = (key, pk)
    return pk


def out_hw(pk: Packed, h: int, w: int):
    return (h + 2 * pk.pad_h - pk.kh) // pk.stride + 1, (w + 2 * pk.pad_w - pk.kw) // pk.stride + 1


def conv(x: Split, pk: Packed, *, relu: bool = False, res: Optional[Split] = None,
         res_f32: Optional[torch.Tensor] = None, out: Optional[Split] = None,
         out_f32: Optional[torch.Tensor] = None, out_planar: Optional[torch.Tensor] = None,
         planar_tf32: bool = False, f32_tf32: bool = False, stats: Optional[torch.Tensor] = None, want: str = "split",
         view=None, bn: int = 0, epi: int = EPI_PLAIN, gru_h: Optional[torch.Tensor] = None,
         gru_z: Optional[torch.Tensor] = None, algo: int = ALGO_AUTO, x2: Optional[Split] = None,
         dot: Optional[tuple] = None):
    """One convolution launch.  x: Split (n, h, w, cin).  Outputs: `out` (Split), `out_f32` (n, OH, OW, >=cout)
    fp32 channels-last (may be a channel slice of a wider buffer), `out_planar` (n, cout, OH, OW); when none is
    given one is allocated according to `want` ("split" | "f32" | "planar").  `view` = (n, h, w, cin, lo_off,
    stride_w, stride_h, stride_n) overrides how the input buffer is addressed (the stem's overlapped window
    view).  `x2`: a second Split of the same pixels whose channels continue x's as the filter's input channels
    (x.c a multiple of 64): the operand the reference builds with torch.cat, read from where it lies.
    `dot` = (dot_w [cout_pad, DOT_PITCH] fp32, dot_out [n, OH, OW, slots, pitch] fp32, dot_n) with `bn` given and
    slots = 2 * cout_pad / bn: the projections of every finished output pixel, one partial sum per slot (see
    `head_predict`); with `dot` and no explicit output nothing else is written.
    Returns the output object(s) in the order split, f32, planar (only those produced)."""
    lib = _lib.load()
    n, h, w = x.shape
    cin = x.c
    lo_off, sw, sh, sn = x.cp, x.pitch, w * x.pitch, h * w * x.pitch
    if view is not None:
        n, h, w, cin, lo_off, sw, sh, sn = view
    cin2 = 0 if x2 is None else x2.c
    if x2 is not None and (tuple(x2.shape) != (n, h, w) or cin % 64 or view is not None):
        raise _lib.ScflowError("conv_tc: the second input must cover the same pixels and follow a multiple of 64 channels")
    if cin + cin2 != pk.cin:
        raise _lib.ScflowError(f"conv_tc: input has {cin + cin2} channels, the filter takes {pk.cin}")
    oh, ow = out_hw(pk, h, w)
    dev = x.buf.device
    if out is None and out_f32 is None and out_planar is None and dot is None:
        if want == "split":
            out = alloc_split(n, oh, ow, pk.cout, dev)
        elif want == "f32":
            out_f32 = torch.empty((n, oh, ow, _rup(pk.cout, 8)), device=dev, dtype=torch.float32)
        else:
            out_planar = torch.empty((n, _rup(pk.cout, 8), oh, ow), device=dev, dtype=torch.float32)
    d = _Desc()
    d.n, d.h, d.w, d.cin, d.cout = n, h, w, cin, pk.cout
    d.kh, d.kw, d.stride, d.pad_h, d.pad_w, d.bk = pk.kh, pk.kw, pk.stride, pk.pad_h, pk.pad_w, pk.bk
    d.x = x.hi_ptr()
    d.x_lo_off, d.x_stride_w, d.x_stride_h, d.x_stride_n = lo_off, sw, sh, sn
    if x2 is not None:
        d.x2, d.cin2, d.x2_lo_off = x2.hi_ptr(), cin2, x2.cp
        d.x2_stride_w, d.x2_stride_h, d.x2_stride_n = x2.pitch, w * x2.pitch, h * w * x2.pitch
    d.wpack = pk.w.data_ptr()
    d.bias = pk.bias.data_ptr() if pk.bias is not None else None
    d.relu = int(relu)
    c8 = _rup(pk.cout, 8)
    d.bn, d.epi, d.algo = bn, epi, algo
    d.mma_terms = MMA_TERMS
    if epi != EPI_PLAIN:
        # ConvGRU epilogues: exact fp32 state + update gate, (n, oh, ow, C) contiguous; out = a C-channel slice
        assert gru_h.is_contiguous() and gru_z.is_contiguous() and gru_h.shape == gru_z.shape
        d.gru_c, d.gru_h, d.gru_z = gru_h.shape[-1], gru_h.data_ptr(), gru_z.data_ptr()
        c8 = gru_h.shape[-1]
    if res is not None:
        assert tuple(res.shape) == (n, oh, ow) and res.c >= pk.cout
        d.res_split, d.res_pitch, d.res_lo = res.hi_ptr(), res.pitch, res.cp
    if res_f32 is not None:
        assert res_f32.shape[:3] == (n, oh, ow) and res_f32.stride(3) == 1 and res_f32.dtype == torch.float32
        d.res_f32, d.resf_pitch = res_f32.data_ptr(), res_f32.stride(2)
    if out is not None:
        assert tuple(out.shape) == (n, oh, ow) and out.off + c8 <= out.cp, (out.shape, (n, oh, ow), out.off, c8, out.cp)
        d.out_split, d.out_pitch, d.out_lo = out.hi_ptr(), out.pitch, out.cp
    if out_f32 is not None:
        assert out_f32.shape[:3] == (n, oh, ow) and out_f32.shape[3] >= c8 and out_f32.stride(3) == 1
        assert out_f32.stride(1) == ow * out_f32.stride(2) and out_f32.stride(0) == oh * out_f32.stride(1)
        d.out_f32, d.outf_pitch, d.f32_tf32 = out_f32.data_ptr(), out_f32.stride(2), int(f32_tf32)
    if out_planar is not None:
        assert out_planar.shape == (n, c8, oh, ow) and out_planar.is_contiguous()
        d.out_planar, d.planar_tf32 = out_planar.data_ptr(), int(planar_tf32)
    if stats is not None:
        assert stats.dtype == torch.float64 and stats.shape == (n, STATS_REPLICAS, c8, 2) and stats.is_contiguous()
        d.stats = stats.data_ptr()
    if dot is not None:
        dw, dout, dn = dot
        assert dw.dtype == torch.float32 and dw.is_contiguous() and dw.shape[1] == DOT_PITCH and dw.shape[0] >= _rup(pk.cout, 16)
        assert dout.dtype == torch.float32 and dout.is_contiguous() and dout.shape[:3] == (n, oh, ow) and dout.shape[4] >= dn
        d.dot_w, d.dot_out, d.dot_n = dw.data_ptr(), dout.data_ptr(), dn
        d.dot_slots, d.dot_pitch = dout.shape[3], dout.shape[4]
    cin = cin + cin2
    flops = 2 * n * oh * ow * pk.cout * pk.cin * pk.kh * pk.kw
    # algorithmic bytes: split-plane input once, packed filter once, every output written once
    nbytes = (4 * n * h * w * cin + pk.w.numel() * 2
              + n * oh * ow * c8 * 4 * sum(o is not None for o in (out, out_f32, out_planar))
              + (4 * n * oh * ow * c8 if (res is not None or res_f32 is not None) else 0))
    name = f"scflow_conv_tc_fwd[{pk.kh}x{pk.kw}/{pk.stride} {cin}->{pk.cout} @{oh}x{ow}]"
    with _timed(name, 1, nbytes, flops):
        rc = lib.scflow_conv_tc_fwd(ctypes.byref(d), _stream())
    _lib.check(rc, f"scflow_conv_tc_fwd({n}x{h}x{w}x{cin} -> {pk.cout}, {pk.kh}x{pk.kw}/{pk.stride})")
    res_ = tuple(o for o in (out, out_f32, out_planar) if o is not None)
    if not res_:
        return None
    return res_[0] if len(res_) == 1 else res_


def set_precision(mode: str) -> None:
    """"bf16x3" (default, fp32-grade products: the mode every parity claim is made in), or the opt-in reduced modes
    "bf16x2" (activations rounded to bf16, filters exact: 2 MMAs per product) and "bf16" (1 MMA).  Applies to every
    later conv() launch; a captured graph keeps the mode it was captured in."""
    global MMA_TERMS
    MMA_TERMS = {"bf16x3": 3, "bf16x2": 2, "bf16": 1}[mode]


def packed_dot(pred: nn.Conv2d, cout_pad: int) -> torch.Tensor:
    """The 1-2 channel predictor W' (cout', c, ks, ks) behind a head's last layer as projection vectors of that
    layer's output pixel: dot_w[c][(ky ks + kx) cout' + co] = W'[co][c][ky][kx], rows padded to cout_pad."""
    key = _key(pred.weight, pred.bias)
    hit = pred.__dict__.get("_scflow_dot")
    if hit is not None and hit[0] == (key, cout_pad):
        return hit[1]
    co, c, kh, kw = pred.weight.shape
    if kh != kw or kh * kw * co > DOT_PITCH:
        raise _lib.ScflowError(f"head predictor {tuple(pred.weight.shape)}: more than {DOT_PITCH} tap products per pixel")
    with torch.no_grad():
        w = pred.weight.detach().float().permute(1, 2, 3, 0).reshape(c, kh * kw * co)
        dw = torch.zeros(cout_pad, DOT_PITCH, device=w.device, dtype=torch.float32)
        dw[:c, :kh * kw * co] = w
    pred.__dict__["_scflow_dot"] = ((key, cout_pad), dw)
    return dw


def head_predict(x: Split, pk: Packed, pred: nn.Conv2d, *, relu: bool = True) -> torch.Tensor:
    """last 3x3 layer of a head (ReLU) + its 1-2 channel predictor (raft_decoder.py:256-294) without the
    (n, 256, h, w) feature map in between: the layer's epilogue accumulates the predictor's per-pixel tap products
    (scflow_conv_tc_desc.dot_out), scflow_tap_sum_fwd adds the taps and the bias -> planar (n, cout', h, w)."""
    lib = _lib.load()
    n, h, w = x.shape
    oh, ow = out_hw(pk, h, w)
    co, _, ks, _ = pred.weight.shape
    dn = ks * ks * co
    dw = packed_dot(pred, _rup(pk.cout, 16))
    cp = _rup(pk.cout, 16)
    bn = cp // 2 if cp % 32 == 0 and cp // 2 >= 64 else cp            # the fuller-wave tile width conv_tc picks itself
    slots = 2 * (cp // bn)
    dout = torch.empty((n, oh, ow, slots, _rup(dn, 4)), device=x.buf.device, dtype=torch.float32)
    conv(x, pk, relu=relu, dot=(dw, dout, dn), bn=bn)
    out = torch.empty((n, co, oh, ow), device=x.buf.device, dtype=torch.float32)
    with _timed("scflow_tap_sum_fwd", 1, 4 * (dout.numel() + out.numel())):
        rc = lib.scflow_tap_sum_fwd(_ptr(dout), slots, dout.shape[4], _ptr(pred.bias), _ptr(out), n, oh, ow, ks, co,
                                    _stream())
    _lib.check(rc, "scflow_tap_sum_fwd")
    return out


def expand_taps(x: torch.Tensor, kw: int) -> Split:
    """planar fp32 (n, c, h, w), c * kw <= 16 -> Split (n, h, w, 16): per pixel its kw x-neighbours of every channel
    (scflow_expand_taps_fwd)."""
    lib = _lib.load()
    n, c, h, w = x.shape
    if not x.is_cuda or x.dtype != torch.float32:
        raise _lib.ScflowError("expand_taps: expected a CUDA fp32 tensor")
    x = x.contiguous()
    out = alloc_split(n, h, w, 16, x.device)
    with _timed("scflow_expand_taps_fwd", 1, 4 * x.numel() + 64 * n * h * w):
        rc = lib.scflow_expand_taps_fwd(_ptr(x), _ptr(out.buf), n, c, kw, h, w, _stream())
    _lib.check(rc, "scflow_expand_taps_fwd")
    return out


def packed_small_cin(conv: nn.Conv2d) -> Packed:
    """A kh x kw filter on 1-2 input channels as the kh x 1 filter on the 16 expanded channels of expand_taps
    (input channel dx * c + ci), cached like packed_conv."""
    key = _key(conv.weight, conv.bias)
    hit = conv.__dict__.get("_scflow_cin")
    if hit is not None and hit[0] == key:
        return hit[1]
    cout, c, kh, kw = conv.weight.shape
    with torch.no_grad():
        w = conv.weight.detach().float().permute(0, 3, 1, 2).reshape(cout, kw * c, kh)     # [co][dx*c + ci][ky]
        w = torch.nn.functional.pad(w, (0, 0, 0, 16 - kw * c)).unsqueeze(-1).contiguous()    # (cout, 16, kh, 1)
        pk = pack_weights(w, conv.bias, 1, (conv.padding[0], 0))
    conv.__dict__["_scflow_cin"] = (key, pk)
    return pk


def split_planar(x: torch.Tensor, out: Split) -> None:
    """planar fp32 (n, cf, h, w) -> the cf channels of the split slice `out` (scflow_split_planar)."""
    lib = _lib.load()
    n, cf, h, w = x.shape
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and tuple(out.shape) == (n, h, w) and out.c == cf
    with _timed("scflow_split_planar", 1, 8 * x.numel()):
        rc = lib.scflow_split_planar(_ptr(x), ctypes.c_void_p(out.hi_ptr()), out.pitch, out.cp, n, cf, h * w, _stream())
    _lib.check(rc, "scflow_split_planar")


def norm_apply(x: torch.Tensor, sums: torch.Tensor, *, relu: bool = True, eps: float = 1e-5,
               res: Optional[Split] = None, res_raw: Optional[torch.Tensor] = None,
               res_sums: Optional[torch.Tensor] = None, out: Optional[Split] = None,
               out_f32: Optional[torch.Tensor] = None) -> Split:
    """InstanceNorm (+ residual) + ReLU of a raw conv output x (n, h, w, c) fp32 with the statistics of the
    conv epilogue -> split planes (scflow_norm_apply_split)."""
    lib = _lib.load()
    n, h, w, c = x.shape
    assert x.is_contiguous() and sums.shape == (n, STATS_REPLICAS, c, 2) and sums.dtype == torch.float64
    if out is None:
        out = alloc_split(n, h, w, c, x.device)
    with _timed("scflow_norm_apply_split", 1, (8 + (4 if res is not None or res_raw is not None else 0)) * x.numel()):
        rc = lib.scflow_norm_apply_split(
            _ptr(x), _ptr(sums), None if res is None else ctypes.c_void_p(res.hi_ptr()),
            0 if res is None else res.pitch, 0 if res is None else res.cp, _ptr(res_raw), _ptr(res_sums),
            ctypes.c_void_p(out.hi_ptr()), out.pitch, out.cp, _ptr(out_f32), n, h * w, c, eps, int(relu), _stream())
    _lib.check(rc, "scflow_norm_apply_split")
    return out
