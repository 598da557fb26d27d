"""Torch-tensor front ends of the C-ABI launchers (device pointers + current stream).

Each function allocates its outputs with torch (the library never allocates), checks
device/dtype/contiguity, and raises on any non-zero return code.  CPU tensors are rejected:
there is no fallback.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from . import _lib

PREC_FP32, PREC_TF32 = 0, 1
DIV_IEEE, DIV_RCP = 0, 1


class LaunchStats:
    """Counts launches of OUR kernels; with `timing` on, brackets each launcher call with CUDA
    events on the launching (current) stream so bench.py can attribute device time per kernel."""

    def __init__(self):
        self.count = 0
        self.timing = False
        self.events = []          # (name, kernels_in_call, start, end, algorithmic bytes or None, flops or None)

    def reset(self, timing: bool = False):
        self.count, self.timing, self.events = 0, timing, []

    def summary(self):
        """{name: (calls, total_ms, total algorithmic bytes or None, total algorithmic flops or None)} -- call
        after a synchronize."""
        out = {}
        for name, _, a, b, nb, fl in self.events:
            c, t, by, f = out.get(name, (0, 0.0, 0, 0))
            out[name] = (c + 1, t + a.elapsed_time(b), None if (nb is None or by is None) else by + nb,
                         None if (fl is None or f is None) else f + fl)
        return out


STATS = LaunchStats()


class _timed:
    def __init__(self, name, kernels=1, nbytes=None, flops=None):
        self.name, self.kernels, self.nbytes, self.flops = name, kernels, nbytes, flops

    def __enter__(self):
        STATS.count += self.kernels
        if STATS.timing:
            self.a = torch.cuda.Event(enable_timing=True)
            self.b = torch.cuda.Event(enable_timing=True)
            self.a.record()
        return self

    def __exit__(self, *exc):
        if STATS.timing:
            self.b.record()
            STATS.events.append((self.name, self.kernels, self.a, self.b, self.nbytes, self.flops))
        return False


def _chk(t: torch.Tensor, name: str, dtype=torch.float32) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.ScflowError(f"{name}: expected a CUDA tensor (scflow_b200 has no CPU path)")
    if t.dtype != dtype:
        raise _lib.ScflowError(f"{name}: expected {dtype}, got {t.dtype}")
    return t.contiguous()


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _planar_tf32(t: torch.Tensor, name: str) -> torch.Tensor:
    """NCHW-contiguous copy of a feature map rounded to the nearest TF32 (one launch); takes
    channels-last tensors (the fused encoder's output) without a separate layout pass."""
    lib = _lib.load()
    if not t.is_cuda or t.dtype != torch.float32:
        _chk(t, name)
    if getattr(t, "_scflow_planar_tf32", False) and t.is_contiguous():
        return t        # the encoder's last convolution already wrote this format (encoder.forward_tc)
    n, c, h, w = t.shape
    out = torch.empty((n, c, h, w), device=t.device, dtype=torch.float32)
    if t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous():
        with _timed("scflow_nhwc_to_nchw", 1, 8 * t.numel()):
            rc = lib.scflow_nhwc_to_nchw(_ptr(t), _ptr(out), n, h * w, c, 1, _stream())
        _lib.check(rc, "scflow_nhwc_to_nchw")
        return out
    t = t.contiguous()
    if t.numel() % 4:
        raise _lib.ScflowError(f"{name}: element count must be a multiple of 4")
    with _timed("scflow_pack_nhwc", 1):
        rc = lib.scflow_pack_nhwc(_ptr(t), 4, None, 0, None, 0, _ptr(out), 4, 0, t.numel() // 4, 1, 1, _stream())
    _lib.check(rc, "scflow_pack_nhwc")
    return out


def corr_pyramid(feat1: torch.Tensor, feat2: torch.Tensor, num_levels: int = 4,
                 precision: int = PREC_TF32) -> List[torch.Tensor]:
    """CorrelationPyramid.forward (raft_decoder.py:35-58): list of (N*H*W, 1, H>>l, W>>l)."""
    lib = _lib.load()
    if precision == PREC_TF32:
        # tcgen05 kind::tf32 truncates its operands; hand it features rounded to nearest instead
        f1, f2 = _planar_tf32(feat1, "feat1"), _planar_tf32(feat2, "feat2")
    else:
        f1, f2 = _chk(feat1, "feat1"), _chk(feat2, "feat2")
    n, c, h, w = f1.shape
    assert f2.shape == f1.shape
    levels = [torch.empty((n * h * w, 1, h >> l, w >> l), device=f1.device, dtype=torch.float32)
              for l in range(num_levels)]
    arr = (ctypes.c_void_p * num_levels)(*[l.data_ptr() for l in levels])
    with _timed("scflow_corr_pyramid_fwd", 1 if precision == PREC_TF32 else 2):
        rc = lib.scflow_corr_pyramid_fwd(_ptr(f1), _ptr(f2), arr, num_levels, n, c, h, w, precision,
                                         None, 0, _stream())
    _lib.check(rc, "scflow_corr_pyramid_fwd")
    return levels


LAYOUT_NCHW, LAYOUT_NHWC, LAYOUT_NHWC_GEN1, LAYOUT_NHWC_SPLIT = 0, 1, 2, 3


def corr_pyramid_bwd(grad_corr: torch.Tensor, feat1: torch.Tensor, feat2: torch.Tensor, want1: bool = True,
                     want2: bool = True, precision: int = PREC_TF32):
    """The two GEMMs behind the backward of `torch.matmul(feat1^T, feat2)` (raft_decoder.py:53):
    grad_corr (n, P, P) = d loss / d (f1^T f2) -> (d feat1, d feat2), one `scflow_corr_pyramid_bwd` launch
    (tcgen05 kind::tf32 at the shipped shapes; fp32 FMA for PREC_FP32).  corr[q,t] = sum_c f1[c,q] f2[c,t], so
    d f1 = f2 . G^T and d f2 = f1 . G."""
    lib = _lib.load()
    f1, f2, g = _chk(feat1, "feat1"), _chk(feat2, "feat2"), _chk(grad_corr, "grad_corr")
    n, c, h, w = f1.shape
    P = h * w
    if tuple(g.shape) != (n, P, P) or f2.shape != f1.shape:
        raise _lib.ScflowError(f"corr_pyramid_bwd: grad_corr {tuple(g.shape)} vs features {tuple(f1.shape)}")
    g1 = torch.empty_like(f1) if want1 else None
    g2 = torch.empty_like(f2) if want2 else None
    if g1 is None and g2 is None:
        return None, None
    nprod = (g1 is not None) + (g2 is not None)
    with _timed("scflow_corr_pyramid_bwd", 1, 4 * n * (P * P + 2 * c * P) * nprod):
        rc = lib.scflow_corr_pyramid_bwd(_ptr(g), _ptr(f1), _ptr(f2), _ptr(g1), _ptr(g2), n, c, h, w, precision,
                                         _stream())
    _lib.check(rc, "scflow_corr_pyramid_bwd")
    return g1, g2


def lookup_split_order(num_levels: int = 4, radius: int = 4) -> torch.Tensor:
    """Channel order of the lookup's split-plane output (SCFLOW_LAYOUT_NHWC_SPLIT): plane channel c' holds the
    reference's channel order[c'] (corr_lookup.py:134: channel = 81 level + 9 a + b).  Lane-major: lane = 9 g + a of
    the kernel owns taps (a, b = 3 g + j) of every level and writes its twelve values side by side, c' = 12 lane +
    3 level + j.  The 1x1 convolution behind the lookup is packed with its input channels in this order."""
    d = 2 * radius + 1
    assert num_levels == 4 and d == 9, "the split layout exists for the shipped lookup (4 levels, radius 4)"
    order = torch.empty(num_levels * d * d, dtype=torch.long)
    for lane in range(27):
        g, a = divmod(lane, 9)
        for l in range(num_levels):
            for j in range(3):
                order[12 * lane + 3 * l + j] = d * d * l + d * a + 3 * g + j
    return order


def corr_lookup(pyramid: Sequence[torch.Tensor], flow: torch.Tensor, radius: int = 4,
                div_mode: int = DIV_IEEE, return_windows: bool = False, channels_last: bool = True,
                gen1: bool = False, split: bool = False):
    """CorrLookup.forward (corr_lookup.py:102-136): (N, L*(2r+1)^2, H, W).

    With `channels_last` (default) the result has the reference's shape but channels-last
    strides (memory (N, H, W, 324)): what the warp-per-query kernel writes and what the
    downstream convolutions consume without a layout pass.  `gen1`: the first-generation kernel on any
    shape (the tests' cross-check; the second generation runs wherever it applies otherwise).
    `split`: return the result as a convtc.Split (split-bf16 planes) instead of an fp32 tensor, its channels in
    the kernel's lane-major order (`lookup_split_order`)."""
    lib = _lib.load()
    fl = _chk(flow, "flow")
    n, two, h, w = fl.shape
    assert two == 2
    lv = [_chk(p, f"pyramid[{i}]") for i, p in enumerate(pyramid)]
    for l, p in enumerate(lv):
        if p.shape != (n * h * w, 1, h >> l, w >> l):
            raise _lib.ScflowError(f"pyramid[{l}] has shape {tuple(p.shape)}")
    L = len(lv)
    d = 2 * radius + 1
    if split:
        # straight into the split-bf16 planes the 1x1 convolution behind the lookup reads (convtc.Split);
        # shapes the second-generation kernel does not take go through fp32 + one split pass
        from . import convtc as C
        sp = C.alloc_split(n, h, w, L * d * d, fl.device)
        arr = (ctypes.c_void_p * L)(*[p.data_ptr() for p in lv])
        with _timed("scflow_corr_lookup_fwd", 1):
            rc = lib.scflow_corr_lookup_fwd(arr, L, _ptr(fl), _ptr(sp.buf), None, n, h, w, radius, div_mode,
                                            LAYOUT_NHWC_SPLIT, _stream())
        if rc == -2:                                   # SCFLOW_E_UNSUPPORTED
            nat = corr_lookup(pyramid, flow, radius, div_mode).permute(0, 2, 3, 1)
            return C.to_split(nat[..., lookup_split_order(L, radius).to(nat.device)].contiguous(), out=sp)
        _lib.check(rc, "scflow_corr_lookup_fwd")
        return sp
    if channels_last:
        out = torch.empty((n, h, w, L * d * d), device=fl.device, dtype=torch.float32).permute(0, 3, 1, 2)
    else:
        out = torch.empty((n, L * d * d, h, w), device=fl.device, dtype=torch.float32)
    win = torch.empty((n * h * w, L, 2, d), device=fl.device, dtype=torch.int16) if return_windows else None
    arr = (ctypes.c_void_p * L)(*[p.data_ptr() for p in lv])
    with _timed("scflow_corr_lookup_fwd", 1):
        rc = lib.scflow_corr_lookup_fwd(arr, L, _ptr(fl), _ptr(out), _ptr(win), n, h, w, radius, div_mode,
                                        (LAYOUT_NHWC_GEN1 if gen1 else LAYOUT_NHWC) if channels_last else LAYOUT_NCHW,
                                        _stream())
    _lib.check(rc, "scflow_corr_lookup_fwd")
    return (out, win) if return_windows else out


def pack_nhwc(dst: torch.Tensor, dst_off: int, a: Optional[torch.Tensor] = None,
              b: Optional[torch.Tensor] = None, planar: Optional[torch.Tensor] = None,
              round_tf32: bool = True) -> torch.Tensor:
    """dst[..., dst_off:dst_off+ca+cb+cf] = cat[a, b, planar] per pixel, optionally rounded to the
    nearest TF32.  dst (N,H,W,C) and a, b (N,H,W,c) are channels-last contiguous; planar is
    (N,cf,H,W).  Replaces the torch.cat + layout pass that builds the ConvGRU input."""
    lib = _lib.load()
    dst = _chk(dst, "dst")
    n, H, W, cd = dst.shape
    ca = cb = cf = 0
    if a is not None:
        a = _chk(a, "a"); ca = a.shape[-1]
        assert a.shape[:3] == dst.shape[:3]
    if b is not None:
        b = _chk(b, "b"); cb = b.shape[-1]
        assert b.shape[:3] == dst.shape[:3]
    if planar is not None:
        planar = _chk(planar, "planar"); cf = planar.shape[1]
        assert planar.shape[0] == n and planar.shape[2:] == (H, W)
    with _timed("scflow_pack_nhwc", 1, 8 * n * H * W * (ca + cb + cf)):
        rc = lib.scflow_pack_nhwc(_ptr(a), ca, _ptr(b), cb, _ptr(planar), cf, _ptr(dst), cd, dst_off,
                                  n * H * W, H * W, 1 if round_tf32 else 0, _stream())
    _lib.check(rc, "scflow_pack_nhwc")
    return dst


def lift_points(depth, K, R0, t0):
    """Dense cal_3d_2d_corr (pose.py:44-64): points3d (N,3,H,W), mask (N,H,W) uint8."""
    lib = _lib.load()
    depth, K, R0, t0 = _chk(depth, "depth"), _chk(K, "K"), _chk(R0, "R0"), _chk(t0, "t0")
    n, H, W = depth.shape
    pts = torch.empty((n, 3, H, W), device=depth.device, dtype=torch.float32)
    mask = torch.empty((n, H, W), device=depth.device, dtype=torch.uint8)
    with _timed("scflow_lift_points", 1):
        rc = lib.scflow_lift_points(_ptr(depth), _ptr(K), _ptr(R0), _ptr(t0), _ptr(pts), _ptr(mask), n, H, W, _stream())
    _lib.check(rc, "scflow_lift_points")
    return pts, mask


def pose_flow(R_prev, t_prev, K, *, dR6=None, dt=None, depth=None, R0=None, t0=None,
              points3d=None, mask=None, H: int = 0, W: int = 0, scale: int = 8,
              invalid_num: float = 0.0, weight: float = 10.0, want_full: bool = True,
              want_lr: bool = True):
    """Fused pose-compose + pose-induced flow + 1/scale down-sample (see scflow_pose_flow_fwd).
    Returns (R, t, flow_full or None, flow_lr or None)."""
    lib = _lib.load()
    R_prev, t_prev, K = _chk(R_prev, "R_prev"), _chk(t_prev, "t_prev"), _chk(K, "K")
    n = R_prev.shape[0]
    dev = R_prev.device
    if dR6 is not None:
        dR6, dt = _chk(dR6, "dR6"), _chk(dt, "dt")
    if points3d is not None:
        points3d, mask = _chk(points3d, "points3d"), _chk(mask, "mask", torch.uint8)
        H, W = points3d.shape[-2:]
    else:
        depth, R0, t0 = _chk(depth, "depth"), _chk(R0, "R0"), _chk(t0, "t0")
        H, W = depth.shape[-2:]
    R = torch.empty((n, 3, 3), device=dev, dtype=torch.float32)
    t = torch.empty((n, 3), device=dev, dtype=torch.float32)
    full = torch.empty((n, 2, H, W), device=dev, dtype=torch.float32) if want_full else None
    lr = torch.empty((n, 2, H // scale, W // scale), device=dev, dtype=torch.float32) if want_lr else None
    with _timed("scflow_pose_flow_fwd", 1):
        rc = lib.scflow_pose_flow_fwd(_ptr(dR6), _ptr(dt), _ptr(R_prev), _ptr(t_prev), _ptr(K), _ptr(R0),
                                      _ptr(t0), _ptr(depth), _ptr(points3d), _ptr(mask), _ptr(R), _ptr(t),
                                      _ptr(full), _ptr(lr), n, H, W, scale, float(invalid_num),
                                      float(weight), _stream())
    _lib.check(rc, "scflow_pose_flow_fwd")
    return R, t, full, lr


def upsample(x, add=None, scale: int = 8, mul: float = 1.0):
    """mul * F.interpolate(x [+ add], scale_factor=scale, bilinear, align_corners=True)."""
    lib = _lib.load()
    x = _chk(x, "x")
    if add is not None:
        add = _chk(add, "add")
        assert add.shape == x.shape
    n, c, h, w = x.shape
    out = torch.empty((n, c, h * scale, w * scale), device=x.device, dtype=torch.float32)
    with _timed("scflow_upsample_fwd", 1, 4 * (x.numel() * (2 if add is not None else 1) + out.numel())):
        rc = lib.scflow_upsample_fwd(_ptr(x), _ptr(add), _ptr(out), n, c, h, w, scale, float(mul), _stream())
    _lib.check(rc, "scflow_upsample_fwd")
    return out


def downsample(x, scale: int = 8, mul: float = 1.0):
    """mul * F.interpolate(x, scale_factor=1/scale, bilinear, align_corners=True)."""
    lib = _lib.load()
    x = _chk(x, "x")
    n, c, H, W = x.shape
    out = torch.empty((n, c, H // scale, W // scale), device=x.device, dtype=torch.float32)
    with _timed("scflow_downsample_fwd", 1, 4 * (x.numel() + out.numel())):
        rc = lib.scflow_downsample_fwd(_ptr(x), _ptr(out), n, c, H, W, scale, float(mul), _stream())
    _lib.check(rc, "scflow_downsample_fwd")
    return out


def _packed_conv_weight(conv: torch.nn.Conv2d) -> torch.Tensor:
    """Conv2d weight (cout, cin, k, k) as [cin][k][k][cout] fp32, cached on the module and rebuilt when
    the parameter changes (version counter / storage), never saved."""
    w = conv.weight
    key = (w._version, w.data_ptr(), w.device)
    hit = conv.__dict__.get("_scflow_wpack")
    if hit is None or hit[0] != key:
        hit = (key, w.detach().float().permute(1, 2, 3, 0).contiguous())
        conv.__dict__["_scflow_wpack"] = hit
    return hit[1]


def conv_small_cin(x: torch.Tensor, conv: torch.nn.Conv2d, relu: bool, split_out: bool = False):
    """act(conv(x) + bias) for cin in {1,2}: returns an NCHW-shaped channels-last tensor, or, with
    `split_out`, a convtc.Split (split-bf16 planes: the operand of the tensor-core convolution that follows).
    x may be planar or channels-last (strides are passed through)."""
    lib = _lib.load()
    if not x.is_cuda or x.dtype != torch.float32:
        _chk(x, "x")
    n, cin, h, w = x.shape
    if x.is_contiguous():
        sn, sc, sp = cin * h * w, h * w, 1
    elif x.is_contiguous(memory_format=torch.channels_last):
        sn, sc, sp = cin * h * w, 1, cin
    else:
        x = x.contiguous()
        sn, sc, sp = cin * h * w, h * w, 1
    cout, k = conv.out_channels, conv.kernel_size[0]
    out = sp_ = None
    spl = (None, 0, 0)
    if split_out:
        from . import convtc as C
        sp_ = C.alloc_split(n, h, w, cout, x.device)
        spl = (ctypes.c_void_p(sp_.hi_ptr()), sp_.pitch, sp_.cp)
    else:
        out = torch.empty((n, h, w, cout), device=x.device, dtype=torch.float32)
    with _timed("scflow_conv_small_cin_fwd", 1, 4 * (x.numel() + n * h * w * cout)):
        rc = lib.scflow_conv_small_cin_fwd(_ptr(x), sn, sc, sp, _ptr(_packed_conv_weight(conv)),
                                           _ptr(conv.bias.detach() if conv.bias is not None else None),
                                           _ptr(out), spl[0], spl[1], spl[2], n, cin, h, w, cout, k,
                                           1 if relu else 0, _stream())
    _lib.check(rc, "scflow_conv_small_cin_fwd")
    return sp_ if split_out else out.permute(0, 3, 1, 2)


def stem_s2d_split(x: torch.Tensor) -> torch.Tensor:
    """The same space-to-depth image as two bf16 planes (2, N, H/2+3, W/2+3, 16), value = hi + lo: the
    operand of the tensor-core stem convolution (scflow_stem_s2d_split_fwd)."""
    lib = _lib.load()
    x = _chk(x, "x")
    n, c, h, w = x.shape
    out = torch.empty((2, n, h // 2 + 3, w // 2 + 3, 16), device=x.device, dtype=torch.bfloat16)
    with _timed("scflow_stem_s2d_split_fwd", 1, 4 * x.numel() + 2 * out.numel()):
        rc = lib.scflow_stem_s2d_split_fwd(_ptr(x), _ptr(out), n, c, h, w, _stream())
    _lib.check(rc, "scflow_stem_s2d_split_fwd")
    return out


def conv_small_cout(x: torch.Tensor, conv: torch.nn.Conv2d) -> torch.Tensor:
    """conv(x) + bias for cout in {1,2} on a channels-last x: planar (n, cout, h, w)."""
    lib = _lib.load()
    n, c, h, w = x.shape
    xn = _chk(x.permute(0, 2, 3, 1), "x")          # no copy for channels-last inputs
    cout, k = conv.out_channels, conv.kernel_size[0]
    out = torch.empty((n, cout, h, w), device=x.device, dtype=torch.float32)
    with _timed("scflow_conv_small_cout_fwd", 1, 4 * (x.numel() + out.numel())):
        rc = lib.scflow_conv_small_cout_fwd(_ptr(xn), _ptr(_packed_conv_weight(conv)),
                                            _ptr(conv.bias.detach() if conv.bias is not None else None),
                                            _ptr(out), n, c, h, w, cout, k, _stream())
    _lib.check(rc, "scflow_conv_small_cout_fwd")
    return out


def group_norm4_act(x: torch.Tensor, gn: torch.nn.GroupNorm, relu: bool, split_out=None):
    """GroupNorm (4 channels per group) + ReLU on a channels-last (N,H,W,C) tensor: in place (returns x), or,
    with `split_out` (a convtc.Split of the same pixels), written as split planes only -- the operand of the
    next tensor-core convolution, without an fp32 round trip (returns split_out)."""
    lib = _lib.load()
    x = _chk(x, "x")
    n, H, W, C = x.shape
    sp = (None, 0, 0) if split_out is None else (ctypes.c_void_p(split_out.hi_ptr()), split_out.pitch, split_out.cp)
    if split_out is not None:
        assert tuple(split_out.shape) == (n, H, W) and split_out.c == C
    with _timed("scflow_group_norm4_act_nhwc", 1, 8 * x.numel()):
        rc = lib.scflow_group_norm4_act_nhwc(_ptr(x), _ptr(gn.weight.detach()), _ptr(gn.bias.detach()),
                                             None if split_out is not None else _ptr(x), sp[0], sp[1], sp[2],
                                             n, H * W, C, float(gn.eps), 1 if relu else 0, _stream())
    _lib.check(rc, "scflow_group_norm4_act_nhwc")
    return x if split_out is None else split_out


def pose_tail(part1: torch.Tensor, head, label: torch.Tensor):
    """ReLU(fc1 pre-activation (M, n1) + bias) -> (dr (M,6), dt (M,3)) of the class of label[0]
    (scflow_pose_tail_fwd)."""
    lib = _lib.load()
    part1 = _chk(part1, "part1")
    M, n1 = part1.shape
    label = _chk(label, "label", torch.int64)
    dr = torch.empty((M, 6), device=part1.device, dtype=torch.float32)
    dt = torch.empty((M, 3), device=part1.device, dtype=torch.float32)
    d = lambda t: _ptr(t.detach())
    with _timed("scflow_pose_tail_fwd", 1):
        rc = lib.scflow_pose_tail_fwd(_ptr(part1), d(head.fc_layers[1][0].bias), d(head.rotation_pred.weight),
                                      d(head.rotation_pred.bias), d(head.translation_pred.weight),
                                      d(head.translation_pred.bias), _ptr(label), _ptr(dr), _ptr(dt), M, 1, n1,
                                      head.num_class, _stream())
    _lib.check(rc, "scflow_pose_tail_fwd")
    return dr, dt


def pose_fc(x2d: torch.Tensor, w0: torch.Tensor, head, label: torch.Tensor, splitk: int = 4):
    """FC tail of MultiClassPoseHead on x2d (M, K) with the (possibly column-permuted) fc0 weight w0:
    returns (dr (M,6), dt (M,3)) of the class of label[0]."""
    lib = _lib.load()
    x2d = _chk(x2d, "x")
    M, K = x2d.shape
    fc0, fc1 = head.fc_layers[0][0], head.fc_layers[1][0]
    n0, n1 = fc0.out_features, fc1.out_features
    label = _chk(label, "label", torch.int64)
    ws = torch.empty(lib.scflow_pose_fc_workspace_floats(M, n0, n1, splitk), device=x2d.device, dtype=torch.float32)
    dr = torch.empty((M, 6), device=x2d.device, dtype=torch.float32)
    dt = torch.empty((M, 3), device=x2d.device, dtype=torch.float32)
    d = lambda t: _ptr(t.detach())
    with _timed("scflow_pose_fc_fwd", 3, 4 * (w0.numel() + x2d.numel())):
        rc = lib.scflow_pose_fc_fwd(_ptr(x2d), d(w0), d(fc0.bias), d(fc1.weight), d(fc1.bias),
                                    d(head.rotation_pred.weight), d(head.rotation_pred.bias),
                                    d(head.translation_pred.weight), d(head.translation_pred.bias),
                                    _ptr(label), _ptr(ws), _ptr(dr), _ptr(dt), M, K, n0, n1, head.num_class,
                                    splitk, _stream())
    _lib.check(rc, "scflow_pose_fc_fwd")
    return dr, dt
