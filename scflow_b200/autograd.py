"""torch.autograd.Function pairs over the C-ABI forward / backward launchers (training, cfg5).

The reference trains through ATen's autograd (bmm / avg_pool2d / grid_sample / interpolate and the
elementwise pose chain).  Here each hot-path operator keeps its single forward launch and gets a
hand-written adjoint (scflow_b200/csrc/backward.cu):

  corr_pyramid   raft_decoder.py:47-56     fold kernel + two plain batched GEMMs (cuBLAS)
  corr_lookup    corr_lookup.py:102-136    one scatter kernel (gradient wrt the pyramid; the flow is
                                           detached on the SCFlow path, scflow_decoder.py:193-194)
  upsample       scflow_decoder.py:223-228 one gather kernel
  pose_flow      pose.py:66-88             one per-crop reduction kernel (wrt R_k, t_k)
  compose_pose   pose.py:124-169           scflow_pose_compose_fwd / _bwd (one launch each)
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence

import torch
import torch.nn.functional as F

from . import _lib, ops
from .ops import _chk, _ptr, _stream, _timed


def _ptr_array(tensors: Sequence[torch.Tensor]):
    return (ctypes.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])


class CorrPyramidFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat1, feat2, num_levels, precision):
        levels = ops.corr_pyramid(feat1, feat2, num_levels, precision)
        ctx.save_for_backward(feat1, feat2)
        ctx.num_levels = num_levels
        ctx.precision = precision
        return tuple(levels)

    @staticmethod
    def backward(ctx, *grads):
        feat1, feat2 = ctx.saved_tensors
        lib = _lib.load()
        n, c, h, w = feat1.shape
        P = h * w
        gs = [None if g is None else _chk(g, "grad_level") for g in grads]
        if all(g is None for g in gs):
            return None, None, None, None
        gcorr = torch.empty((n, P, P), device=feat1.device, dtype=torch.float32)
        with _timed("scflow_corr_fold_bwd", 1, 4 * n * P * P * 2):
            rc = lib.scflow_corr_fold_bwd(_ptr_array(gs), ctx.num_levels, _ptr(gcorr), n, c, h, w, _stream())
        _lib.check(rc, "scflow_corr_fold_bwd")
        g1, g2 = ops.corr_pyramid_bwd(gcorr, feat1, feat2, ctx.needs_input_grad[0], ctx.needs_input_grad[1],
                                      ctx.precision)
        return g1, g2, None, None


class CorrLookupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, flow, radius, div_mode, channels_last, *levels):
        out = ops.corr_lookup(levels, flow, radius, div_mode, channels_last=channels_last)
        ctx.save_for_backward(flow)
        ctx.meta = (radius, div_mode, [tuple(l.shape) for l in levels])
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (flow,) = ctx.saved_tensors
        radius, div_mode, shapes = ctx.meta
        lib = _lib.load()
        n, _, h, w = flow.shape
        if grad_out.is_contiguous():
            layout = ops.LAYOUT_NCHW
        else:
            grad_out = grad_out.contiguous(memory_format=torch.channels_last)
            layout = ops.LAYOUT_NHWC
        if not grad_out.is_cuda or grad_out.dtype != torch.float32:
            raise _lib.ScflowError("corr_lookup backward: expected a CUDA fp32 gradient")
        glv = [torch.empty(s, device=flow.device, dtype=torch.float32) for s in shapes]
        nbytes = 4 * (grad_out.numel() + sum(g.numel() for g in glv))
        with _timed("scflow_corr_lookup_bwd", 1, nbytes):
            rc = lib.scflow_corr_lookup_bwd(_ptr(grad_out), layout, _ptr(flow), _ptr_array(glv), len(glv),
                                            n, h, w, radius, div_mode, _stream())
        _lib.check(rc, "scflow_corr_lookup_bwd")
        return (None, None, None, None, *glv)


class UpsampleFn(torch.autograd.Function):
    """mul * up(x [+ add]); both inputs receive the same gradient."""

    @staticmethod
    def forward(ctx, x, add, scale, mul):
        ctx.meta = (scale, mul, tuple(x.shape))
        return ops.upsample(x, add, scale, mul)

    @staticmethod
    def backward(ctx, grad_out):
        scale, mul, shape = ctx.meta
        lib = _lib.load()
        g = _chk(grad_out, "grad_out")
        n, c, h, w = shape
        gin = torch.empty(shape, device=g.device, dtype=torch.float32)
        with _timed("scflow_upsample_bwd", 1, 4 * (2 * g.numel() + gin.numel())):
            rc = lib.scflow_upsample_bwd(_ptr(g), _ptr(gin), n, c, h, w, scale, float(mul), _stream())
        _lib.check(rc, "scflow_upsample_bwd")
        return (gin if ctx.needs_input_grad[0] else None,
                gin if ctx.needs_input_grad[1] else None, None, None)


class PoseFlowFn(torch.autograd.Function):
    """Pose-induced flow of an explicit pose (R, t): returns (flow_full, flow_lr); flow_lr is the
    detached 1/scale flow the next iteration starts from."""

    @staticmethod
    def forward(ctx, R, t, K, depth, R0, t0, scale, invalid_num):
        _, _, full, lr = ops.pose_flow(R, t, K, depth=depth, R0=R0, t0=t0, scale=scale,
                                       invalid_num=invalid_num)
        ctx.save_for_backward(R, t, K, depth, R0, t0)
        ctx.mark_non_differentiable(lr)
        return full, lr

    @staticmethod
    def backward(ctx, grad_full, _grad_lr):
        R, t, K, depth, R0, t0 = ctx.saved_tensors
        lib = _lib.load()
        g = _chk(grad_full, "grad_full")
        n, H, W = depth.shape
        gp = torch.empty((n, 12), device=g.device, dtype=torch.float32)
        with _timed("scflow_pose_flow_bwd", 1, 4 * (g.numel() + depth.numel())):
            rc = lib.scflow_pose_flow_bwd(_ptr(g), _ptr(depth), _ptr(K), _ptr(R0), _ptr(t0),
                                          _ptr(R.contiguous()), _ptr(t.contiguous()), _ptr(gp), n, H, W, _stream())
        _lib.check(rc, "scflow_pose_flow_bwd")
        return gp[:, :9].reshape(n, 3, 3), gp[:, 9:].contiguous(), None, None, None, None, None, None


def corr_pyramid(feat1, feat2, num_levels=4, precision=ops.PREC_TF32) -> List[torch.Tensor]:
    return list(CorrPyramidFn.apply(feat1, feat2, num_levels, precision))


def corr_lookup(pyramid, flow, radius=4, div_mode=ops.DIV_IEEE, channels_last=True):
    return CorrLookupFn.apply(flow.detach(), radius, div_mode, channels_last, *pyramid)


def upsample(x, add=None, scale=8, mul=1.0):
    return UpsampleFn.apply(x, add, scale, mul)


def pose_flow(R, t, K, depth, R0, t0, scale=8, invalid_num=0.0):
    return PoseFlowFn.apply(R, t, K, depth, R0, t0, scale, invalid_num)


class GruResetFn(torch.autograd.Function):
    """rh = sigmoid(a_r) * h (raft_decoder.py:249-250): scflow_gru_reset_fwd / _bwd."""

    @staticmethod
    def forward(ctx, a_r, h):
        lib = _lib.load()
        a, hh = _chk(a_r, "a_r"), _chk(h, "h")
        if a.shape != hh.shape:
            raise _lib.ScflowError("gru_reset: a_r and h differ in shape")
        out = torch.empty_like(hh)
        with _timed("scflow_gru_reset_fwd", 1, 12 * hh.numel()):
            rc = lib.scflow_gru_reset_fwd(_ptr(a), _ptr(hh), _ptr(out), hh.numel(), _stream())
        _lib.check(rc, "scflow_gru_reset_fwd")
        ctx.save_for_backward(a, hh)
        return out

    @staticmethod
    def backward(ctx, g):
        a, hh = ctx.saved_tensors
        lib = _lib.load()
        g = _chk(g, "grad_rh")
        ga, gh = torch.empty_like(a), torch.empty_like(hh)
        with _timed("scflow_gru_reset_bwd", 1, 20 * hh.numel()):
            rc = lib.scflow_gru_reset_bwd(_ptr(g), _ptr(a), _ptr(hh), _ptr(ga), _ptr(gh), hh.numel(), _stream())
        _lib.check(rc, "scflow_gru_reset_bwd")
        return ga, gh


class GruBlendFn(torch.autograd.Function):
    """h' = (1 - sigmoid(a_z)) h + sigmoid(a_z) tanh(a_q) (raft_decoder.py:248-252): scflow_gru_blend_fwd / _bwd."""

    @staticmethod
    def forward(ctx, a_z, a_q, h):
        lib = _lib.load()
        az, aq, hh = _chk(a_z, "a_z"), _chk(a_q, "a_q"), _chk(h, "h")
        if az.shape != hh.shape or aq.shape != hh.shape:
            raise _lib.ScflowError("gru_blend: a_z, a_q and h differ in shape")
        out = torch.empty_like(hh)
        with _timed("scflow_gru_blend_fwd", 1, 16 * hh.numel()):
            rc = lib.scflow_gru_blend_fwd(_ptr(az), _ptr(aq), _ptr(hh), _ptr(out), hh.numel(), _stream())
        _lib.check(rc, "scflow_gru_blend_fwd")
        ctx.save_for_backward(az, aq, hh)
        return out

    @staticmethod
    def backward(ctx, g):
        az, aq, hh = ctx.saved_tensors
        lib = _lib.load()
        g = _chk(g, "grad_h")
        gz, gq, gh = torch.empty_like(az), torch.empty_like(aq), torch.empty_like(hh)
        with _timed("scflow_gru_blend_bwd", 1, 28 * hh.numel()):
            rc = lib.scflow_gru_blend_bwd(_ptr(g), _ptr(az), _ptr(aq), _ptr(hh), _ptr(gz), _ptr(gq), _ptr(gh),
                                          hh.numel(), _stream())
        _lib.check(rc, "scflow_gru_blend_bwd")
        return gz, gq, gh


def gru_reset(a_r, h):
    return GruResetFn.apply(a_r, h)


def gru_blend(a_z, a_q, h):
    return GruBlendFn.apply(a_z, a_q, h)


class PoseComposeFn(torch.autograd.Function):
    """get_pose_from_delta_pose (pose.py:124-149, ortho6d + 'exp'): scflow_pose_compose_fwd / _bwd, one launch each."""

    @staticmethod
    def forward(ctx, rotation_delta, translation_delta, rotation_src, translation_src, weight, detach_depth_for_xy):
        lib = _lib.load()
        r6, dt = _chk(rotation_delta, "rotation_delta"), _chk(translation_delta, "translation_delta")
        Rs, ts = _chk(rotation_src, "rotation_src"), _chk(translation_src, "translation_src")
        n = r6.shape[0]
        if r6.shape != (n, 6) or dt.shape != (n, 3) or Rs.shape != (n, 3, 3) or ts.shape != (n, 3):
            raise _lib.ScflowError("compose_pose: expected (n,6) ortho6d, (n,3), (n,3,3), (n,3)")
        R, t = torch.empty_like(Rs), torch.empty_like(ts)
        with _timed("scflow_pose_compose_fwd", 1, 4 * n * 33):
            rc = lib.scflow_pose_compose_fwd(_ptr(r6), _ptr(dt), _ptr(Rs), _ptr(ts), _ptr(R), _ptr(t), n, float(weight),
                                             _stream())
        _lib.check(rc, "scflow_pose_compose_fwd")
        ctx.save_for_backward(r6, dt, Rs, ts)
        ctx.meta = (float(weight), bool(detach_depth_for_xy))
        return R, t

    @staticmethod
    def backward(ctx, gR, gt):
        r6, dt, Rs, ts = ctx.saved_tensors
        weight, detach = ctx.meta
        lib = _lib.load()
        n = r6.shape[0]
        if gR is None and gt is None:
            return None, None, None, None, None, None
        gR = None if gR is None else _chk(gR, "grad_R")
        gt = None if gt is None else _chk(gt, "grad_t")
        g6, gdt = torch.empty_like(r6), torch.empty_like(dt)
        gRs = torch.empty_like(Rs) if ctx.needs_input_grad[2] else None
        gts = torch.empty_like(ts) if ctx.needs_input_grad[3] else None
        with _timed("scflow_pose_compose_bwd", 1, 4 * n * 54):
            rc = lib.scflow_pose_compose_bwd(_ptr(gR), _ptr(gt), _ptr(r6), _ptr(dt), _ptr(Rs), _ptr(ts), _ptr(g6),
                                             _ptr(gdt), _ptr(gRs), _ptr(gts), n, weight, int(detach), _stream())
        _lib.check(rc, "scflow_pose_compose_bwd")
        return g6, gdt, gRs, gts, None, None


def compose_pose(rotation_delta, translation_delta, rotation_src, translation_src, weight=10.0,
                 detach_depth_for_xy=False):
    """get_pose_from_delta_pose (pose.py:124-149, ortho6d + 'exp') on the training path: one kernel forward, one
    backward (`PoseComposeFn`)."""
    return PoseComposeFn.apply(rotation_delta.float(), translation_delta.float(), rotation_src.float(),
                               translation_src.float(), weight, detach_depth_for_xy)


def compose_pose_torch(rotation_delta, translation_delta, rotation_src, translation_src, weight=10.0,
                       detach_depth_for_xy=False):
    """The same function with torch ops (differentiable as written): what `compose_pose` replaced; kept as the
    tests' second implementation."""
    x_raw, y_raw = rotation_delta[:, 0:3], rotation_delta[:, 3:6]
    x = F.normalize(x_raw, p=2, dim=1)
    z = F.normalize(torch.cross(x, y_raw, dim=1), p=2, dim=1)
    y = torch.cross(z, x, dim=1)
    dR = torch.stack((x, y, z), dim=2)                      # columns (x, y, z)
    R = torch.bmm(dR, rotation_src)
    tz_src = translation_src[:, 2]
    tz = tz_src / torch.exp(translation_delta[:, 2])
    tz_xy = tz.detach() if detach_depth_for_xy else tz
    tx = tz_xy * (translation_delta[:, 0] / weight + translation_src[:, 0] / tz_src)
    ty = tz_xy * (translation_delta[:, 1] / weight + translation_src[:, 1] / tz_src)
    return R, torch.stack((tx, ty, tz), dim=1)
