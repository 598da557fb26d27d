"""GPU tier, collected last on purpose: the no-allocation entry point
scflow_pose_fc_fwd_static (split-K scratch in a file-scope `static __device__` array of the library; no
counterpart in the reference) against the existing
scflow_pose_fc_fwd (caller-supplied workspace) through the C ABI, on two calls in a row with different
inputs -- the array keeps the previous call's contents, which must not matter."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from scflow_b200 import _lib

DEV = "cuda"
MAX_FLOATS = 1 << 20            # SCFLOW_POSE_FC_STATIC_MAX_FLOATS (include/scflow_b200.h)


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _problem(seed, M, K, n0, n1, classes):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s, sc=1.0: (torch.randn(*s, generator=g) * sc).to(DEV)
    return dict(x=r(M, K), w0=r(n0, K, sc=K ** -0.5), b0=r(n0, sc=0.1), w1=r(n1, n0, sc=n0 ** -0.5), b1=r(n1, sc=0.1),
                wr=r(6 * classes, n1, sc=n1 ** -0.5), br=r(6 * classes, sc=0.1), wt=r(3 * classes, n1, sc=n1 ** -0.5),
                bt=r(3 * classes, sc=0.1), label=torch.randint(0, classes, (M,), generator=g).to(DEV))


def _run(lib, q, M, K, n0, n1, classes, splitk, static):
    dr = torch.full((M, 6), float("nan"), device=DEV)
    dt = torch.full((M, 3), float("nan"), device=DEV)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    head = [_p(q[k]) for k in ("x", "w0", "b0", "w1", "b1", "wr", "br", "wt", "bt", "label")]
    tail = [_p(dr), _p(dt), M, K, n0, n1, classes, splitk, stream]
    if static:
        before = torch.cuda.memory_allocated()
        rc = lib.scflow_pose_fc_fwd_static(*head, *tail)
        assert torch.cuda.memory_allocated() == before          # nothing came from torch's allocator either
    else:
        ws = torch.empty(lib.scflow_pose_fc_workspace_floats(M, n0, n1, splitk), device=DEV)
        rc = lib.scflow_pose_fc_fwd(*head, _p(ws), *tail)
    assert rc == 0, lib.scflow_strerror(rc)
    torch.cuda.synchronize()
    return dr, dt


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def test_pose_fc_static_scratch_matches_workspace_entry_point():
    lib = _lib.load()
    # two calls in a row: the shipped head at the bench batch, then a different shape and different data
    for seed, (M, K, n0, n1, classes, splitk) in enumerate([(64, 2048, 1024, 256, 21, 4), (5, 512, 256, 128, 3, 2),
                                                            (64, 2048, 1024, 256, 21, 4)]):
        q = _problem(100 + seed, M, K, n0, n1, classes)
        dr_w, dt_w = _run(lib, q, M, K, n0, n1, classes, splitk, static=False)
        dr_s, dt_s = _run(lib, q, M, K, n0, n1, classes, splitk, static=True)
        assert torch.isfinite(dr_s).all() and torch.isfinite(dt_s).all()
        # same kernels, same deterministic summation order: the pose-head tests' tolerance is 1e-4, these agree
        # to rounding (bit-identical in practice)
        assert _rel(dr_s, dr_w) < 1e-6 and _rel(dt_s, dt_w) < 1e-6, (seed, _rel(dr_s, dr_w), _rel(dt_s, dt_w))
        # and against the fp64 evaluation of models/head/pose_head.py:201-211 (class of label[0] for every crop)
        a0 = torch.relu(q["x"].double() @ q["w0"].double().t() + q["b0"].double())
        a1 = torch.relu(a0 @ q["w1"].double().t() + q["b1"].double())
        c = int(q["label"][0])
        want_r = (a1 @ q["wr"].double().t() + q["br"].double()).view(M, classes, 6)[:, c]
        want_t = (a1 @ q["wt"].double().t() + q["bt"].double()).view(M, classes, 3)[:, c]
        assert _rel(dr_s, want_r) < 1e-4 and _rel(dt_s, want_t) < 1e-4


def test_pose_fc_static_rejects_beyond_the_maximum():
    lib = _lib.load()
    M, K, n0, n1, classes, splitk = 256, 2048, 1024, 256, 21, 4          # 4 * 256 * 1280 floats > 2^20
    assert lib.scflow_pose_fc_workspace_floats(M, n0, n1, splitk) > MAX_FLOATS
    q = _problem(7, 8, K, n0, n1, classes)          # pointers only need to be valid: the size check comes first
    dr, dt = torch.zeros(M, 6, device=DEV), torch.zeros(M, 3, device=DEV)
    rc = lib.scflow_pose_fc_fwd_static(*[_p(q[k]) for k in ("x", "w0", "b0", "w1", "b1", "wr", "br", "wt", "bt", "label")],
                                       _p(dr), _p(dt), M, K, n0, n1, classes, splitk, None)
    assert rc == -9 and b"static device scratch" in lib.scflow_strerror(rc)
    torch.cuda.synchronize()
    assert float(dr.abs().max()) == 0.0             # nothing was launched
