"""Per-kernel micro-benchmark at cfg2 shapes (N=64, 256x256 crops -> 32x32 features).
    python tools/kbench.py [lookup] [corr] [pose] [small] [posefc] [--n 64] [--s 256]   (convolutions: tools/convbench.py)
Each kernel is timed with CUDA events over `reps` launches on rotating buffers larger than L2."""
import os, sys, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from scflow_b200 import ops, modules

ap = argparse.ArgumentParser()
ap.add_argument("which", nargs="*", default=["lookup", "corr", "pose", "small"])
ap.add_argument("--n", type=int, default=64)
ap.add_argument("--s", type=int, default=256)
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
dev = "cuda"
N, S = a.n, a.s
h = w = S // 8
P = h * w
HBM = 6553.3


def timeit(fn, reps=a.reps, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3   # us


g = torch.Generator(device=dev).manual_seed(0)
if "corr" in a.which:
    f1 = [torch.randn(N, 256, h, w, device=dev, generator=g) for _ in range(2)]
    f2 = [torch.randn(N, 256, h, w, device=dev, generator=g) for _ in range(2)]
    import ctypes
    from scflow_b200 import _lib
    lib = _lib.load()
    f1 = [ops._planar_tf32(t, "f") for t in f1]; f2 = [ops._planar_tf32(t, "f") for t in f2]
    lv = [torch.empty((N * P, 1, h >> l, w >> l), device=dev) for l in range(4)]
    arr = (ctypes.c_void_p * 4)(*[l.data_ptr() for l in lv])
    st = torch.cuda.current_stream().cuda_stream
    def run(i):       # the kernel alone: operands already rounded, outputs preallocated
        rc = lib.scflow_corr_pyramid_fwd(f1[i % 2].data_ptr(), f2[i % 2].data_ptr(), arr, 4, N, 256, h, w,
                                         ops.PREC_TF32, None, 0, st)
        assert rc == 0, rc
    us = timeit(run)
    by = N * (2 * P * 256 * 4 + 85 * P * P * 4 // 64)
    fl = N * 2.0 * P * P * 256
    print(f"corr_pyramid tf32   {us:8.1f} us  {by/us/1e3:7.1f} GB/s ({by/us/1e3/HBM*100:4.1f}% HBM)  {fl/us/1e6:6.1f} TFLOP/s")
if "lookup" in a.which:
    sets = 2 if N * P * P * 4 * 85 // 64 > 100e6 else 4
    pyrs = [[torch.randn(N * P, 1, h >> l, w >> l, device=dev, generator=g) for l in range(4)] for _ in range(sets)]
    flow = [2.0 * torch.randn(N, 2, h, w, device=dev, generator=g) for _ in range(sets)]
    by = N * (P * 324 * 4 + P * 4 * 100 * 4 + 8 * P)
    for tag, g1 in (("gen 1", True), ("gen 2", False)):
        us = timeit(lambda i: ops.corr_lookup(pyrs[i % sets], flow[i % sets], 4, channels_last=True, gen1=g1))
        print(f"corr_lookup nhwc {tag:7s} {us:8.1f} us  {by/us/1e3:7.1f} GB/s ({by/us/1e3/HBM*100:4.1f}% HBM)")
    zero = torch.zeros_like(flow[0])
    us = timeit(lambda i: ops.corr_lookup(pyrs[i % sets], zero, 4, channels_last=True))
    print(f"corr_lookup nhwc zero flow (per-tap rows at every integer level) {us:8.1f} us")
    half = [f * (torch.rand(N, 1, h, w, device=dev, generator=g) > 0.57) for f in flow]     # 57 % background as in the bench crops
    us = timeit(lambda i: ops.corr_lookup(pyrs[i % sets], half[i % sets], 4, channels_last=True))
    print(f"corr_lookup nhwc 57 % zero-flow queries {us:8.1f} us  {by/us/1e3:7.1f} GB/s ({by/us/1e3/HBM*100:4.1f}% HBM)")
    us = timeit(lambda i: ops.corr_lookup(pyrs[i % sets], flow[i % sets], 4, channels_last=False))
    print(f"corr_lookup nchw    {us:8.1f} us  {by/us/1e3:7.1f} GB/s ({by/us/1e3/HBM*100:4.1f}% HBM)")
if "pose" in a.which:
    from scflow_b200 import synthetic as syn
    cr = {k: v.to(dev) for k, v in syn.make_crops(4, S, seed=1).items()}
    rep = lambda t: t.repeat((N // 4,) + (1,) * (t.dim() - 1)).contiguous()
    c = {k: rep(v) for k, v in cr.items()}
    dR6 = torch.tensor([1., 0, 0, 0, 1., 0], device=dev).repeat(N, 1) + 0.01 * torch.randn(N, 6, device=dev)
    dt = 0.01 * torch.randn(N, 3, device=dev)
    us = timeit(lambda i: ops.pose_flow(c["R0"], c["t0"], c["K"], dR6=dR6, dt=dt, depth=c["depth"], R0=c["R0"], t0=c["t0"]))
    by = N * (4 * S * S + 8 * S * S + 8 * P)
    print(f"pose_flow           {us:8.1f} us  {by/us/1e3:7.1f} GB/s ({by/us/1e3/HBM*100:4.1f}% HBM)  (includes 4 output allocations)")

if "small" in a.which:
    from scflow_b200.modules import ConvModule, XHead
    for cin, k, cout in ((2, 7, 128), (1, 3, 64)):
        m = ConvModule(cin, cout, k, padding=k // 2).to(dev)
        xs = [torch.randn(N, cin, h, w, device=dev, generator=g) for _ in range(4)]
        with torch.no_grad():
            us = timeit(lambda i: m(xs[i % 4]))
        by = N * P * (cin + cout) * 4
        print(f"conv_small_cin {cin}->{cout} {k}x{k}  {us:8.1f} us  {by/us/1e3:7.1f} GB/s")
    for k, cout in ((3, 2), (1, 1)):
        hd = XHead(128, [256], cout, x="flow" if k == 3 else "mask").to(dev)
        ys = [torch.randn(N, 256, h, w, device=dev, generator=g).contiguous(memory_format=torch.channels_last) for _ in range(4)]
        by = N * P * (256 + cout) * 4
        with torch.no_grad():
            us = timeit(lambda i: ops.conv_small_cout(ys[i % 4], hd.predict_layer))
        print(f"conv_small_cout 256->{cout} {k}x{k} {us:8.1f} us  {by/us/1e3:7.1f} GB/s ({by/us/1e3/HBM*100:4.1f}% HBM)")
if "posefc" in a.which:
    from scflow_b200.pose_head import MultiClassPoseHead
    head = MultiClassPoseHead(21, 224, "Basic", dict(type="GN", num_groups=32), dict(type="ReLU"),
                              feat_size=(h, w), rotation_mode="ortho6d").to(dev)
    fc_in = head.fc_layers[0][0].in_features
    xs = [torch.randn(N, fc_in, device=dev, generator=g) for _ in range(4)]
    lab = torch.zeros(N, dtype=torch.int64, device=dev)
    w0 = head.fc_layers[0][0].weight.detach()
    with torch.no_grad():
        us = timeit(lambda i: ops.pose_fc(xs[i % 4], w0, head, lab))
        def ref(i):
            y = head.fc_layers(xs[i % 4]); return head.rotation_pred(y), head.translation_pred(y)
        r = timeit(ref)
    print(f"pose_fc (3 launches) {us:8.1f} us   (cuBLAS fp32 chain {r:.1f} us)")
