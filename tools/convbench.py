"""Micro-benchmark of the tensor-core convolution family at the refiner's layer shapes (cfg2: 64 crops of 256x256).
    python tools/convbench.py [--n 64] [--reps 10] [--only SUBSTR] [--once]
Every shape is timed with CUDA events over `reps` launches on rotating operand / output sets larger than L2 and
printed with its algorithmic TFLOP/s, executed-MMA fraction of the bf16 peak and operand bytes a tile pulls
through L2.  --once: a single launch per shape (the form to run under ncu)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from scflow_b200 import convtc as C

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=64)
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--only", default="")
ap.add_argument("--once", action="store_true")
ap.add_argument("--bn", type=int, default=0, help="output-channel tile width override (0: widest)")
ap.add_argument("--algo", type=int, default=0, help="0 auto, 1 per-tap boxes, 2 patch, 3 patch on CTA pairs")
ap.add_argument("--hw-scale", type=int, default=1, help="k: the same pixels as k x k larger images (cfg4 = 2), n / k^2 of them")
ap.add_argument("--precision", default="bf16x3", choices=["bf16x3", "bf16x2", "bf16"])
a = ap.parse_args()
C.set_precision(a.precision)
dev = "cuda"
N = a.n
PEAK = 1400.0
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    PEAK = float(json.load(open(p)).get("bf16_tflops_sustained", PEAK))

# name, h, w, cin, cout, kh, kw, stride, pad_h, pad_w, out ("split" | "f32+stats" | "f32" | "planar"), calls per step
SHAPES = [
    ("enc 3x3 64->64 @128 (IN raw+stats)", 128, 128, 64, 64, 3, 3, 1, 1, 1, "f32+stats", 8),
    ("enc 3x3 64->64 @128 (no stats, f32)", 128, 128, 64, 64, 3, 3, 1, 1, 1, "f32", 0),
    ("enc 3x3 64->64 @128 (BN split)", 128, 128, 64, 64, 3, 3, 1, 1, 1, "split", 4),
    ("enc 3x3/2 64->96 @128->64", 128, 128, 64, 96, 3, 3, 2, 1, 1, "f32+stats", 3),
    ("enc 1x1/2 64->96 @128->64", 128, 128, 64, 96, 1, 1, 2, 0, 0, "f32+stats", 3),
    ("enc 3x3 96->96 @64", 64, 64, 96, 96, 3, 3, 1, 1, 1, "f32+stats", 9),
    ("enc 3x3/2 96->128 @64->32", 64, 64, 96, 128, 3, 3, 2, 1, 1, "f32+stats", 3),
    ("enc 3x3 128->128 @32", 32, 32, 128, 128, 3, 3, 1, 1, 1, "f32+stats", 9),
    ("enc 1x1 128->256 @32", 32, 32, 128, 256, 1, 1, 1, 0, 0, "planar", 3),
    ("mot 1x1 324->256 @32", 32, 32, 324, 256, 1, 1, 1, 0, 0, "split", 8),
    ("mot 3x3 256->192 @32", 32, 32, 256, 192, 3, 3, 1, 1, 1, "split", 8),
    ("mot 3x3 128->64 @32", 32, 32, 128, 64, 3, 3, 1, 1, 1, "split", 16),
    ("mot 3x3 256->126 @32", 32, 32, 256, 126, 3, 3, 1, 1, 1, "split", 8),
    ("gru 1x5 384->256 @32", 32, 32, 384, 256, 1, 5, 1, 0, 2, "split", 8),
    ("gru 5x1 384->256 @32", 32, 32, 384, 256, 5, 1, 1, 2, 0, "split", 8),
    ("gru 1x5 256->128 @32", 32, 32, 256, 128, 1, 5, 1, 0, 2, "f32", 8),
    ("gru 1x5 128->128 @32", 32, 32, 128, 128, 1, 5, 1, 0, 2, "split", 8),
    ("head 3x3 128->256 @32", 32, 32, 128, 256, 3, 3, 1, 1, 1, "f32", 16),
    ("head 3x3 128->256 @32 + 3x3 ->2 predictor", 32, 32, 128, 256, 3, 3, 1, 1, 1, "dot18", 8),
    ("head 3x3 128->256 @32 + 1x1 ->1 predictor", 32, 32, 128, 256, 3, 3, 1, 1, 1, "dot1", 8),
    ("small-cin 7x1 16->128 @32 (flow / mask encoders)", 32, 32, 16, 128, 7, 1, 1, 3, 0, "split", 16),
    ("small-cin 3x1 16->64 @32", 32, 32, 16, 64, 3, 1, 1, 1, 0, "split", 8),
    ("head 3x3 64->32 @32", 32, 32, 64, 32, 3, 3, 1, 1, 1, "split", 8),
    ("pose 3x3/2 224->128 @32->16", 32, 32, 224, 128, 3, 3, 2, 1, 1, "f32", 8),
    ("pose 3x3/2 128->128 @16->8", 16, 16, 128, 128, 3, 3, 2, 1, 1, "f32", 8),
    ("GRU zr epilogue 1x5 256->256 @32 (context precomputed)", 32, 32, 256, 256, 1, 5, 1, 0, 2, "gru_zr", 16),
    ("GRU gate epilogue 1x5 128+128->128 @32 (context precomputed)", 32, 32, 256, 128, 1, 5, 1, 0, 2, "gru_gate", 16),
    ("GRU zr epilogue 1x5 384->256 @32", 32, 32, 384, 256, 1, 5, 1, 0, 2, "gru_zr", 16),
    ("GRU gate epilogue 1x5 128+256->128 @32", 32, 32, 384, 128, 1, 5, 1, 0, 2, "gru_gate", 16),
]


def timeit(fn, reps, warm=2):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


g = torch.Generator(device=dev).manual_seed(0)
total_us = 0.0
for name, h, w, cin, cout, kh, kw, s, ph, pw, out, calls in SHAPES:
    if a.only and a.only not in name:
        continue
    h, w, N = h * a.hw_scale, w * a.hw_scale, a.n // (a.hw_scale * a.hw_scale)
    wt = torch.randn(cout, cin, kh, kw, device=dev, generator=g) / (cin * kh * kw) ** 0.5
    b = torch.randn(cout, device=dev, generator=g)
    pk = C.pack_weights(wt, None if "stats" in out else b, s, (ph, pw))
    oh, ow = C.out_hw(pk, h, w)
    in_bytes = N * h * w * cin * 4
    sets = max(2, int(200e6 // max(in_bytes, 1)) + 1)
    sets = min(sets, 6)
    xs = [C.to_split(torch.randn(N, h, w, cin, device=dev, generator=g)) for _ in range(sets)]
    c8 = (cout + 7) // 8 * 8
    kw_ = []
    for i in range(sets):
        if out == "split":
            kw_.append(dict(out=C.alloc_split(N, oh, ow, cout, dev), relu=True))
        elif out == "f32":
            kw_.append(dict(out_f32=torch.empty(N, oh, ow, c8, device=dev), relu=True))
        elif out == "planar":
            kw_.append(dict(out_planar=torch.empty(N, c8, oh, ow, device=dev), planar_tf32=True))
        elif out.startswith("dot"):
            dn = int(out[3:])
            kw_.append(dict(relu=True, bn=128, dot=(torch.randn(256, C.DOT_PITCH, device=dev),
                                                    torch.empty(N, oh, ow, 4, (dn + 3) // 4 * 4, device=dev), dn)))
        elif out == "gru_zr":
            kw_.append(dict(out=C.alloc_split(N, oh, ow, 128, dev), epi=C.EPI_GRU_ZR, bn=128,
                            gru_h=torch.randn(N, oh, ow, 128, device=dev), gru_z=torch.empty(N, oh, ow, 128, device=dev)))
        elif out == "gru_gate":
            kw_.append(dict(out=C.alloc_split(N, oh, ow, 128, dev), epi=C.EPI_GRU_GATE,
                            gru_h=torch.randn(N, oh, ow, 128, device=dev), gru_z=torch.rand(N, oh, ow, 128, device=dev)))
        else:
            kw_.append(dict(out_f32=torch.empty(N, oh, ow, c8, device=dev),
                            stats=torch.zeros(N, C.STATS_REPLICAS, c8, 2, device=dev, dtype=torch.float64)))
    algo = a.algo if (s == 1 or a.algo <= 1) else 1
    if a.bn and cout % a.bn == 0 and cout > a.bn and "bn" not in kw_[0]:
        for k_ in kw_:
            k_["bn"] = a.bn
    run = lambda i: C.conv(xs[i % sets], pk, algo=algo, **kw_[i % sets])
    if a.once:
        run(0)
        torch.cuda.synchronize()
        continue
    us = timeit(run, a.reps)
    fl = 2.0 * N * oh * ow * cout * cin * kh * kw
    dram = in_bytes + N * oh * ow * c8 * 4 + pk.w.numel() * 2
    total_us += us * calls
    print(f"{name:36s} {us:8.1f} us  {fl/us/1e6:6.1f} TFLOP/s alg  pipe {3*fl/us/1e6/PEAK*100:5.1f}%  "
          f"dram {dram/us/1e3:6.0f} GB/s  x{calls:2d} = {us*calls/1e3:6.3f} ms")
if not a.once:
    print(f"sum over a step's calls: {total_us/1e3:.3f} ms")
