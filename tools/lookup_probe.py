"""Why the lookup is slower inside the refiner than alone: time it alone on (a) random flows, fp32 and split output,
(b) the pyramid and the eight flows the refiner actually passes (captured from one forward of the bench's model)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from scflow_b200 import ops
from scflow_b200.refiner import SCFlowRefinerCore

dev = torch.device("cuda:0")
def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3

N, h, w = 64, 32, 32
P = h * w
g = torch.Generator(device=dev).manual_seed(0)
pyrs = [[torch.randn(N * P, 1, h >> l, w >> l, device=dev, generator=g) for l in range(4)] for _ in range(2)]
flow = 2.0 * torch.randn(N, 2, h, w, device=dev, generator=g)
i = [0]
def alt(split):
    i[0] += 1
    return ops.corr_lookup(pyrs[i[0] % 2], flow, 4, split=split)
print("random flow, fp32 out  %.1f us" % timeit(lambda: alt(False)))
print("random flow, split out %.1f us" % timeit(lambda: alt(True)))

core = bench.seeded_weights(SCFlowRefinerCore(256)).eval().to(dev)
batch = {k: v.to(dev) for k, v in bench.tile_batch(bench.make_batch(4, 1234, 256), N).items()}
cap = []
orig = ops.corr_lookup
def spy(pyramid, flow, *a, **kw):
    cap.append(([p for p in pyramid], flow.clone()))
    return orig(pyramid, flow, *a, **kw)
ops.corr_lookup = spy
with torch.no_grad():
    core.refine(batch)
ops.corr_lookup = orig
print("captured", len(cap), "lookups")
for it, (pyr, fl) in enumerate(cap):
    big = float((fl.abs() > 4).float().mean()); zero = float((fl == 0).float().mean())
    us = timeit(lambda: orig(pyr, fl, 4, split=True))
    us2 = timeit(lambda: orig(pyrs[0], fl, 4, split=True))
    print(f"iter {it}: refiner pyramid+flow {us:.1f} us | random pyramid, refiner flow {us2:.1f} us | |flow|>4: {big:.2f}, ==0: {zero:.2f}, max {float(fl.abs().max()):.1f}")
fl = cap[-1][1]
print("refiner pyramid, random flow %.1f us" % timeit(lambda: orig(cap[-1][0], flow, 4, split=True)))
