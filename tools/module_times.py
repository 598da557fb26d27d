"""Per-module device time of one refiner forward (eager, CUDA events around every leaf module and
every launcher of ours): where the cuDNN / torch-glue time of the step goes.
    python tools/module_times.py [--batch 64] [--image 256]"""
import os, sys, argparse, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from scflow_b200 import ops
from scflow_b200.refiner import SCFlowRefinerCore

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--image", type=int, default=256)
a = ap.parse_args()
dev = torch.device("cuda")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
core = bench.seeded_weights(SCFlowRefinerCore(a.image)).eval().to(dev)
batch = {k: v.to(dev) for k, v in bench.tile_batch(bench.make_batch(4, 1234, a.image), a.batch).items()}
for _ in range(3):
    core.refine(batch)
torch.cuda.synchronize()
ev = []
def pre(name):
    def f(m, i):
        e = torch.cuda.Event(enable_timing=True); e.record(); m.__dict__["_t0"] = e
    return f
def post(name):
    def f(m, i, o):
        e = torch.cuda.Event(enable_timing=True); e.record(); ev.append((name, type(m).__name__, m.__dict__["_t0"], e))
    return f
for name, m in core.named_modules():
    if len(list(m.children())) == 0 or type(m).__name__ in ("ConvModule",):
        m.register_forward_pre_hook(pre(name)); m.register_forward_hook(post(name))
ops.STATS.reset(timing=True)
a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a0.record(); core.refine(batch); a1.record()
torch.cuda.synchronize()
tot = a0.elapsed_time(a1)
agg = collections.OrderedDict()
for name, ty, s, e in ev:
    c, t = agg.get((name, ty), (0, 0.0)); agg[(name, ty)] = (c + 1, t + s.elapsed_time(e))
print(f"eager step {tot:.2f} ms")
for (name, ty), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{t:8.3f} ms  {100*t/tot:5.1f}%  x{c:3d}  {t/c*1e3:7.1f} us  {ty:14s} {name}")
print("-- our launchers")
for name, (c, t, nb, _fl) in sorted(ops.STATS.summary().items(), key=lambda kv: -kv[1][1]):
    print(f"{t:8.3f} ms  {100*t/tot:5.1f}%  x{c:3d}  {t/c*1e3:7.1f} us  {name}")
