"""Per-iteration durations of the lookup (and the GRU) inside one eager refiner step, GPU kept
behind the host with a device-side sleep: which iterations cost what."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from scflow_b200 import ops
from scflow_b200.refiner import SCFlowRefinerCore
dev = torch.device("cuda")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
core = bench.seeded_weights(SCFlowRefinerCore(256)).eval().to(dev)
batch = {k: v.to(dev) for k, v in bench.tile_batch(bench.make_batch(4, 1234), 64).items()}
for _ in range(3):
    core.refine(batch)
torch.cuda.synchronize()
ops.STATS.reset(timing=True)
torch.cuda._sleep(int(40e-3 * 1.9e9))
core.refine(batch)
torch.cuda.synchronize()
for want in ("scflow_corr_lookup_fwd", "scflow_convgru_fwd", "scflow_pose_flow_fwd"):
    ts = [a.elapsed_time(b) * 1e3 for name, _, a, b, _ in ops.STATS.events if name == want]
    print(want, " ".join(f"{t:6.1f}" for t in ts))
d = batch["depth"]
print("background fraction", float((d <= 0).float().mean()))
