"""torch.profiler view of one bench step (which ATen / cuDNN / scflow kernels the time goes to).
    python tools/torch_profile.py [batch]  -> prints the top CUDA kernels and top ATen ops."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from scflow_b200.refiner import SCFlowRefinerCore

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
core = bench.seeded_weights(SCFlowRefinerCore(256)).eval().to(dev)
batch = {k: v.to(dev) for k, v in bench.tile_batch(bench.make_batch(4, 1234), B).items()}
for _ in range(3):
    core.refine(batch)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    core.refine(batch)
    torch.cuda.synchronize()
ka = prof.key_averages()
print(ka.table(sort_by="self_cuda_time_total", row_limit=45, max_name_column_width=70))
