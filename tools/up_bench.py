import os,sys,torch
sys.path.insert(0,'/root/repo')
from scflow_b200 import ops
dev='cuda'
g=torch.Generator(device=dev).manual_seed(0)
for c in (2,1):
    xs=[torch.randn(256,c,32,32,device=dev,generator=g) for _ in range(4)]
    ys=[torch.randn(256,c,32,32,device=dev,generator=g) for _ in range(4)]
    for i in range(3): ops.upsample(xs[i%4], ys[i%4], 8, 8.0)
    torch.cuda.synchronize()
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(20): ops.upsample(xs[i%4], ys[i%4] if c==2 else None, 8, 8.0)
    b.record(); torch.cuda.synchronize()
    print("upsample c=%d: %.1f us"%(c, a.elapsed_time(b)/20*1e3))
