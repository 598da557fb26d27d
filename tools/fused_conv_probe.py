"""Probe: cuDNN fused conv+bias+ReLU (torch.cudnn_convolution_relu / _add_relu) against conv + our bias/ReLU pass,
channels-last fp32 (TF32), at the decoder's and encoder's shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from scflow_b200 import ops
dev = "cuda"
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
def timeit(fn, reps=20):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): fn(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3
g = torch.Generator(device=dev).manual_seed(0)
for (n, cin, cout, hw, k, s) in ((64, 256, 192, 32, 3, 1), (64, 128, 256, 32, 3, 1), (64, 324, 256, 32, 1, 1), (64, 256, 128, 32, 3, 1),
                                 (64, 64, 64, 128, 3, 1), (64, 96, 96, 64, 3, 1), (64, 128, 128, 32, 3, 1), (64, 64, 96, 128, 3, 2),
                                 (64, 3, 64, 256, 7, 2)):
    xs = [torch.randn(n, cin, hw, hw, device=dev, generator=g).contiguous(memory_format=torch.channels_last) for _ in range(3)]
    w = (torch.randn(cout, cin, k, k, device=dev, generator=g) * 0.05).contiguous(memory_format=torch.channels_last)
    b = torch.randn(cout, device=dev, generator=g)
    ones = torch.ones(cout, device=dev)
    pad = k // 2
    def ours(i):
        y = F.conv2d(xs[i % 3], w, None, s, pad).contiguous(memory_format=torch.channels_last)
        ops.norm_act(y.permute(0, 2, 3, 1), affine=(ones, b), relu=True)
        return y
    def fused(i):
        return torch.cudnn_convolution_relu(xs[i % 3], w, b, (s, s), (pad, pad), (1, 1), 1)
    def plain(i):
        return F.conv2d(xs[i % 3], w, None, s, pad)
    with torch.no_grad():
        t0, t1, t2 = timeit(plain), timeit(ours), timeit(fused)
        y0, y1 = ours(0), fused(0)
        z = torch.randn_like(y0)
        def addrelu(i):
            return torch.cudnn_convolution_add_relu(xs[i % 3], w, z, 1.0, b, (s, s), (pad, pad), (1, 1), 1)
        t3 = timeit(addrelu)
    err = float((y0 - y1).abs().max() / y0.abs().max())
    print(f"{cin:3d}->{cout:3d} k{k} s{s} @{hw:3d}: conv {t0:6.1f}  conv+ours {t1:6.1f}  cudnn fused {t2:6.1f}  add_relu {t3:6.1f} us   rel diff {err:.1e}  cl={y1.is_contiguous(memory_format=torch.channels_last)}")
