"""Time scflow_corr_pyramid_bwd against the two torch.bmm calls it replaced (round 1), same inputs, CUDA events."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from scflow_b200 import ops

def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

def main():
    n, c, h, w = (int(v) for v in (sys.argv[1:5] or (32, 256, 32, 32)))
    P = h * w
    dev = torch.device("cuda:0")
    f1 = torch.randn(n, c, h, w, device=dev); f2 = torch.randn(n, c, h, w, device=dev)
    G = torch.randn(n, P, P, device=dev)
    flops = 2 * 2.0 * n * c * P * P
    out = {"shape": [n, c, h, w]}
    ms = timeit(lambda: ops.corr_pyramid_bwd(G, f1, f2, True, True, ops.PREC_TF32))
    out["tcgen05_tf32"] = {"ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1), "gbps": round(4.0 * n * (2 * P * P + 4 * c * P) / ms / 1e6, 1)}
    ms = timeit(lambda: ops.corr_pyramid_bwd(G, f1, f2, True, True, ops.PREC_FP32), reps=5)
    out["fma_fp32"] = {"ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1)}
    a1, a2 = f1.reshape(n, c, P), f2.reshape(n, c, P)
    for name, allow in (("bmm_tf32", True), ("bmm_fp32", False)):
        torch.backends.cuda.matmul.allow_tf32 = allow
        ms = timeit(lambda: (torch.bmm(a2, G.transpose(1, 2)), torch.bmm(a1, G)))
        out[name] = {"ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1)}
    print(json.dumps(out))

if __name__ == "__main__":
    main()
