"""Time scflow_corr_lookup_bwd alone (CUDA events, inputs larger than L2): algorithmic bytes = the pyramid gradient
written once (1360 floats per query at 32x32) + the 324-channel gradient read once."""
import sys, os, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from scflow_b200 import _lib, ops
from scflow_b200.ops import _ptr, _stream
from scflow_b200.autograd import _ptr_array

def main():
    n, h, w = (int(v) for v in (sys.argv[1:4] or (32, 32, 32)))
    dev = torch.device("cuda:0")
    lib = _lib.load()
    P = h * w
    flow = (torch.rand(n, 2, h, w, device=dev) - 0.5) * 8
    out = {"shape": [n, h, w]}
    for name, layout in (("nchw", ops.LAYOUT_NCHW), ("nhwc", ops.LAYOUT_NHWC)):
        g = torch.randn(n, 324, h, w, device=dev) if layout == ops.LAYOUT_NCHW else torch.randn(n, h, w, 324, device=dev)
        glv = [torch.empty(n * P, 1, h >> l, w >> l, device=dev) for l in range(4)]
        nbytes = 4 * (g.numel() + sum(t.numel() for t in glv))
        arr = _ptr_array(glv)
        def run():
            rc = lib.scflow_corr_lookup_bwd(_ptr(g), layout, _ptr(flow), arr, 4, n, h, w, 4, ops.DIV_IEEE, _stream())
            assert rc == 0, rc
        for _ in range(3): run()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20): run()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 20
        out[name] = {"ms": round(ms, 4), "gbps": round(nbytes / ms / 1e6, 1), "bytes": nbytes}
    print(json.dumps(out))

if __name__ == "__main__":
    main()
