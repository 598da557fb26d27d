"""Debug helper: run the NHWC lookup on one small case and compare with the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import scflow_oracle as O
from oracle.gen_golden import lookup_flow
from scflow_b200 import ops, synthetic as syn

n, h, w = [int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (2, 8, 16))]
ft = syn.make_features(n, h, w, seed=11, c=16, hc=4)
pyr = O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)
flow = lookup_flow(n, h, w, 12)
want, wwin = O.corr_lookup(pyr, flow, 4, 0, return_windows=True)
for cl in (True, False):
    got, gwin = ops.corr_lookup([p.cuda() for p in pyr], flow.cuda(), 4, 0, return_windows=True, channels_last=cl)
    torch.cuda.synchronize()
    err = (got.cpu() - want).abs().max().item()
    print("channels_last", cl, "max abs err", err, "ref max", want.abs().max().item())
    for l in range(4):
        for ax, size in ((0, w >> l), (1, h >> l)):
            a = wwin[:, l, ax].clamp(-32, size + 32).to(torch.int16)
            print("  level", l, "axis", ax, "windows equal", torch.equal(gwin[:, l, ax].cpu(), a))
    if err > 1e-5:
        d = (got.cpu() - want).abs()
        idx = (d > 1e-5).nonzero()
        print("  bad count", idx.shape[0], "first", idx[:8].tolist())
