"""clock64 stamps of CTA 0 for its first items of the fused ConvGRU launch (SCFLOW_GRU_DBG_TS):
where an item's time goes.  python tools/gru_stamps.py [bk]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SCFLOW_GRU_DBG_TS"] = "1"
if len(sys.argv) > 1:
    os.environ["SCFLOW_GRU_BK"] = sys.argv[1]
import torch
from scflow_b200 import gru_tc, modules
dev = "cuda"; N, h, w = 64, 32, 32
g = torch.Generator(device=dev).manual_seed(0)
mod = modules.ConvGRU(128, 256).to(dev)
hs = torch.randn(N, h, w, 128, device=dev, generator=g)
xs = gru_tc.round_tf32(torch.randn(N, h, w, 256, device=dev, generator=g))
hr = gru_tc.round_tf32(hs); rh = torch.empty_like(hs)
for _ in range(3):
    gru_tc.update_inplace(mod, hs, hr, xs, rh)
torch.cuda.synchronize()
ws = mod.__dict__["_scflow_gru_ws"]
off = ((N * 4 + 255) // 256) * 256 // 4
ts = ws[off: off + 64].view(torch.int64).cpu().reshape(4, 8)
names = ["item start (TMEM free)", "S1 issued", "s1_done seen by epilogue", "E1 done (rh published)",
         "rh_ready seen by producer", "S2 issued", "s2_done seen by epilogue", "TMEM handed back"]
t0 = int(ts[0, 0])
for i in range(4):
    print(f"item {i}:")
    for k in (0, 1, 2, 3, 4, 5, 6, 7):
        print(f"   {names[k]:28s} {(int(ts[i, k]) - t0) / 1.965e3:9.2f} us")
