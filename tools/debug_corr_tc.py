"""Diagnostics for the tcgen05 correlation kernel (dev aid, not a test)."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scflow_b200 import _lib, ops, synthetic as syn

lib = _lib.load()
n, c, h, w = 1, 256, 32, 32
ft = syn.make_features(n, h, w, seed=5, c=c, hc=4)
f1, f2 = ft["feat_render"].cuda(), ft["feat_real"].cuda()
P = h * w
ref = torch.matmul(f1.reshape(n, c, P).transpose(1, 2), f2.reshape(n, c, P)) / (c ** 0.5)
levels = [torch.full((n * P, 1, h >> l, w >> l), 777.0, device="cuda") for l in range(4)]
arr = (ctypes.c_void_p * 4)(*[l.data_ptr() for l in levels])
rc = lib.scflow_corr_pyramid_fwd(f1.data_ptr(), f2.data_ptr(), arr, 4, n, c, h, w, 1, None, 0,
                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
torch.cuda.synchronize()
print("rc", rc)
l0 = levels[0].reshape(n, P, P)
print("sentinel frac", float((l0 == 777.0).float().mean()), "zero frac", float((l0 == 0).float().mean()))
print("got[0,:2,:8]", l0[0, :2, :8].tolist())
print("ref[0,:2,:8]", ref[0, :2, :8].tolist())
err = (l0 - ref).abs().max() / ref.abs().max()
print("rel err", float(err))
# does got match ref under some permutation?  check row/col correlation
g = l0[0]; r = ref[0]
print("got abs mean", float(g.abs().mean()), "ref abs mean", float(r.abs().mean()))
if float(g.abs().mean()) > 0:
    # try transposed
    print("err vs ref^T", float((g - r.t()).abs().max() / r.abs().max()))
