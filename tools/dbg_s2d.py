import os, sys, torch, torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scflow_b200 import encoder as E
torch.manual_seed(0)
dev = "cuda"
for norm in ("IN", "BN"):
    enc = E.RAFTEncoder(3, 256, norm_cfg=dict(type=norm)).to(dev).eval()
    if norm == "BN":
        for m in enc.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.normal_(); m.running_var.uniform_(0.5, 2.0)
    x = torch.rand(2, 3, 256, 256, device=dev)
    for tf in (False, True):
        torch.backends.cudnn.allow_tf32 = True
        outs = {}
        for s in ("0", "1"):
            os.environ["SCFLOW_STEM_S2D"] = s
            with torch.no_grad():
                outs[s] = enc(x).clone()
        d = (outs["0"] - outs["1"]).abs().max().item(); m = outs["0"].abs().max().item()
        print(norm, "enc out diff", d, "max", m)
    # stem only
    w = enc.conv1.weight.detach()
    ref = F.conv2d(x, w, None, 2, 3)
    got = F.conv2d(E._s2d_input(x), E._s2d_filter(w))
    print(norm, "stem diff", (ref - got).abs().max().item(), ref.abs().max().item(), got.shape, ref.shape)
    torch.backends.cudnn.allow_tf32 = False
    ref = F.conv2d(x, w, None, 2, 3)
    got = F.conv2d(E._s2d_input(x), E._s2d_filter(w))
    print(norm, "stem diff fp32", (ref - got).abs().max().item())
