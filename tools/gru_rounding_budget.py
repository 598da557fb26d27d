"""Where does the TF32 ConvGRU's pose error come from: rounding the weights or rounding the activations?
The GRU runs through cuDNN in plain fp32 with operands rounded to TF32 by hand (emulating the tensor-core kernel);
everything else is fp32 (correlation in its exact mode)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["SCFLOW_GRU_BACKEND"] = "cudnn"
import numpy as np, torch
from oracle import scflow_oracle as O
from scflow_b200 import synthetic as syn, gru
from scflow_b200.refiner import SCFlowRefinerCore, decoder_cfg
gd = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
es = json.load(open(os.path.join(gd, "encoder_state_shapes.json")))
torch.backends.cudnn.allow_tf32 = False
rn = lambda t: ((t.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)
def make_forward(round_w, round_a, split_a=False):
    import torch.nn.functional as F
    def conv(cm, x):
        c = cm.conv
        w = rn(c.weight.detach()) if round_w else c.weight
        if round_a and split_a:
            hi = rn(x); lo = rn(x - hi)
            y = F.conv2d(hi, w, c.bias, 1, c.padding) + F.conv2d(lo, w, None, 1, c.padding)
        else:
            y = F.conv2d(rn(x) if round_a else x, w, c.bias, 1, c.padding)
        return cm.activate(y)
    def fwd(mod, h, x):
        for cz, cr, cq in zip(mod.conv_z, mod.conv_r, mod.conv_q):
            hx = torch.cat([h, x], dim=1)
            z = conv(cz, hx); r = conv(cr, hx)
            q = conv(cq, torch.cat([r * h, x], dim=1))
            h = (1 - z) * h + z * q
        return h
    return fwd
for tag, s, seed in (("cfg1", 256, 63), ("cfg4", 512, 65)):
    g = np.load(os.path.join(gd, f"refiner_{tag}.npz"))
    core = SCFlowRefinerCore(s, decoder=decoder_cfg(feat_size=(s // 8, s // 8), corr_precision="fp32")).eval()
    core.encoder.load_state_dict(syn.fill_state_dict(es["IN"], seed=61, gain=1.7))
    core.context.load_state_dict(syn.fill_state_dict(es["BN"], seed=62, gain=1.7))
    shp = {k: tuple(v.shape) for k, v in core.decoder.state_dict().items()}
    core.decoder.load_state_dict(syn.fill_state_dict(shp, seed=64))
    core.cuda()
    cr = {k: v.cuda() for k, v in syn.make_crops(1, s, seed=seed).items()}
    Rg, tg = torch.from_numpy(g["rotations"]), torch.from_numpy(g["translations"])
    for name, args in (("fp32", (False, False)), ("weights rounded", (True, False)), ("activations rounded", (False, True)),
                       ("both rounded", (True, True)), ("weights rounded, activations hi+lo", (True, True, True))):
        gru._cudnn_forward = make_forward(*args)
        with torch.no_grad():
            out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"])
        er = float(O.rotation_geodesic(out[2][-1].cpu(), Rg[-1]).max()); et = float((out[3][-1].cpu() - tg[-1]).abs().max())
        print(f"{tag} GRU {name:36s}: {er:.3e} rad, {et:.3e} mm")
