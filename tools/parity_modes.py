"""Final-pose error of the whole refiner against the golden fixtures (unmodified reference, fp32 CPU) in the two
cuDNN modes: strict fp32 (what tests/ enforce) and TF32 allowed (PyTorch's GPU default, what bench.py times)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import scflow_oracle as O
from scflow_b200 import synthetic as syn
from scflow_b200.refiner import SCFlowRefinerCore
gd = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
es = json.load(open(os.path.join(gd, "encoder_state_shapes.json")))
for tag, s, seed in (("cfg1", 256, 63), ("cfg4", 512, 65)):
    g = np.load(os.path.join(gd, f"refiner_{tag}.npz"))
    core = SCFlowRefinerCore(s).eval()
    core.encoder.load_state_dict(syn.fill_state_dict(es["IN"], seed=61, gain=1.7))
    core.context.load_state_dict(syn.fill_state_dict(es["BN"], seed=62, gain=1.7))
    shp = {k: tuple(v.shape) for k, v in core.decoder.state_dict().items()}
    core.decoder.load_state_dict(syn.fill_state_dict(shp, seed=64))
    core.cuda()
    cr = {k: v.cuda() for k, v in syn.make_crops(1, s, seed=seed).items()}
    Rg, tg = torch.from_numpy(g["rotations"]), torch.from_numpy(g["translations"])
    import torch.nn.functional as F
    from scflow_b200 import gru_tc
    orig = (F.conv2d, torch.cudnn_convolution_relu, torch.cudnn_convolution_add_relu)
    rn = lambda t: ((t.detach().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)   # cvt.rna.tf32
    for mode in (False, True, "rn"):
        torch.backends.cudnn.allow_tf32 = bool(mode)
        if mode == "rn":      # experiment: operands rounded to nearest TF32 before every cuDNN convolution
            F.conv2d = lambda x, w, *a, **k: orig[0](rn(x), rn(w), *a, **k)
            torch.cudnn_convolution_relu = lambda x, w, *a, **k: orig[1](rn(x), rn(w), *a, **k)
            torch.cudnn_convolution_add_relu = lambda x, w, *a, **k: orig[2](rn(x), rn(w), *a, **k)
        with torch.no_grad():
            out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"])
        er = float(O.rotation_geodesic(out[2][-1].cpu(), Rg[-1]).max()); et = float((out[3][-1].cpu() - tg[-1]).abs().max())
        print(f"{tag} cudnn.allow_tf32={mode}: {er:.3e} rad, {et:.3e} mm")
    F.conv2d, torch.cudnn_convolution_relu, torch.cudnn_convolution_add_relu = orig
