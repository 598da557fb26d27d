"""Where the pose error of the cfg1 refiner fixture comes from: final-pose error vs the golden
reference for combinations of {fused encoder glue, TF32 correlation, tcgen05 GRU}."""
import os, sys, json, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import scflow_oracle as O
from scflow_b200 import synthetic as syn, gru, encoder
from scflow_b200.refiner import SCFlowRefinerCore, decoder_cfg

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
gd = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
g = np.load(os.path.join(gd, "refiner_cfg1.npz"))
es = json.load(open(os.path.join(gd, "encoder_state_shapes.json")))
Rg, tg = torch.from_numpy(g["rotations"]), torch.from_numpy(g["translations"])
cr = {k: v.cuda() for k, v in syn.make_crops(1, 256, seed=63).items()}
orig_ok = encoder.RAFTEncoder._fused_ok
for fused, prec, backend in itertools.product((False, True), ("fp32", "tf32"), ("cudnn", "tcgen05")):
    encoder.RAFTEncoder._fused_ok = orig_ok if fused else (lambda self, x: False)
    gru.BACKEND = backend
    core = SCFlowRefinerCore(256, decoder=decoder_cfg(feat_size=(32, 32), corr_precision=prec)).eval()
    core.encoder.load_state_dict(syn.fill_state_dict(es["IN"], seed=61, gain=1.7))
    core.context.load_state_dict(syn.fill_state_dict(es["BN"], seed=62, gain=1.7))
    shp = {k: tuple(v.shape) for k, v in core.decoder.state_dict().items()}
    core.decoder.load_state_dict(syn.fill_state_dict(shp, seed=64))
    core.cuda()
    with torch.no_grad():
      out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"])
    er = [float(O.rotation_geodesic(out[2][i].cpu(), Rg[i]).max()) for i in range(8)]
    et = float((out[3][-1].cpu() - tg[-1]).abs().max())
    print(f"fused_enc={fused!s:5} corr={prec} gru={backend:8s} rot err per iter: " + " ".join(f"{e:.1e}" for e in er) + f"  | t err {et:.2e} mm")
