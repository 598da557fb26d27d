"""Which cuDNN layer groups' TF32 arithmetic moves the final pose?  Everything runs with cuDNN in fp32 except one
group at a time (TF32 allowed inside that group's forward only); golden fixtures as in tools/parity_modes.py."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import scflow_oracle as O
from scflow_b200 import synthetic as syn
from scflow_b200.refiner import SCFlowRefinerCore
gd = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
es = json.load(open(os.path.join(gd, "encoder_state_shapes.json")))
GROUPS = {"none": [], "encoder(IN)": ["encoder"], "context(BN)": ["context"], "motion": ["decoder.encoder"],
          "xheads": ["decoder.flow_pred", "decoder.mask_pred"], "dflow/mask enc": ["decoder.delta_flow_encoder", "decoder.mask_encoder"],
          "pose head": ["decoder.pose_pred"], "all decoder": ["decoder"], "all": ["encoder", "context", "decoder"]}
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
    mods = dict(core.named_modules())
    for name, prefixes in GROUPS.items():
        hooks = []
        depth = [0]
        def pre(m, i):
            depth[0] += 1; torch.backends.cudnn.allow_tf32 = True
        def post(m, i, o):
            depth[0] -= 1
            if depth[0] == 0: torch.backends.cudnn.allow_tf32 = False
        for p in prefixes:
            hooks.append(mods[p].register_forward_pre_hook(pre)); hooks.append(mods[p].register_forward_hook(post))
        torch.backends.cudnn.allow_tf32 = False
        with torch.no_grad():
            out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"])
        for h in hooks: h.remove()
        er = float(O.rotation_geodesic(out[2][-1].cpu(), Rg[-1]).max()); et = float((out[3][-1].cpu() - tg[-1]).abs().max())
        print(f"{tag} TF32 in {name:16s}: {er:.3e} rad, {et:.3e} mm")
