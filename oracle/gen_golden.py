"""Generate tests/golden/*.npz by running the UNMODIFIED reference (dev container only).

    python -m oracle.gen_golden

Imports the reference's own hot-path files from /root/reference through oracle/_ref_shims.py,
feeds them the seeded inputs of scflow_b200/synthetic.py and stores the outputs.  The fixtures
hold only what cannot be regenerated from a seed (reference OUTPUTS, plus small inputs where
that is cheaper than a seed), so they stay small.  TEST INFRASTRUCTURE.
"""
import copy
import json
import os
import sys

import numpy as np
import torch

from oracle import _ref_shims as S
from scflow_b200 import synthetic as syn

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def lookup_flow(n, h, w, seed):
    """Flow field exercising integer coordinates, near-integers, far out-of-bounds."""
    rs = np.random.RandomState(seed)
    f = rs.uniform(-6, 6, size=(n, 2, h, w))
    f[:, :, 0, :] = np.round(f[:, :, 0, :])            # exact integer coords
    f[:, :, 1, :] = 0.0                                 # zero flow
    f[:, :, 2, :] = np.round(f[:, :, 2, :] * 2) / 2     # half-integers
    f[:, :, 3, : w // 2] = rs.uniform(-40, 40, size=(n, 2, w // 2))  # far out of bounds
    return torch.from_numpy(f.astype(np.float32))


def shapes_of(m):
    return {k: tuple(v.shape) for k, v in m.state_dict().items()}


def refiner_fixture(R, enc, cxt, s, tag, seed):
    """Whole refiner forward of the unmodified reference on one seeded s x s crop."""
    cr = syn.make_crops(1, s, seed=seed)
    cfg = copy.deepcopy(S.DECODER_CFG)
    cfg["pose_head_cfg"]["feat_size"] = (s // 8, s // 8)
    dec = R.SCFlowDecoder(**cfg).eval()
    dec.load_state_dict(syn.fill_state_dict(shapes_of(dec), seed=64))
    with torch.no_grad():
        fr = enc(cr["real_images"]); fn = enc(cr["rendered_images"]); cx = cxt(cr["rendered_images"])
        hf, cf = torch.split(cx, [128, 128], 1)
        o = dec(fn, fr, torch.tanh(hf), torch.relu(cf), cr["R0"], cr["t0"], cr["depth"], cr["K"],
                cr["labels"], torch.zeros(1, 2, s, s), 0.)
    st = 4 * s // 256
    np.savez_compressed(os.path.join(OUT, f"refiner_{tag}.npz"),
                        feat_real=fr[:, ::16, ::st, ::st].numpy(), feat_render=fn[:, ::16, ::st, ::st].numpy(),
                        cxt=cx[:, ::16, ::st, ::st].numpy(),
                        rotations=torch.stack(o[2]).numpy(), translations=torch.stack(o[3]).numpy())


GRAD_KEYS = ("gru.conv_z.0.conv.bias", "gru.conv_q.1.conv.bias", "flow_pred.predict_layer.weight",
             "encoder.corr_net.0.conv.bias", "pose_pred.rotation_pred.weight")


def decoder_grad_fixture(R, n=2, s=64, seed=41):
    """Gradients of scflow_b200.synthetic.synthetic_loss through the unmodified reference decoder
    (shipped detach flags) wrt its four feature inputs and its parameters."""
    cfg = copy.deepcopy(S.DECODER_CFG)
    hh = s // 8
    cfg["pose_head_cfg"]["feat_size"] = (hh, hh)
    dec = R.SCFlowDecoder(**cfg).train()
    dec.load_state_dict(syn.fill_state_dict(shapes_of(dec), seed=seed))
    cr = syn.make_crops(n, s, seed=seed + 1)
    ft = syn.make_features(n, hh, hh, seed=seed + 2)
    ins = [ft[k].clone().requires_grad_(True) for k in ("feat_render", "feat_real", "h_feat", "cxt_feat")]
    o = dec(ins[0], ins[1], ins[2], ins[3], cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"],
            torch.zeros(n, 2, s, s), 0.)
    loss = syn.synthetic_loss(o)
    loss.backward()
    d = {"loss": np.float64(loss.item())}
    for k, t in zip(("feat_render", "feat_real", "h_feat", "cxt_feat"), ins):
        d["g_" + k] = t.grad[:, ::4].numpy()
    names, norms = [], []
    for k, p in dec.named_parameters():
        names.append(k)
        norms.append(0.0 if p.grad is None else float(p.grad.double().norm()))
        if k in GRAD_KEYS:
            d["p_" + k] = p.grad.numpy()
    d["param_names"] = np.array(names)
    d["param_grad_norms"] = np.array(norms)
    np.savez_compressed(os.path.join(OUT, "decoder_s64_grads.npz"), **d)


def encoders(R):
    enc = R.RAFTEncoder(in_channels=3, out_channels=256, net_type="Basic", norm_cfg=dict(type="IN")).eval()
    cxt = R.RAFTEncoder(in_channels=3, out_channels=256, net_type="Basic", norm_cfg=dict(type="BN")).eval()
    se, sc = shapes_of(enc), shapes_of(cxt)
    enc.load_state_dict(syn.fill_state_dict(se, seed=61, gain=1.7))
    cxt.load_state_dict(syn.fill_state_dict(sc, seed=62, gain=1.7))
    return enc, cxt, se, sc


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    R = S.load_reference()
    if sys.argv[1:] == ["grads"]:         # only the training-gradient fixture
        decoder_grad_fixture(R)
        return
    if sys.argv[1:] == ["cfg4"]:          # only the 512^2 refiner fixture (the others stay as committed)
        enc, cxt, _, _ = encoders(R)
        refiner_fixture(R, enc, cxt, 512, "cfg4", 65)
        return

    # ---- 1. pyramid + lookup, small non-square map (H=8 -> last level H=1: max(H-1,1) branch)
    for tag, (n, c, h, w) in {"small": (2, 8, 8, 16), "sq32": (1, 16, 32, 32)}.items():
        ft = syn.make_features(n, h, w, seed=11, c=c, hc=4)
        flow = lookup_flow(n, h, w, 12)
        pyr = R.CorrelationPyramid(4)(ft["feat_render"], ft["feat_real"])
        out = R.CorrLookup(radius=4, align_corners=True)(pyr, flow.clone())
        if tag == "small":
            extra = {f"pyr{i}": p.numpy() for i, p in enumerate(pyr)}
            o = out.numpy()
        else:   # larger map: keep the fixture small (inputs come from the seed)
            extra = {}
            o = out[:, :, ::2, ::2].numpy()
        np.savez_compressed(os.path.join(OUT, f"lookup_{tag}.npz"),
                            shape=np.array([n, c, h, w]), flow=flow.numpy(), out=o, **extra)

    # ---- 2. ConvGRU, small channel counts
    gru = R.ConvGRU(8, 16, net_type="SeqConv")
    sd = syn.fill_state_dict(shapes_of(gru), seed=21, gain=2.0)
    gru.load_state_dict(sd)
    rs = np.random.RandomState(22)
    h = torch.from_numpy(np.tanh(rs.standard_normal((2, 8, 6, 10))).astype(np.float32))
    x = torch.from_numpy(rs.standard_normal((2, 16, 6, 10)).astype(np.float32))
    with torch.no_grad():
        hn = gru(h, x)
    np.savez_compressed(os.path.join(OUT, "gru_small.npz"), h=h.numpy(), x=x.numpy(), out=hn.numpy())

    # ---- 3. pose utilities
    cr = syn.make_crops(3, 32, seed=31)
    P = R.pose
    p2, p3 = [], []
    dense3 = np.zeros((3, 3, 32, 32), np.float32)
    for i in range(3):
        a, b = P.cal_3d_2d_corr(cr["depth"][i], cr["K"][i], cr["R0"][i], cr["t0"][i])
        p2.append(a); p3.append(b)
        dense3[i][:, a[:, 1].long(), a[:, 0].long()] = b.t().numpy()
    rs = np.random.RandomState(32)
    dR6 = torch.from_numpy((np.tile([1., 0, 0, 0, 1., 0], (3, 1)) + 0.05 * rs.standard_normal((3, 6))).astype(np.float32))
    dt = torch.from_numpy((0.05 * rs.standard_normal((3, 3))).astype(np.float32))
    Rn, tn = P.get_pose_from_delta_pose(dR6, dt, cr["R0"], cr["t0"], depth_transform="exp", detach_depth_for_xy=True)
    flow = P.get_flow_from_delta_pose_and_points(Rn, tn, cr["K"], p2, p3, 32, 32, invalid_num=0.)
    flow400 = P.get_flow_from_delta_pose_and_points(Rn, tn, cr["K"], p2, p3, 32, 32, invalid_num=400.)
    np.savez_compressed(os.path.join(OUT, "pose_small.npz"), points3d=dense3, dR6=dR6.numpy(), dt=dt.numpy(),
                        R=Rn.numpy(), t=tn.numpy(), flow=flow.numpy(), flow400=flow400.numpy())

    # ---- 4. full decoder, (a) S=64 N=2 (pose head feat_size 8x8)  (b) cfg1: S=256 N=1
    def run_decoder(n, s, tag, seed):
        cfg = copy.deepcopy(S.DECODER_CFG)
        hh = s // 8
        cfg["pose_head_cfg"]["feat_size"] = (hh, hh)
        dec = R.SCFlowDecoder(**cfg).eval()
        shp = shapes_of(dec)
        dec.load_state_dict(syn.fill_state_dict(shp, seed=seed))
        cr = syn.make_crops(n, s, seed=seed + 1)
        ft = syn.make_features(n, hh, hh, seed=seed + 2)
        with torch.no_grad():
            o = dec(ft["feat_render"], ft["feat_real"], ft["h_feat"], ft["cxt_feat"], cr["R0"], cr["t0"],
                    cr["depth"], cr["K"], cr["labels"], torch.zeros(n, 2, s, s), 0.)
        names = ["flow_from_pose", "flow_from_pred", "rotations", "translations", "masks",
                 "delta_rotations", "delta_translations"]
        d = {}
        for nm, lst in zip(names, o):
            st = torch.stack(lst).numpy()
            if st.ndim == 5:
                st = st[[0, -1]][..., ::4, ::4]      # first/last iteration, every 4th pixel
            d[nm] = st
        np.savez_compressed(os.path.join(OUT, f"decoder_{tag}.npz"), n=n, s=s, seed=seed, **d)
        return shp

    run_decoder(2, 64, "s64", 41)
    shp = run_decoder(1, 256, "cfg1", 51)
    with open(os.path.join(OUT, "decoder_state_shapes.json"), "w") as f:
        json.dump({k: list(v) for k, v in shp.items()}, f, indent=0, sort_keys=True)

    # ---- 5. RAFT encoders + whole refiner forward (cfg1: N=1, 256^2)
    enc, cxt, se, sc = encoders(R)
    with open(os.path.join(OUT, "encoder_state_shapes.json"), "w") as f:
        json.dump({"IN": {k: list(v) for k, v in se.items()}, "BN": {k: list(v) for k, v in sc.items()}},
                  f, indent=0, sort_keys=True)
    refiner_fixture(R, enc, cxt, 256, "cfg1", 63)
    refiner_fixture(R, enc, cxt, 512, "cfg4", 65)
    decoder_grad_fixture(R)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
