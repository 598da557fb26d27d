"""Pin oracle/scflow_oracle.py (the CPU restatement) against outputs of the UNMODIFIED
reference stored in tests/golden/ by oracle/gen_golden.py.  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import scflow_oracle as O
from oracle.gen_golden import lookup_flow
from scflow_b200 import synthetic as syn


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def _t(a):
    return torch.from_numpy(np.asarray(a))


def rel_err(a, b):
    a, b = _t(a).double(), _t(b).double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def test_pyramid_and_lookup_small(golden_dir):
    g = _load(golden_dir, "lookup_small.npz")
    n, c, h, w = g["shape"]
    ft = syn.make_features(n, h, w, seed=11, c=c, hc=4)
    pyr = O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)
    for i, p in enumerate(pyr):
        assert p.shape == g[f"pyr{i}"].shape
        assert rel_err(p, g[f"pyr{i}"]) < 1e-6
    flow = _t(g["flow"])
    assert torch.equal(flow, lookup_flow(n, h, w, 12))
    # on the reference's own pyramid: the restated bilinear gather must agree to fp32 rounding
    ref_pyr = [_t(g[f"pyr{i}"]) for i in range(4)]
    out = O.corr_lookup(ref_pyr, flow, 4)
    assert out.shape == g["out"].shape
    assert rel_err(out, g["out"]) < 2e-6


def test_lookup_sq32(golden_dir):
    g = _load(golden_dir, "lookup_sq32.npz")
    n, c, h, w = g["shape"]
    ft = syn.make_features(n, h, w, seed=11, c=c, hc=4)
    pyr = O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)
    out = O.corr_lookup(pyr, _t(g["flow"]), 4)
    assert rel_err(out[:, :, ::2, ::2], g["out"]) < 5e-6


def test_lookup_impulse_channel_order():
    """SURVEY 3.4 probe: a unit impulse at (y+1, x+3) must light channel 9*a+b = 9*7+5 = 68."""
    h = w = 16
    q_y, q_x = 5, 6
    lvl0 = torch.zeros(h * w, 1, h, w)
    lvl0[q_y * w + q_x, 0, q_y + 1, q_x + 3] = 1.0
    pyr = [lvl0] + [torch.zeros(h * w, 1, h >> i, w >> i) for i in (1, 2, 3)]
    out = O.corr_lookup(pyr, torch.zeros(1, 2, h, w), 4)
    ch = out[0, :81, q_y, q_x]
    assert int(ch.argmax()) == 68 and abs(float(ch[68]) - 1.0) < 1e-6


def test_window_roundtrip_not_identity():
    """at W=16 the normalise/un-normalise round trip lands integer coordinates
    1,2,3 just below the integer; the oracle must replay that, not 'fix' it."""
    c = torch.arange(0, 16, dtype=torch.float32)
    ix = O.lookup_axis_coords(c, 16, 0, 0)[:, 0]
    assert (torch.floor(ix) != c).any()
    assert float((ix - c).abs().max()) < 2e-6


def test_conv_gru_small(golden_dir):
    g = _load(golden_dir, "gru_small.npz")
    shapes = {}
    for nm in "zrq":
        for p, k in ((0, (1, 5)), (1, (5, 1))):
            shapes[f"conv_{nm}.{p}.conv.weight"] = (8, 24) + k
            shapes[f"conv_{nm}.{p}.conv.bias"] = (8,)
    sd = syn.fill_state_dict(shapes, seed=21, gain=2.0)
    out = O.conv_gru(_t(g["h"]), _t(g["x"]), sd)
    assert rel_err(out, g["out"]) < 1e-6


def test_pose_small(golden_dir):
    g = _load(golden_dir, "pose_small.npz")
    cr = syn.make_crops(3, 32, seed=31)
    pts, mask = O.lift_points_dense(cr["depth"], cr["K"], cr["R0"], cr["t0"])
    assert rel_err(pts, g["points3d"]) < 1e-5
    for i in range(3):   # ragged form agrees with the dense one
        p2, p3 = O.lift_points(cr["depth"][i], cr["K"][i], cr["R0"][i], cr["t0"][i])
        assert int(mask[i].sum()) == p2.shape[0]
        d = pts[i][:, p2[:, 1].long(), p2[:, 0].long()].t()
        assert rel_err(d, p3) < 1e-5
    R, t = O.pose_compose(_t(g["dR6"]), _t(g["dt"]), cr["R0"], cr["t0"])
    assert float((R - _t(g["R"])).abs().max()) < 1e-6
    assert rel_err(t, g["t"]) < 1e-6
    for inv, key in ((0.0, "flow"), (400.0, "flow400")):
        fl = O.pose_flow_dense(R, t, cr["K"], pts, mask, inv)
        assert float((fl - _t(g[key])).abs().max()) < 2e-3 * max(1.0, float(np.abs(g["flow"]).max()))


def test_resize_matches_torch():
    x = torch.randn(2, 2, 64, 64, generator=torch.Generator().manual_seed(0))
    import torch.nn.functional as F
    d = O.resize_bilinear_ac(x, 8, 8)
    assert torch.allclose(d, F.interpolate(x, scale_factor=(1 / 8, 1 / 8), mode="bilinear", align_corners=True), atol=1e-6)
    u = O.resize_bilinear_ac(d, 64, 64)
    assert torch.allclose(u, F.interpolate(d, scale_factor=(8, 8), mode="bilinear", align_corners=True), atol=1e-6)


def _decoder_inputs(n, s, seed, golden_dir):
    with open(os.path.join(golden_dir, "decoder_state_shapes.json")) as f:
        shp = {k: tuple(v) for k, v in json.load(f).items()}
    hh = s // 8
    shp["pose_pred.fc_layers.0.0.weight"] = (1024, 128 * (hh // 8) ** 2)
    sd = syn.fill_state_dict(shp, seed=seed)
    cr = syn.make_crops(n, s, seed=seed + 1)
    ft = syn.make_features(n, hh, hh, seed=seed + 2)
    return sd, cr, ft


@pytest.mark.parametrize("tag,n,s,seed", [("s64", 2, 64, 41), ("cfg1", 1, 256, 51)])
def test_decoder_forward(golden_dir, tag, n, s, seed):
    g = _load(golden_dir, f"decoder_{tag}.npz")
    sd, cr, ft = _decoder_inputs(n, s, seed, golden_dir)
    o = O.decoder_forward(sd, ft["feat_render"], ft["feat_real"], ft["h_feat"], ft["cxt_feat"],
                          cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"],
                          torch.zeros(n, 2, s, s), 0.0, iters=8)
    Rg, tg = _t(g["rotations"]), _t(g["translations"])
    # the pose must actually move, otherwise the comparison proves nothing
    assert float(O.rotation_geodesic(Rg[-1], cr["R0"]).max()) > 1e-2
    R, t = torch.stack(o["rotations"]), torch.stack(o["translations"])
    assert float(O.rotation_geodesic(R[-1], Rg[-1]).max()) < 2e-5
    assert float((t[-1] - tg[-1]).abs().max()) < 2e-2          # mm
    assert rel_err(torch.stack(o["delta_rotations"]), g["delta_rotations"]) < 1e-4
    for nm in ("flow_from_pose", "flow_from_pred", "masks"):
        st = torch.stack(o[nm])[[0, -1]][..., ::4, ::4]
        assert rel_err(st, g[nm]) < 1e-3, nm


def test_refiner_cfg1(golden_dir):
    g = _load(golden_dir, "refiner_cfg1.npz")
    with open(os.path.join(golden_dir, "encoder_state_shapes.json")) as f:
        es = json.load(f)
    with open(os.path.join(golden_dir, "decoder_state_shapes.json")) as f:
        ds = {k: tuple(v) for k, v in json.load(f).items()}
    se = syn.fill_state_dict(es["IN"], seed=61, gain=1.7)
    sc = syn.fill_state_dict(es["BN"], seed=62, gain=1.7)
    sd = syn.fill_state_dict(ds, seed=64)
    cr = syn.make_crops(1, 256, seed=63)
    fr = O.raft_encoder(cr["real_images"], se, "IN")
    assert rel_err(fr[:, ::16, ::4, ::4], g["feat_real"]) < 1e-4
    cx = O.raft_encoder(cr["rendered_images"], sc, "BN")
    assert rel_err(cx[:, ::16, ::4, ::4], g["cxt"]) < 1e-4
    o = O.refiner_forward(se, sc, sd, cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"],
                          cr["depth"], cr["K"], cr["labels"])
    Rg, tg = _t(g["rotations"]), _t(g["translations"])
    assert float(O.rotation_geodesic(torch.stack(o["rotations"])[-1], Rg[-1]).max()) < 5e-5
    assert float((torch.stack(o["translations"])[-1] - tg[-1]).abs().max()) < 5e-2


def test_refiner_cfg4_512(golden_dir):
    """BASELINE.json configs[3]: 512x512 crop, 64x64 feature map, pose head fed 8x8 (FC in 8192)."""
    g = _load(golden_dir, "refiner_cfg4.npz")
    with open(os.path.join(golden_dir, "encoder_state_shapes.json")) as f:
        es = json.load(f)
    with open(os.path.join(golden_dir, "decoder_state_shapes.json")) as f:
        ds = {k: tuple(v) for k, v in json.load(f).items()}
    ds["pose_pred.fc_layers.0.0.weight"] = (1024, 8192)
    se = syn.fill_state_dict(es["IN"], seed=61, gain=1.7)
    sc = syn.fill_state_dict(es["BN"], seed=62, gain=1.7)
    sd = syn.fill_state_dict(ds, seed=64)
    cr = syn.make_crops(1, 512, seed=65)
    o = O.refiner_forward(se, sc, sd, cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"],
                          cr["depth"], cr["K"], cr["labels"])
    Rg, tg = _t(g["rotations"]), _t(g["translations"])
    assert float(O.rotation_geodesic(Rg[-1], cr["R0"]).max()) > 1e-2          # the pose moved
    assert float(O.rotation_geodesic(torch.stack(o["rotations"])[-1], Rg[-1]).max()) < 5e-5
    assert float((torch.stack(o["translations"])[-1] - tg[-1]).abs().max()) < 5e-2


def test_decoder_gradients_s64(golden_dir):
    """Oracle autograd (train_detach = the shipped detach flags) against gradients taken through the
    unmodified reference decoder (oracle/gen_golden.py: decoder_grad_fixture)."""
    g = _load(golden_dir, "decoder_s64_grads.npz")
    with open(os.path.join(golden_dir, "decoder_state_shapes.json")) as f:
        ds = {k: tuple(v) for k, v in json.load(f).items()}
    ds["pose_pred.fc_layers.0.0.weight"] = (1024, 128)          # feat_size 8x8 -> 1x1 after 3 stride-2 convs
    sd = {k: v.clone().requires_grad_(True) for k, v in syn.fill_state_dict(ds, seed=41).items()}
    cr = syn.make_crops(2, 64, seed=42)
    ft = syn.make_features(2, 8, 8, seed=43)
    ins = {k: ft[k].clone().requires_grad_(True) for k in ("feat_render", "feat_real", "h_feat", "cxt_feat")}
    o = O.decoder_forward(sd, ins["feat_render"], ins["feat_real"], ins["h_feat"], ins["cxt_feat"],
                          cr["R0"], cr["t0"], cr["depth"], cr["K"], cr["labels"],
                          torch.zeros(2, 2, 64, 64), train_detach=True)
    loss = syn.synthetic_loss(o)
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * abs(float(g["loss"]))
    loss.backward()
    for k, t in ins.items():
        assert rel_err(t.grad[:, ::4], g["g_" + k]) < 2e-3, k
    for k in g.files:
        if k.startswith("p_"):
            assert rel_err(sd[k[2:]].grad, g[k]) < 2e-3, k
    norms = dict(zip([str(x) for x in g["param_names"]], g["param_grad_norms"]))
    for k, want in norms.items():
        got = 0.0 if sd[k].grad is None else float(sd[k].grad.double().norm())
        assert abs(got - want) <= 2e-3 * max(want, 1e-6), (k, got, want)


def test_add_oracle_equals_reference_golden(golden_dir):
    """oracle/render_oracle.add_errors (restatement of metrics/add.py:354-394 + datasets/pose.py:18-70) against the
    outputs of the UNMODIFIED reference method (oracle/gen_golden_add.py): identical in float64."""
    import numpy as np
    from oracle import render_oracle as RO
    g = np.load(os.path.join(golden_dir, "add_eval.npz"))
    vl = [g[f"verts{i}"].astype(np.float64) for i in range(3)]
    f = lambda k: g[k].astype(np.float64)
    e3n, e2, e3 = RO.add_errors(vl, f("gt_t"), f("gt_r"), f("pred_t"), f("pred_r"), g["labels"], f("k"),
                                {"cls_2": True}, list(g["diam"]))
    assert np.abs(e3 - g["e3"]).max() < 1e-12 and np.abs(e2 - g["e2"]).max() < 1e-12
    assert np.abs(e3n - g["e3n"]).max() < 1e-12


def test_render_oracle_known_answers():
    """A fronto-parallel square at depth 500: constant z-buffer, coverage equal to the projected area, ambient-only
    colour where the light (placed by rendering.py:203-206 behind the surface) adds nothing."""
    import numpy as np
    from oracle import render_oracle as RO
    verts = np.array([[-50, -50, 0], [50, -50, 0], [50, 50, 0], [-50, 50, 0]], np.float32)
    faces = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    colors = np.full((4, 3), 0.8, np.float32)
    K = np.array([[320.0, 0, 32], [0, 320.0, 32], [0, 0, 1]])
    img, z, fi = RO.render(verts, faces, colors, np.eye(3), np.array([0, 0, 500.0]), K, 64, 64)
    hit = z > 0
    assert abs(hit.sum() - 64 * 64) <= 2 * 64 + 4          # the square projects to exactly 64 x 64 image units
    assert np.allclose(z[hit], 500.0)
    assert np.allclose(img[:, ~hit], 0.5)
