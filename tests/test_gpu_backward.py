"""GPU tier, training path (BASELINE.json configs[4]): every hand-written backward kernel, called
through the C ABI via scflow_b200.autograd, against autograd through the CPU oracle on the same
seeded inputs, and the whole decoder against gradients taken through the unmodified reference
(tests/golden/decoder_s64_grads.npz)."""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import scflow_oracle as O
from oracle.gen_golden import lookup_flow
from scflow_b200 import autograd as ag, ops, synthetic as syn
import scflow_b200 as sb
from scflow_b200.refiner import SCFlowRefinerCore, decoder_cfg

DEV = "cuda"


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(autouse=True, scope="module")
def _strict_fp32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("channels_last_grad", [False, True])
@pytest.mark.parametrize("n,h,w", [(2, 8, 16), (1, 32, 32), (3, 8, 8), (1, 64, 64), (3, 16, 32)])
def test_lookup_backward_vs_oracle(n, h, w, channels_last_grad):
    ft = syn.make_features(n, h, w, seed=11, c=16, hc=4)
    flow = lookup_flow(n, h, w, 12)
    pyr = [p.clone().requires_grad_(True) for p in O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)]
    out = O.corr_lookup(pyr, flow.clone())
    go = torch.from_numpy(np.random.RandomState(5).standard_normal(tuple(out.shape)).astype(np.float32))
    out.backward(go)
    pyr_d = [p.detach().to(DEV).requires_grad_(True) for p in pyr]
    got = ag.corr_lookup(pyr_d, flow.to(DEV), 4, ops.DIV_IEEE, channels_last=channels_last_grad)
    assert rel(got, out) < 1e-5
    gd = go.to(DEV)
    if channels_last_grad:
        gd = gd.contiguous(memory_format=torch.channels_last)
    got.backward(gd)
    for l, (a, b) in enumerate(zip(pyr_d, pyr)):
        assert a.grad.shape == b.grad.shape
        assert rel(a.grad, b.grad) < 1e-5, l
        # the scatter touches exactly the cells the reference's gather read
        assert torch.equal(a.grad.cpu() != 0, b.grad != 0) or rel(a.grad, b.grad) < 1e-6


@pytest.mark.parametrize("prec,tol", [(ops.PREC_FP32, 1e-5), (ops.PREC_TF32, 2e-3)])
@pytest.mark.parametrize("n,c,h,w", [(2, 64, 8, 32), (1, 256, 32, 32)])
def test_corr_pyramid_backward_vs_oracle(prec, tol, n, c, h, w):
    ft = syn.make_features(n, h, w, seed=3, c=c, hc=4)
    f1 = ft["feat_render"].clone().requires_grad_(True)
    f2 = ft["feat_real"].clone().requires_grad_(True)
    pyr = O.correlation_pyramid(f1, f2, 4)
    rs = np.random.RandomState(9)
    gs = [torch.from_numpy(rs.standard_normal(tuple(p.shape)).astype(np.float32)) for p in pyr]
    torch.autograd.backward(pyr, gs)
    d1 = ft["feat_render"].to(DEV).requires_grad_(True)
    d2 = ft["feat_real"].to(DEV).requires_grad_(True)
    got = ag.corr_pyramid(d1, d2, 4, prec)
    torch.autograd.backward(got, [g.to(DEV) for g in gs])
    assert rel(d1.grad, f1.grad) < tol
    assert rel(d2.grad, f2.grad) < tol


@pytest.mark.parametrize("prec,tol", [(ops.PREC_FP32, 2e-6), (ops.PREC_TF32, 2e-3)])
@pytest.mark.parametrize("n,c,h,w", [(3, 256, 32, 32), (2, 128, 16, 32), (2, 96, 8, 12)])
def test_corr_pyramid_bwd_gemms_vs_fp64(prec, tol, n, c, h, w):
    """scflow_corr_pyramid_bwd alone against the fp64 contraction; (3, 256, 32, 32) and (2, 128, 16, 32) are on the
    tcgen05 tiling in TF32 mode, (2, 96, 8, 12) is a ragged shape on the fp32 kernel in both modes."""
    g = torch.Generator().manual_seed(21)
    P = h * w
    f1 = torch.randn(n, c, h, w, generator=g)
    f2 = torch.randn(n, c, h, w, generator=g)
    G = torch.randn(n, P, P, generator=g)
    e1 = torch.einsum("nct,nqt->ncq", f2.double().reshape(n, c, P), G.double()).reshape(n, c, h, w)
    e2 = torch.einsum("ncq,nqt->nct", f1.double().reshape(n, c, P), G.double()).reshape(n, c, h, w)
    g1, g2 = ops.corr_pyramid_bwd(G.to(DEV), f1.to(DEV), f2.to(DEV), True, True, prec)
    assert rel(g1, e1.float()) < tol and rel(g2, e2.float()) < tol
    only1, none2 = ops.corr_pyramid_bwd(G.to(DEV), f1.to(DEV), f2.to(DEV), True, False, prec)
    assert none2 is None and torch.equal(only1, g1)
    none1, only2 = ops.corr_pyramid_bwd(G.to(DEV), f1.to(DEV), f2.to(DEV), False, True, prec)
    assert none1 is None and torch.equal(only2, g2)


def test_corr_pyramid_backward_partial_levels():
    """Levels that received no gradient (None) are skipped, as autograd would."""
    ft = syn.make_features(1, 8, 32, seed=4, c=32, hc=4)
    f1 = ft["feat_render"].clone().requires_grad_(True)
    pyr = O.correlation_pyramid(f1, ft["feat_real"], 4)
    (pyr[2] * 0.37).sum().backward()
    d1 = ft["feat_render"].to(DEV).requires_grad_(True)
    got = ag.corr_pyramid(d1, ft["feat_real"].to(DEV), 4, ops.PREC_FP32)
    (got[2] * 0.37).sum().backward()
    assert rel(d1.grad, f1.grad) < 1e-5


@pytest.mark.parametrize("n,c,h,w,scale", [(2, 2, 8, 8, 8), (1, 1, 32, 32, 8), (2, 2, 5, 7, 4)])
def test_upsample_backward_vs_torch(n, c, h, w, scale):
    g = torch.Generator().manual_seed(2)
    x = torch.randn(n, c, h, w, generator=g, requires_grad=True)
    y = torch.randn(n, c, h, w, generator=g, requires_grad=True)
    ref = 3.0 * F.interpolate(x + y, scale_factor=scale, mode="bilinear", align_corners=True)
    go = torch.randn(ref.shape, generator=g)
    ref.backward(go)
    xd = x.detach().to(DEV).requires_grad_(True)
    yd = y.detach().to(DEV).requires_grad_(True)
    out = ag.upsample(xd, yd, scale, 3.0)
    assert rel(out, ref) < 1e-6
    out.backward(go.to(DEV))
    assert rel(xd.grad, x.grad) < 1e-5 and rel(yd.grad, y.grad) < 1e-5


def test_pose_flow_backward_vs_oracle():
    cr = syn.make_crops(3, 64, seed=21)
    rs = np.random.RandomState(22)
    dR6 = torch.from_numpy((np.array([[1, 0, 0, 0, 1, 0]]) + 0.05 * rs.standard_normal((3, 6))).astype(np.float32))
    dt = torch.from_numpy((0.05 * rs.standard_normal((3, 3))).astype(np.float32))
    R, t = O.pose_compose(dR6, dt, cr["R0"], cr["t0"])
    R = R.detach().requires_grad_(True)
    t = t.detach().requires_grad_(True)
    pts, mask = O.lift_points_dense(cr["depth"], cr["K"], cr["R0"], cr["t0"])
    flow = O.pose_flow_dense(R, t, cr["K"], pts, mask, 0.0)
    go = torch.from_numpy(rs.standard_normal(tuple(flow.shape)).astype(np.float32))
    flow.backward(go)
    c = {k: v.to(DEV) for k, v in cr.items()}
    Rd = R.detach().to(DEV).requires_grad_(True)
    td = t.detach().to(DEV).requires_grad_(True)
    full, lr = ag.pose_flow(Rd, td, c["K"], c["depth"], c["R0"], c["t0"], 8, 0.0)
    assert rel(full, flow) < 1e-4 and not lr.requires_grad
    full.backward(go.to(DEV))
    assert rel(Rd.grad, R.grad) < 1e-4
    assert rel(td.grad, t.grad) < 1e-4


@pytest.mark.parametrize("detach", [False, True])
def test_pose_compose_kernels_all_gradients_vs_fp64(detach):
    """scflow_pose_compose_fwd / _bwd against autograd through the same algebra in fp64 (the torch form the kernels
    replaced), gradients wrt all four inputs, both detach_depth_for_xy settings, one-sided output gradients."""
    rs = np.random.RandomState(41)
    n = 37
    cr = syn.make_crops(4, 64, seed=42)
    R0 = cr["R0"].repeat(10, 1, 1)[:n]; t0 = cr["t0"].repeat(10, 1)[:n] * torch.from_numpy(1 + 0.1 * rs.rand(n, 1)).float()
    dR6 = torch.from_numpy((np.array([[1, 0, 0, 0, 1, 0]]) + 0.3 * rs.standard_normal((n, 6))).astype(np.float32))
    dt = torch.from_numpy((0.2 * rs.standard_normal((n, 3))).astype(np.float32))
    gR = torch.from_numpy(rs.standard_normal((n, 3, 3)).astype(np.float32))
    gt = torch.from_numpy(rs.standard_normal((n, 3)).astype(np.float32))
    ins64 = [v.double().requires_grad_(True) for v in (dR6, dt, R0, t0)]
    R, t = ag.compose_pose_torch(*ins64, detach_depth_for_xy=detach)
    torch.autograd.backward([R, t], [gR.double(), gt.double()])
    insd = [v.to(DEV).requires_grad_(True) for v in (dR6, dt, R0, t0)]
    Rd, td = ag.compose_pose(*insd, detach_depth_for_xy=detach)
    assert rel(Rd, R) < 1e-6 and rel(td, t) < 1e-6
    torch.autograd.backward([Rd, td], [gR.to(DEV), gt.to(DEV)])
    for a, b in zip(insd, ins64):
        assert rel(a.grad, b.grad) < 2e-6
    only = [v.to(DEV).requires_grad_(True) for v in (dR6, dt)]
    Rd, td = ag.compose_pose(only[0], only[1], R0.to(DEV), t0.to(DEV), detach_depth_for_xy=detach)
    (td * gt.to(DEV)).sum().backward()                                  # no gradient reaches R'
    ref = [v.double().requires_grad_(True) for v in (dR6, dt)]
    _, t2 = ag.compose_pose_torch(ref[0], ref[1], R0.double(), t0.double(), detach_depth_for_xy=detach)
    (t2 * gt.double()).sum().backward()
    assert float(only[0].grad.abs().max()) == 0 and rel(only[1].grad, ref[1].grad) < 2e-6


def test_convgru_training_path_vs_fp64_autograd():
    """ConvGRU under autograd (raft_decoder.py:235-253): cuDNN gate convolutions + the fused gate kernels
    (scflow_gru_reset / _blend fwd + bwd) against the plain torch chain in fp64: output and every gradient."""
    import copy
    g = torch.Generator().manual_seed(51)
    gru = sb.ConvGRU(128, 256, "SeqConv")
    ref = copy.deepcopy(gru).double()
    h = torch.randn(3, 128, 16, 24, generator=g) * 0.5
    x = torch.randn(3, 256, 16, 24, generator=g) * 0.5
    go = torch.randn(3, 128, 16, 24, generator=g)
    hr, xr = h.double().requires_grad_(True), x.double().requires_grad_(True)
    y = hr
    for cz, cr, cq in zip(ref.conv_z, ref.conv_r, ref.conv_q):
        hx = torch.cat([y, xr], 1)
        z, r = torch.sigmoid(cz.conv(hx)), torch.sigmoid(cr.conv(hx))
        q = torch.tanh(cq.conv(torch.cat([r * y, xr], 1)))
        y = (1 - z) * y + z * q
    y.backward(go.double())
    gru = gru.to(DEV)
    hd, xd = h.to(DEV).requires_grad_(True), x.to(DEV).requires_grad_(True)
    ops.STATS.reset()
    yd = gru(hd, xd)
    assert ops.STATS.count == 4                      # reset + blend, two passes
    yd.backward(go.to(DEV))
    assert rel(yd, y) < 2e-5
    assert rel(hd.grad, hr.grad) < 2e-5 and rel(xd.grad, xr.grad) < 2e-5
    for (k, p), (_, pr) in zip(gru.named_parameters(), ref.named_parameters()):
        assert rel(p.grad, pr.grad) < 5e-5, k


def test_compose_pose_matches_oracle_and_kernel():
    rs = np.random.RandomState(31)
    cr = syn.make_crops(4, 64, seed=32)
    dR6 = torch.from_numpy((np.array([[1, 0, 0, 0, 1, 0]]) + 0.1 * rs.standard_normal((4, 6))).astype(np.float32))
    dt = torch.from_numpy((0.1 * rs.standard_normal((4, 3))).astype(np.float32))
    a = dR6.clone().requires_grad_(True); b = dt.clone().requires_grad_(True)
    R, t = O.pose_compose(a, b, cr["R0"], cr["t0"], detach_depth_for_xy=True)
    ((R * R.detach().roll(1, 0)).sum() + (t * 0.01).sum()).backward()
    ad = dR6.to(DEV).requires_grad_(True); bd = dt.to(DEV).requires_grad_(True)
    Rd, td = ag.compose_pose(ad, bd, cr["R0"].to(DEV), cr["t0"].to(DEV), detach_depth_for_xy=True)
    ((Rd * Rd.detach().roll(1, 0)).sum() + (td * 0.01).sum()).backward()
    assert rel(Rd, R) < 1e-6 and rel(td, t) < 1e-6
    assert rel(ad.grad, a.grad) < 1e-5 and rel(bd.grad, b.grad) < 1e-5
    Rk, tk = sb.pose.get_pose_from_delta_pose(dR6.to(DEV), dt.to(DEV), cr["R0"].to(DEV), cr["t0"].to(DEV))
    assert rel(Rk, R) < 1e-6 and rel(tk, t) < 1e-6


def _decoder(s, seed, **over):
    hh = s // 8
    dec = sb.SCFlowDecoder(**decoder_cfg(feat_size=(hh, hh), **over)).train()
    shp = {k: tuple(v.shape) for k, v in dec.state_dict().items()}
    sd = syn.fill_state_dict(shp, seed=seed)
    dec.load_state_dict(sd, strict=True)
    return dec.to(DEV), sd


def _run_decoder_grads(dec, n, s, seed):
    cr = {k: v.to(DEV) for k, v in syn.make_crops(n, s, seed=seed + 1).items()}
    ft = syn.make_features(n, s // 8, s // 8, seed=seed + 2)
    ins = {k: ft[k].to(DEV).requires_grad_(True) for k in ("feat_render", "feat_real", "h_feat", "cxt_feat")}
    out = dec(ins["feat_render"], ins["feat_real"], ins["h_feat"], ins["cxt_feat"], cr["R0"], cr["t0"],
              cr["depth"], cr["K"], cr["labels"], torch.zeros(n, 2, s, s, device=DEV), 0.)
    loss = syn.synthetic_loss(out)
    loss.backward()
    return loss, ins, out


def test_decoder_gradients_s64_vs_reference_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "decoder_s64_grads.npz"))
    dec, _ = _decoder(64, 41, corr_precision="fp32")
    ops.STATS.reset()
    loss, ins, out = _run_decoder_grads(dec, 2, 64, 41)
    assert ops.STATS.count > 8 * 6, "the training path must launch the scflow kernels"
    assert abs(loss.item() - float(g["loss"])) < 1e-4 * abs(float(g["loss"]))
    for k, t in ins.items():
        assert rel(t.grad[:, ::4], torch.from_numpy(g["g_" + k])) < 2e-3, k
    params = dict(dec.named_parameters())
    for k in g.files:
        if k.startswith("p_"):
            assert rel(params[k[2:]].grad, torch.from_numpy(g[k])) < 2e-3, k
    norms = dict(zip([str(x) for x in g["param_names"]], g["param_grad_norms"]))
    for k, want in norms.items():
        got = 0.0 if params[k].grad is None else float(params[k].grad.double().norm())
        assert abs(got - want) <= 3e-3 * max(want, 1e-6), (k, got, want)


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("prec,tol", [("fp32", 2e-3), ("tf32", 5e-2)])
def test_decoder_gradients_cfg1_vs_oracle(prec, tol):
    """256x256 crop against fp32 oracle autograd.  fp32 correlation: tight.  With the tcgen05 TF32
    correlation in the forward the 8-iteration trajectory moves by ~1e-3 relative, and so do the
    (ReLU-gated) gradients: compared in the L2 norm at a TF32 tolerance."""
    dec, sd = _decoder(256, 51, corr_precision=prec)
    loss, ins, _ = _run_decoder_grads(dec, 1, 256, 51)
    cr = syn.make_crops(1, 256, seed=52)
    ft = syn.make_features(1, 32, 32, seed=53)
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    oi = {k: ft[k].clone().requires_grad_(True) for k in ("feat_render", "feat_real", "h_feat", "cxt_feat")}
    o = O.decoder_forward(sdo, oi["feat_render"], oi["feat_real"], oi["h_feat"], oi["cxt_feat"], cr["R0"],
                          cr["t0"], cr["depth"], cr["K"], cr["labels"], torch.zeros(1, 2, 256, 256),
                          train_detach=True)
    lo = syn.synthetic_loss(o)
    lo.backward()
    assert abs(loss.item() - lo.item()) < 2e-3 * abs(lo.item())
    for k in oi:
        e = rel_l2(ins[k].grad, oi[k].grad)
        print(f"{prec} d/d{k}: rel L2 {e:.2e}, max-norm {rel(ins[k].grad, oi[k].grad):.2e}")
        assert e < tol, k
    params = dict(dec.named_parameters())
    for k in ("gru.conv_q.1.conv.bias", "flow_pred.predict_layer.weight", "encoder.corr_net.0.conv.weight"):
        assert rel_l2(params[k].grad, sdo[k].grad) < max(tol, 5e-3), k      # fp32 wgrad sums cancel


def test_refiner_training_step_runs_and_updates():
    """Encoders + decoder under autograd, one SGD step: the loss is finite and the features'
    encoder receives a gradient through the correlation / lookup backward kernels."""
    core = SCFlowRefinerCore(256).train().to(DEV)
    es = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "encoder_state_shapes.json")))
    core.encoder.load_state_dict(syn.fill_state_dict(es["IN"], seed=61, gain=1.7))
    core.context.load_state_dict(syn.fill_state_dict(es["BN"], seed=62, gain=1.7))
    shp = {k: tuple(v.shape) for k, v in core.decoder.state_dict().items()}
    core.decoder.load_state_dict(syn.fill_state_dict(shp, seed=64))
    cr = {k: v.to(DEV) for k, v in syn.make_crops(2, 256, seed=63).items()}
    out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"], cr["K"],
                        cr["labels"])
    loss = syn.synthetic_loss(out)
    loss.backward()
    assert torch.isfinite(loss)
    gw = next(core.encoder.parameters()).grad
    assert gw is not None and torch.isfinite(gw).all() and float(gw.abs().max()) > 0
