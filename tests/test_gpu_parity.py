"""GPU tier: every CUDA kernel, called through the C ABI, against the CPU oracle on the same
seeded inputs, against the committed golden fixtures (outputs of the unmodified reference), and
-- at full BASELINE sizes -- through size-independent properties."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import scflow_oracle as O
from oracle.gen_golden import lookup_flow
from scflow_b200 import ops, synthetic as syn
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


# ------------------------------------------------------------------ a1 correlation pyramid
@pytest.mark.parametrize("prec,tol", [(ops.PREC_FP32, 2e-6), (ops.PREC_TF32, 1e-3)])
@pytest.mark.parametrize("n,c,h,w", [(2, 64, 8, 32), (1, 256, 32, 32), (1, 32, 16, 64)])
def test_corr_pyramid_vs_oracle(prec, tol, n, c, h, w):
    ft = syn.make_features(n, h, w, seed=3, c=c, hc=4)
    want = O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)
    got = ops.corr_pyramid(ft["feat_render"].to(DEV), ft["feat_real"].to(DEV), 4, prec)
    torch.cuda.synchronize()
    for l in range(4):
        assert got[l].shape == want[l].shape
        # tolerance relative to the level-0 scale: pooled levels average noise away
        err = float((got[l].cpu().double() - want[l].double()).abs().max() / want[0].abs().max())
        assert err < tol, (l, err)


def test_corr_pyramid_tf32_vs_tf32_rounded_oracle():
    """Tight check of the tcgen05 path: with inputs pre-rounded to TF32 the only difference left
    is fp32 accumulation order."""
    ft = syn.make_features(1, 32, 32, seed=5, c=256, hc=4)
    rnd = lambda t: (t.view(torch.int32) & ~0x1FFF).view(torch.float32)
    f1, f2 = rnd(ft["feat_render"].clone()), rnd(ft["feat_real"].clone())
    want = O.correlation_pyramid(f1, f2, 4)
    got = ops.corr_pyramid(f1.to(DEV), f2.to(DEV), 4, ops.PREC_TF32)
    for l in range(4):
        assert rel(got[l], want[l]) < 2e-5, l


def test_corr_pyramid_properties_full_size():
    """cfg2 size (N=64, 256 ch, 32x32): level l+1 is the exact 2x2 mean of level l, and the
    fp32 and tf32 kernels agree to TF32 precision."""
    g = torch.Generator(device=DEV).manual_seed(0)
    f1 = torch.randn(64, 256, 32, 32, device=DEV, generator=g)
    f2 = torch.randn(64, 256, 32, 32, device=DEV, generator=g)
    tf = ops.corr_pyramid(f1, f2, 4, ops.PREC_TF32)
    for l in range(3):
        pooled = torch.nn.functional.avg_pool2d(tf[l], 2, 2)
        assert float((pooled - tf[l + 1]).abs().max()) < 1e-5
    fp = ops.corr_pyramid(f1[:8], f2[:8], 4, ops.PREC_FP32)
    assert rel(tf[0][: 8 * 1024], fp[0]) < 1e-3
    ref = torch.matmul(f1[63].reshape(256, -1).t().double(), f2[63].reshape(256, -1).double()) / 16.0
    assert rel(tf[0][63 * 1024:].reshape(1024, 1024), ref) < 1e-3


# ------------------------------------------------------------------ a2 lookup
@pytest.mark.parametrize("channels_last", [True, False])
@pytest.mark.parametrize("div_mode", [ops.DIV_IEEE, ops.DIV_RCP])
@pytest.mark.parametrize("n,h,w", [(2, 8, 16), (1, 32, 32), (1, 16, 64), (3, 8, 8), (1, 64, 64)])
def test_lookup_vs_oracle_bit_exact_windows(div_mode, n, h, w, channels_last):
    ft = syn.make_features(n, h, w, seed=11, c=16, hc=4)
    pyr = O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)
    flow = lookup_flow(n, h, w, 12)
    want, wwin = O.corr_lookup(pyr, flow, 4, div_mode, return_windows=True)
    got, gwin = ops.corr_lookup([p.to(DEV) for p in pyr], flow.to(DEV), 4, div_mode, return_windows=True,
                                channels_last=channels_last)
    assert got.shape == want.shape
    assert got.is_contiguous(memory_format=torch.channels_last if channels_last else torch.contiguous_format)
    # window selection: integer corner indices, bit-exact (kernel clamps far-outside windows to
    # [-32, size+32], where every tap is zero padding anyway)
    for l in range(4):
        for ax, size in ((0, w >> l), (1, h >> l)):
            a = wwin[:, l, ax].clamp(-32, size + 32).to(torch.int16)
            assert torch.equal(gwin[:, l, ax].cpu(), a), (l, ax)
    assert rel(got, want) < 2e-6


@pytest.mark.parametrize("channels_last", [True, False])
@pytest.mark.parametrize("name", ["lookup_small", "lookup_sq32"])
def test_lookup_golden_fixture(golden_dir, name, channels_last):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    if name == "lookup_small":
        pyr = [torch.from_numpy(g[f"pyr{i}"]) for i in range(4)]
        sub = slice(None)
    else:      # inputs come from the seed; the fixture holds every other output pixel
        n, c, h, w = (int(v) for v in g["shape"])
        ft = syn.make_features(n, h, w, seed=11, c=c, hc=4)
        pyr = O.correlation_pyramid(ft["feat_render"], ft["feat_real"], 4)
        sub = slice(None, None, 2)
    got = ops.corr_lookup([p.to(DEV) for p in pyr], torch.from_numpy(g["flow"]).to(DEV), 4, ops.DIV_IEEE,
                          channels_last=channels_last)
    assert rel(got[:, :, sub, sub], torch.from_numpy(g["out"])) < 5e-6


def test_lookup_layouts_agree_full_size():
    """cfg2 size (N=64, 32x32): the warp-per-query NHWC kernel and the NCHW kernel replay the
    same op sequence, so values and window selection are identical; large flows push windows
    partly and wholly outside the levels (zero padding)."""
    g = torch.Generator(device=DEV).manual_seed(2)
    n, h, w = 64, 32, 32
    pyr = [torch.randn(n * h * w, 1, h >> l, w >> l, device=DEV, generator=g) for l in range(4)]
    flow = 9.0 * torch.randn(n, 2, h, w, device=DEV, generator=g)
    flow[0] = 0.0                      # integer coordinates: the round-trip edge case
    flow[1] = 500.0                    # everything out of range
    a, wa = ops.corr_lookup(pyr, flow, 4, return_windows=True, channels_last=True)
    b, wb = ops.corr_lookup(pyr, flow, 4, return_windows=True, channels_last=False)
    assert torch.equal(wa, wb)
    # same windows and weights; the NHWC kernel sums the four corners row-wise (separable), the
    # NCHW kernel in ATen's nw, ne, sw, se order: equal up to fp32 rounding order
    assert float((a - b).abs().max()) < 4e-6 * float(b.abs().max())
    assert float(a[1].abs().max()) == 0.0


def test_lookup_generations_identical():
    """The second-generation NHWC kernel (16x12 TMA box at every level, immediate-offset taps,
    w0 = 1 - w1) replays the first generation's arithmetic: identical windows AND identical
    values, on random, zero (round-trip edge case -> slow taps), far-out and mixed flows, for both
    division modes; 512^2-crop shape included.  (gen1=True: SCFLOW_LAYOUT_NHWC_GEN1.)"""
    g = torch.Generator(device=DEV).manual_seed(5)
    for n, h, w in ((8, 32, 32), (2, 64, 64), (2, 24, 32)):
        pyr = [torch.randn(n * h * w, 1, h >> l, w >> l, device=DEV, generator=g) for l in range(4)]
        flow = 6.0 * torch.randn(n, 2, h, w, device=DEV, generator=g)
        flow[0] = 0.0
        flow[1, :, : h // 2] = 300.0
        flow[1, :, h // 2:] = torch.round(flow[1, :, h // 2:])       # integer coordinates, non-zero
        for div in (ops.DIV_IEEE, ops.DIV_RCP):
            a, wa = ops.corr_lookup(pyr, flow, 4, div, return_windows=True, gen1=True)
            a2 = ops.corr_lookup(pyr, flow, 4, div, gen1=True)
            b, wb = ops.corr_lookup(pyr, flow, 4, div, return_windows=True)
            b2 = ops.corr_lookup(pyr, flow, 4, div)
            assert torch.equal(wa, wb), (n, h, w, div)
            assert torch.equal(a, b) and torch.equal(a2, b2) and torch.equal(a, a2), (n, h, w, div)


def test_lookup_split_planes_equal_fp32_result():
    """SCFLOW_LAYOUT_NHWC_SPLIT: the lookup written as split-bf16 planes (328 hi | 328 lo per query, channels in
    the kernel's lane-major order `ops.lookup_split_order`) carries the fp32 result to 16 significand bits, with
    zeros in the four padding channels; shapes the second-generation
    kernel does not take go through fp32 + one split pass and give the same planes."""
    g = torch.Generator(device=DEV).manual_seed(9)
    for n, h, w in ((4, 32, 32), (1, 64, 64), (2, 8, 16)):
        pyr = [torch.randn(n * h * w, 1, h >> l, w >> l, device=DEV, generator=g) for l in range(4)]
        flow = 5.0 * torch.randn(n, 2, h, w, device=DEV, generator=g)
        flow[0, :, : h // 2] = 0.0
        order = ops.lookup_split_order().to(DEV)
        assert sorted(order.tolist()) == list(range(324))
        want = ops.corr_lookup(pyr, flow, 4).permute(0, 2, 3, 1)[..., order]      # the planes' lane-major channel order
        sp = ops.corr_lookup(pyr, flow, 4, split=True)
        assert sp.c == 324 and sp.cp == 328 and tuple(sp.shape) == (n, h, w)
        got = sp.float()
        assert float((got - want).abs().max()) <= 2.0 ** -16 * float(want.abs().max())
        assert float(sp.buf[..., 324:328].float().abs().max()) == 0 and float(sp.buf[..., 652:656].float().abs().max()) == 0
        from scflow_b200 import convtc as C
        assert torch.equal(sp.buf[..., :324], C.to_split(want).buf[..., :324])          # same hi plane


def test_lookup_impulse_and_zero_flow():
    h = w = 32
    qy, qx = 5, 6
    lvl0 = torch.zeros(h * w, 1, h, w, device=DEV)
    lvl0[qy * w + qx, 0, qy + 1, qx + 3] = 1.0
    pyr = [lvl0] + [torch.zeros(h * w, 1, h >> i, w >> i, device=DEV) for i in (1, 2, 3)]
    out = ops.corr_lookup(pyr, torch.zeros(1, 2, h, w, device=DEV), 4)
    ch = out[0, :81, qy, qx]
    assert int(ch.argmax()) == 68 and abs(float(ch[68]) - 1.0) < 1e-6     # 9*a + b, a = x offset
    # zero flow: centre tap (a=b=4 -> channel 40) returns corr[q][q] up to the round-trip noise
    lvl = torch.randn(h * w, 1, h, w, device=DEV)
    pyr = [lvl] + pyr[1:]
    out = ops.corr_lookup(pyr, torch.zeros(1, 2, h, w, device=DEV), 4)
    diag = lvl.reshape(h * w, h * w).diagonal().reshape(h, w)
    assert float((out[0, 40] - diag).abs().max()) < 1e-4


# ------------------------------------------------------------------ a4-a7 pose / flow / resize
def _cuda(d):
    return {k: v.to(DEV) for k, v in d.items()}


def test_lift_and_pose_flow_vs_oracle_and_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "pose_small.npz"))
    cr = syn.make_crops(3, 32, seed=31)
    c = _cuda(cr)
    pts, mask = ops.lift_points(c["depth"], c["K"], c["R0"], c["t0"])
    wpts, wmask = O.lift_points_dense(cr["depth"], cr["K"], cr["R0"], cr["t0"])
    assert torch.equal(mask.cpu().bool(), wmask)
    assert rel(pts, wpts) < 1e-5 and rel(pts, torch.from_numpy(g["points3d"])) < 1e-5
    dR6, dt = torch.from_numpy(g["dR6"]), torch.from_numpy(g["dt"])
    wR, wt = O.pose_compose(dR6, dt, cr["R0"], cr["t0"])
    for inv, key in ((0.0, "flow"), (400.0, "flow400")):
        R, t, full, lr = ops.pose_flow(c["R0"], c["t0"], c["K"], dR6=dR6.to(DEV), dt=dt.to(DEV),
                                       depth=c["depth"], R0=c["R0"], t0=c["t0"], scale=8, invalid_num=inv)
        assert float((R.cpu() - torch.from_numpy(g["R"])).abs().max()) < 1e-6
        assert rel(t, torch.from_numpy(g["t"])) < 1e-6
        assert rel(R, wR) < 1e-6 and rel(t, wt) < 1e-6
        wfull = O.pose_flow_dense(wR, wt, cr["K"], wpts, wmask, inv)
        assert rel(full, wfull) < 1e-3                       # north_star: flow within 1e-3 relative
        assert rel(full, torch.from_numpy(g[key])) < 1e-3
        wlr = (1.0 / 8) * O.resize_bilinear_ac(wfull, 4, 4)
        assert rel(lr, wlr) < 1e-3
    # the stored-points variant (drop-in for get_flow_from_delta_pose_and_points) agrees
    ps, _ = sb.cal_3d_2d_corr(c["depth"], c["K"], c["R0"], c["t0"])
    f2 = sb.get_flow_from_delta_pose_and_points(R, t, c["K"], ps, ps, 32, 32, invalid_num=400.)
    assert rel(f2, full) < 1e-5
    R3, t3 = sb.get_pose_from_delta_pose(dR6.to(DEV), dt.to(DEV), c["R0"], c["t0"], detach_depth_for_xy=True)
    assert torch.equal(R3, R) and torch.equal(t3, t)


def test_pose_flow_accepts_the_reference_list_form():
    """get_flow_from_delta_pose_and_points with the reference's ragged arguments (pose.py:66-88: per image (Ni, 2)
    pixel coordinates and (Ni, 3) points, as the reference's own cal_3d_2d_corr builds them) equals the dense
    PointSet form and the oracle."""
    from scflow_b200 import pose as P
    cr = syn.make_crops(3, 64, seed=7)
    c = _cuda(cr)
    ps, _ = P.cal_3d_2d_corr(c["depth"], c["K"], c["R0"], c["t0"])
    R1 = c["R0"] @ torch.tensor([[0.995, -0.0998, 0.0], [0.0998, 0.995, 0.0], [0.0, 0.0, 1.0]], device=DEV)
    t1 = c["t0"] + torch.tensor([2.0, -1.0, 6.0], device=DEV)
    dense = P.get_flow_from_delta_pose_and_points(R1, t1, c["K"], ps, ps, 64, 64, invalid_num=400.)
    p2l, p3l = [], []
    for i in range(3):
        ys, xs = torch.nonzero(ps.mask[i], as_tuple=True)
        p2l.append(torch.stack([xs, ys], -1).float())
        p3l.append(ps.points3d[i][:, ys, xs].t().contiguous())
    ragged = P.get_flow_from_delta_pose_and_points(R1, t1, c["K"], p2l, p3l, 64, 64, invalid_num=400.)
    assert torch.equal(dense, ragged)
    pts, mask = O.lift_points_dense(cr["depth"], cr["K"], cr["R0"], cr["t0"])
    want = O.pose_flow_dense(R1.cpu(), t1.cpu(), cr["K"], pts, mask, 400.)
    assert float((ragged.cpu() - want).abs().max()) < 1e-3


def test_identity_delta_pose_keeps_pose_and_zero_flow():
    """KAT from the zero-initialised pose head (pose_head.py:187-198): dR6 = [1,0,0,0,1,0], dt = 0
    => R_k = R_0, t_k = t_0 and the pose-induced flow vanishes on the foreground."""
    cr = _cuda(syn.make_crops(4, 256, seed=5))
    dR6 = torch.tensor([1., 0, 0, 0, 1., 0], device=DEV).repeat(4, 1)
    dt = torch.zeros(4, 3, device=DEV)
    R, t, full, lr = ops.pose_flow(cr["R0"], cr["t0"], cr["K"], dR6=dR6, dt=dt, depth=cr["depth"],
                                   R0=cr["R0"], t0=cr["t0"])
    assert float((R - cr["R0"]).abs().max()) < 1e-6 and float((t - cr["t0"]).abs().max()) < 1e-4
    assert float(full.abs().max()) < 2e-2      # fp32 cancellation of ~1e3 mm coordinates
    assert float(lr.abs().max()) < 1e-2


def test_resize_vs_oracle():
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 2, 8, 8, generator=g)
    y = torch.randn(2, 2, 8, 8, generator=g)
    up = ops.upsample(x.to(DEV), y.to(DEV), 8, 8.0)
    assert rel(up, 8.0 * O.resize_bilinear_ac(x + y, 64, 64)) < 1e-6
    big = torch.randn(2, 2, 64, 64, generator=g)
    dn = ops.downsample(big.to(DEV), 8, 0.125)
    assert rel(dn, 0.125 * O.resize_bilinear_ac(big, 8, 8)) < 1e-6


# ------------------------------------------------------------------ a8 whole decoder / refiner
def _build_decoder(s, seed, golden_dir, **over):
    hh = s // 8
    dec = sb.SCFlowDecoder(**decoder_cfg(feat_size=(hh, hh), **over)).eval()
    shp = {k: tuple(v.shape) for k, v in dec.state_dict().items()}
    sd = syn.fill_state_dict(shp, seed=seed)
    dec.load_state_dict(sd, strict=True)
    return dec.to(DEV), sd


@pytest.mark.parametrize("tag,n,s,seed", [("s64", 2, 64, 41), ("cfg1", 1, 256, 51)])
@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_decoder_vs_oracle_and_golden(golden_dir, tag, n, s, seed, prec):
    # (S = 64: 8 x 8 features are outside the tcgen05 tiling; the TF32 request runs the fp32 FMA kernel)
    dec, sd = _build_decoder(s, seed, golden_dir, corr_precision=prec)
    cr = syn.make_crops(n, s, seed=seed + 1)
    ft = syn.make_features(n, s // 8, s // 8, seed=seed + 2)
    c, f = _cuda(cr), _cuda(ft)
    with torch.no_grad():
        out = dec(f["feat_render"], f["feat_real"], f["h_feat"], f["cxt_feat"], c["R0"], c["t0"],
                  c["depth"], c["K"], c["labels"], torch.zeros(n, 2, s, s, device=DEV), 0.)
    assert len(out) == 7 and all(len(o) == 8 for o in out)
    g = np.load(os.path.join(golden_dir, f"decoder_{tag}.npz"))
    Rg, tg = torch.from_numpy(g["rotations"]), torch.from_numpy(g["translations"])
    assert float(O.rotation_geodesic(Rg[-1], cr["R0"]).max()) > 1e-2       # the pose really moved
    R8, t8 = out[2][-1].cpu(), out[3][-1].cpu()
    # north_star: final 6D pose within 1e-4 rad / 1e-4 m (= 0.1 mm) after 8 iterations
    assert float(O.rotation_geodesic(R8, Rg[-1]).max()) < 1e-4
    assert float((t8 - tg[-1]).abs().max()) < 0.1
    names = ["flow_from_pose", "flow_from_pred", None, None, "masks"]
    for i in (0, 1, 4):
        st = torch.stack(out[i])[[0, -1]][..., ::4, ::4].cpu()
        assert rel(st, torch.from_numpy(g[names[i]])) < 1e-3, names[i]


def test_zero_init_pose_head_is_identity(golden_dir):
    """As shipped the pose head outputs the identity delta: every iteration returns (R0, t0)."""
    dec = sb.SCFlowDecoder(**decoder_cfg()).eval().to(DEV)
    cr = _cuda(syn.make_crops(2, 256, seed=9))
    ft = _cuda(syn.make_features(2, 32, 32, seed=10))
    with torch.no_grad():
        out = dec(ft["feat_render"], ft["feat_real"], ft["h_feat"], ft["cxt_feat"], cr["R0"], cr["t0"],
                  cr["depth"], cr["K"], cr["labels"], torch.zeros(2, 2, 256, 256, device=DEV), 0.)
    for R, t in zip(out[2], out[3]):
        assert float((R - cr["R0"]).abs().max()) < 1e-5
        assert float((t - cr["t0"]).abs().max()) < 1e-3


@pytest.mark.parametrize("tag,s,seed", [("cfg1", 256, 63), ("cfg4", 512, 65)])
def test_refiner_vs_golden(golden_dir, tag, s, seed):
    """Whole refiner (3 encoder passes + 8 iterations) against the unmodified reference's poses:
    cfg1 = one 256x256 crop, cfg4 = one 512x512 crop (64x64 features, 4096^2 correlation)."""
    g = np.load(os.path.join(golden_dir, f"refiner_{tag}.npz"))
    es = json.load(open(os.path.join(golden_dir, "encoder_state_shapes.json")))
    core = SCFlowRefinerCore(s).eval()
    core.encoder.load_state_dict(syn.fill_state_dict(es["IN"], seed=61, gain=1.7))
    core.context.load_state_dict(syn.fill_state_dict(es["BN"], seed=62, gain=1.7))
    shp = {k: tuple(v.shape) for k, v in core.decoder.state_dict().items()}
    core.decoder.load_state_dict(syn.fill_state_dict(shp, seed=64))
    core.to(DEV)
    cr = _cuda(syn.make_crops(1, s, seed=seed))
    with torch.no_grad():
        out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"],
                            cr["K"], cr["labels"])
    Rg, tg = torch.from_numpy(g["rotations"]), torch.from_numpy(g["translations"])
    er = float(O.rotation_geodesic(out[2][-1].cpu(), Rg[-1]).max())
    et = float((out[3][-1].cpu() - tg[-1]).abs().max())
    print(f"refiner {tag}: pose err {er:.3e} rad, {et:.3e} mm")
    assert er < 1e-4 and et < 0.1


def test_refiner_batch64_vs_oracle():
    """Whole-refiner parity at the bench's size (BASELINE.json configs[1]: 64 crops of 256 x 256, 8 iterations):
    every one of the 64 final poses against the fp32 CPU oracle, within the north_star budget."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from scflow_b200.refiner import unpack_pose
    torch.set_num_threads(os.cpu_count() or 1)
    host = bench.tile_batch(bench.make_batch(4, 1234, 256), 64)
    core = bench.seeded_weights(SCFlowRefinerCore(256)).eval().to(DEV)
    with torch.no_grad():
        Rg, tg = unpack_pose(core.refine({k: v.to(DEV) for k, v in host.items()}).cpu())
        se, sc, sd = bench.cpu_refiner_state(256)
        o = O.refiner_forward(se, sc, sd, host["rendered_images"], host["real_images"], host["R0"], host["t0"],
                              host["depth"], host["K"], host["labels"])
    er = float(O.rotation_geodesic(Rg, o["rotations"][-1]).max())
    et = float((tg - o["translations"][-1]).abs().max())
    moved = float(O.rotation_geodesic(o["rotations"][-1], host["R0"]).max())
    print(f"64 crops: pose err {er:.3e} rad, {et:.3e} mm (the poses moved by up to {moved:.3f} rad)")
    assert moved > 1e-2 and er < 1e-4 and et < 0.1


def test_forked_upsampling_changes_nothing_and_pose_gatherer_cuda():
    """The side-stream branch of the inference loop (decoder.FORK_UPSAMPLE) produces bit-identical outputs, eagerly
    and through the captured graph; PoseGatherer on CUDA (world 1: staging + side stream + host read) returns the
    poses it was given."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    import scflow_b200.decoder as dec_mod
    from scflow_b200.refiner import GraphedRefiner, PoseGatherer
    host = bench.tile_batch(bench.make_batch(4, 1234, 256), 8)
    dev_batch = {k: v.to(DEV) for k, v in host.items()}
    core = bench.seeded_weights(SCFlowRefinerCore(256)).eval().to(DEV)
    outs = {}
    old = dec_mod.FORK_UPSAMPLE
    try:
        for flag in (False, True):
            dec_mod.FORK_UPSAMPLE = flag
            with torch.no_grad():
                full = core.get_pose(dev_batch["rendered_images"], dev_batch["real_images"], dev_batch["R0"],
                                     dev_batch["t0"], dev_batch["depth"], dev_batch["K"], dev_batch["labels"])
                torch.cuda.synchronize()
                outs[flag] = [t.clone() for lst in full for t in lst]
        graphed = GraphedRefiner(core, dev_batch)                       # captured with the fork on
        poses_g = graphed().clone()
        with torch.no_grad():
            poses_e = core.refine(dev_batch)
    finally:
        dec_mod.FORK_UPSAMPLE = old
    assert len(outs[False]) == len(outs[True]) and all(torch.equal(a, b) for a, b in zip(outs[False], outs[True]))
    assert torch.equal(poses_g, poses_e)
    pg = PoseGatherer(8, DEV)
    pinned = torch.empty(pg.out[0].shape, pin_memory=True)
    for step in range(3):
        slot = pg.submit(poses_e + step)
        ev = pg.to_host(slot, pinned)
        assert torch.equal(pg.result(slot), poses_e + step)
        ev.synchronize()
        assert torch.equal(pinned[:8], (poses_e + step).cpu())


def test_batch64_matches_per_crop_runs():
    """Crops are independent: refining 64 crops at once equals refining them 8 at a time."""
    dec, _ = _build_decoder(256, 77, None)
    cr = _cuda(syn.make_crops(16, 256, seed=78))
    ft = _cuda(syn.make_features(16, 32, 32, seed=79))
    cr["labels"][:] = cr["labels"][0]     # pose_head.py:209 reads label[0] for the whole batch
    def run(sl):
        with torch.no_grad():
            o = dec(ft["feat_render"][sl], ft["feat_real"][sl], ft["h_feat"][sl], ft["cxt_feat"][sl],
                    cr["R0"][sl], cr["t0"][sl], cr["depth"][sl], cr["K"][sl], cr["labels"][sl],
                    torch.zeros(cr["R0"][sl].shape[0], 2, 256, 256, device=DEV), 0.)
        return o[2][-1], o[3][-1]
    Ra, ta = run(slice(0, 16))
    Rb = torch.cat([run(slice(i, i + 8))[0] for i in (0, 8)])
    assert float(O.rotation_geodesic(Ra.cpu(), Rb.cpu()).max()) < 1e-4


# ------------------------------------------------------------------ f3 encoder stem (space-to-depth form)
@pytest.mark.parametrize("n,c,h,w", [(2, 3, 256, 256), (1, 3, 64, 96), (3, 1, 16, 8)])
def test_stem_space_to_depth_input(n, c, h, w):
    """scflow_stem_s2d_split_fwd == pixel_unshuffle + channel / border padding as hi + lo bf16 planes (16
    significand bits), and the 4x4 stride-1 filter over it == the shipped 7x7 stride-2 convolution
    (raft_encoder.py:286-288), evaluated in fp64."""
    from scflow_b200 import encoder as E
    g = torch.Generator(device=DEV).manual_seed(31)
    x = torch.rand(n, c, h, w, device=DEV, generator=g)
    planes = ops.stem_s2d_split(x)
    got = planes[0].float() + planes[1].float()                       # (n, h/2+3, w/2+3, 16)
    want = torch.nn.functional.pad(torch.nn.functional.pixel_unshuffle(x, 2), (2, 1, 2, 1, 0, 16 - 4 * c))
    want = want.permute(0, 2, 3, 1)
    assert got.shape == want.shape
    assert float((got - want).abs().max()) <= 2.0 ** -17 * float(want.abs().max())
    assert bool(((want == 0) <= (got == 0)).all())                     # padding and unused channels are exact zeros
    wgt = torch.randn(64, c, 7, 7, device=DEV, generator=g)
    ref = torch.nn.functional.conv2d(x.double(), wgt.double(), None, 2, 3)
    out = torch.nn.functional.conv2d(want.permute(0, 3, 1, 2).double(), E._s2d_filter(wgt).double())
    assert rel(out, ref) < 1e-12


# ------------------------------------------------------------------ a3 ConvGRU (tcgen05)
def _gru_sd(seed):
    shapes = {}
    for nm in "zrq":
        for p, k in ((0, (1, 5)), (1, (5, 1))):
            shapes[f"conv_{nm}.{p}.conv.weight"] = (128, 384) + k
            shapes[f"conv_{nm}.{p}.conv.bias"] = (128,)
    return syn.fill_state_dict(shapes, seed=seed, gain=2.0)


@pytest.mark.parametrize("n,h,w", [(2, 32, 32), (1, 64, 64), (3, 8, 32), (5, 8, 8), (64, 32, 32)])
def test_convgru_tcgen05_vs_oracle(n, h, w):
    """ConvGRU on the split-bf16 convolution family (3 launches per pass, gate epilogues) against the oracle's
    restatement of raft_decoder.py:245-252 in fp32: 5e-5 of the state's magnitude (the TF32 kernels of round 1
    needed 3e-3 here)."""
    from scflow_b200 import gru_tc
    sd = _gru_sd(21)
    mod = sb.ConvGRU(128, 256)
    mod.load_state_dict(sd)
    mod.to(DEV)
    rs = np.random.RandomState(22)
    hh = torch.from_numpy(np.tanh(rs.standard_normal((n, 128, h, w))).astype(np.float32))
    xx = torch.from_numpy(rs.standard_normal((n, 256, h, w)).astype(np.float32))
    want = O.conv_gru(hh[:4], xx[:4], sd)
    hd = hh.to(DEV)
    got = gru_tc.forward(mod, hd, xx.to(DEV))
    assert torch.equal(hd.cpu(), hh)              # the caller's state is not modified
    assert got.shape == hh.shape
    err = rel(got[:4], want)
    assert err < 5e-5, err
    # the module entry point (reference signature) is the same path
    with torch.no_grad():
        again = mod(hd, xx.to(DEV))
    assert torch.equal(again, got)


def test_convgru_context_precompute_equals_full_convolution():
    """gru_tc.precompute: the context channels' share of the gate pre-activations (the same in every iteration) is
    convolved once and added in the gate epilogues; the recurrent convolutions then cover [motion | h] only.  Same
    result as the convolution over all 384 channels (fp32 accumulation split in two: 2e-5), over two successive updates, and
    against the oracle."""
    from scflow_b200 import gru_tc, convtc as C
    sd = _gru_sd(23)
    mod = sb.ConvGRU(128, 256)
    mod.load_state_dict(sd)
    mod.to(DEV)
    rs = np.random.RandomState(24)
    n, h, w = 3, 32, 32
    hh = torch.from_numpy(np.tanh(rs.standard_normal((n, 128, h, w))).astype(np.float32))
    xx = torch.from_numpy(rs.standard_normal((n, 256, h, w)).astype(np.float32))
    outs = []
    for use_pre in (False, True):
        h_state = hh.to(DEV).permute(0, 2, 3, 1).contiguous().clone()
        xh = C.alloc_split(n, h, w, 384, DEV)
        C.to_split(xx.to(DEV).permute(0, 2, 3, 1), out=xh.slice(0, 256))
        C.to_split(h_state, out=xh.slice(256, 128))
        pre = gru_tc.precompute(mod, xh.slice(0, 128)) if use_pre else None
        for p_ in (0, 1):                                   # filter packs are cached: build them before counting
            gru_tc.packs_split(mod, p_, 128) if use_pre else gru_tc.packs(mod, p_)
        ops.STATS.reset()
        for _ in range(2):
            gru_tc.update_inplace(mod, xh, h_state, pre=pre)
        assert ops.STATS.count == 8
        outs.append((h_state.clone(), xh.slice(256, 128).float().clone()))
    assert rel(outs[1][0], outs[0][0]) < 2e-5 and rel(outs[1][1], outs[0][1]) < 2e-5
    want = O.conv_gru(O.conv_gru(hh, xx, sd), xx, sd)
    assert rel(outs[1][0].permute(0, 3, 1, 2), want) < 5e-5


def test_convgru_weight_cache_tracks_parameter_updates():
    from scflow_b200 import gru_tc
    mod = sb.ConvGRU(128, 256).to(DEV)
    a = gru_tc.packs(mod, 0)
    assert gru_tc.packs(mod, 0) is a
    with torch.no_grad():
        mod.conv_z[0].conv.weight.mul_(2.0)
    b = gru_tc.packs(mod, 0)
    assert b is not a and not torch.equal(a[0].w, b[0].w)


def test_decoder_with_tcgen05_gru(golden_dir):
    dec, sd = _build_decoder(256, 51, golden_dir, corr_precision="tf32")
    cr = syn.make_crops(1, 256, seed=52)
    ft = syn.make_features(1, 32, 32, seed=53)
    c, f = _cuda(cr), _cuda(ft)
    with torch.no_grad():
        out = dec(f["feat_render"], f["feat_real"], f["h_feat"], f["cxt_feat"], c["R0"], c["t0"],
                  c["depth"], c["K"], c["labels"], torch.zeros(1, 2, 256, 256, device=DEV), 0.)
    gd = np.load(os.path.join(golden_dir, "decoder_cfg1.npz"))
    Rg, tg = torch.from_numpy(gd["rotations"]), torch.from_numpy(gd["translations"])
    er = float(O.rotation_geodesic(out[2][-1].cpu(), Rg[-1]).max())
    et = float((out[3][-1].cpu() - tg[-1]).abs().max())
    print(f"split-bf16 GRU + TF32 corr: pose err {er:.3e} rad, {et:.3e} mm")
    assert er < 1e-4 and et < 0.1


# ------------------------------------------------------------------ f3 encoders: tensor-core path vs nn.Module path
@pytest.mark.parametrize("norm", ["IN", "BN"])
def test_encoder_fused_path_matches_module_path(norm):
    """The tensor-core inference path of RAFTEncoder (split-bf16 convolutions) equals the plain nn.Module
    path (cuDNN fp32; the strict-fp32 fixture disables TF32)."""
    from scflow_b200.encoder import RAFTEncoder
    torch.manual_seed(3)
    enc = RAFTEncoder(3, 256, net_type="Basic", norm_cfg=dict(type=norm)).to(DEV).eval()
    if norm == "BN":
        for m in enc.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.normal_(0, 0.3); m.running_var.uniform_(0.5, 2.0)
                m.weight.data.uniform_(0.5, 1.5); m.bias.data.normal_(0, 0.2)
    x = torch.rand(2, 3, 64, 64, device=DEV)
    with torch.no_grad():
        got = enc(x)
        with torch.enable_grad():            # grad mode on -> the module path
            want = enc(x.clone().requires_grad_(False)).detach()
    assert got.shape == want.shape
    assert rel(got, want) < 1e-4          # 13 normalised layers deep; cuDNN TF32 reads ~3e-3 here


# ------------------------------------------------------------------ f2: degenerate convolutions
@pytest.mark.parametrize("cin,k,cout,n,h,w", [(2, 7, 128, 3, 32, 32), (1, 3, 64, 2, 32, 32), (2, 7, 128, 2, 8, 8),
                                              (2, 7, 128, 1, 64, 64), (1, 3, 64, 2, 16, 40)])
@pytest.mark.parametrize("channels_last", [False, True])
def test_conv_small_cin_vs_torch(cin, k, cout, n, h, w, channels_last):
    from scflow_b200.modules import ConvModule
    g = torch.Generator().manual_seed(5)
    m = ConvModule(cin, cout, k, padding=k // 2).to(DEV)
    x = torch.randn(n, cin, h, w, generator=g).to(DEV)
    if channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
    want = torch.relu(torch.nn.functional.conv2d(x.double(), m.conv.weight.double(), m.conv.bias.double(), padding=k // 2))
    ops.STATS.reset()
    with torch.no_grad():
        got = m(x)
    assert ops.STATS.count == 1 and got.shape == want.shape
    assert rel(got, want) < 1e-5
    # the same launch writing split-bf16 planes (the operand of the tensor-core convolution that follows):
    # hi + lo carries 16 significand bits of the fp32 result
    with torch.no_grad():
        sp = m.forward_small_cin_split(x)
    assert rel(sp.float().permute(0, 3, 1, 2), want) < 3e-5        # tensor-core form (h % 16 == 0) or the FMA kernel


@pytest.mark.parametrize("k,cout,n,h,w", [(3, 2, 3, 32, 32), (1, 1, 2, 32, 32), (3, 2, 2, 8, 8), (3, 2, 1, 64, 64),
                                          (1, 1, 1, 64, 64), (3, 2, 1, 10, 16)])
def test_conv_small_cout_vs_torch(k, cout, n, h, w):
    from scflow_b200.modules import XHead
    g = torch.Generator().manual_seed(6)
    head = XHead(128, [256], cout, x="flow" if k == 3 else "mask").to(DEV)
    y = torch.randn(n, 256, h, w, generator=g).to(DEV).contiguous(memory_format=torch.channels_last)
    p = head.predict_layer
    want = torch.nn.functional.conv2d(y.double(), p.weight.double(), p.bias.double(), padding=k // 2)
    with torch.no_grad():
        got = ops.conv_small_cout(y, p)
    assert got.is_contiguous() and got.shape == want.shape
    assert rel(got, want) < 1e-5
    if (h * w) % 128 == 0 or 128 % (h * w) == 0:
        with torch.no_grad():                   # the module path: split pass + tensor-core 3x3 + the predictor
            xin = torch.randn(n, 128, h, w, generator=g).to(DEV).contiguous(memory_format=torch.channels_last)
            head(xin)
            ops.STATS.reset()
            got = head(xin)
            assert ops.STATS.count == 3
        want = head.predict_layer(torch.relu(torch.nn.functional.conv2d(
            xin.double(), head.layers[0].conv.weight.double(), head.layers[0].conv.bias.double(), padding=1)).float())
        assert rel(got, want) < 2e-5


def test_group_norm_relu_vs_torch():
    from scflow_b200.modules import ConvModule
    g = torch.Generator().manual_seed(8)
    for (n, cin, h, w) in ((3, 224, 32, 32), (2, 128, 8, 8), (5, 128, 16, 16)):
        m = ConvModule(cin, 128, 3, stride=2, padding=1, norm_cfg=dict(type="GN", num_groups=32), act_cfg=dict(type="ReLU")).to(DEV)
        with torch.no_grad():
            m.gn.weight.copy_(torch.randn(128, generator=g)); m.gn.bias.copy_(torch.randn(128, generator=g))
        x = torch.randn(n, cin, h, w, generator=g).to(DEV)
        y = torch.nn.functional.conv2d(x, m.conv.weight, None, stride=2, padding=1)
        want = torch.relu(torch.nn.functional.group_norm(y.double(), 32, m.gn.weight.double(), m.gn.bias.double(), m.gn.eps))
        with torch.no_grad():
            m(x.contiguous(memory_format=torch.channels_last))            # packs the filter
            ops.STATS.reset()
            got = m(x.contiguous(memory_format=torch.channels_last))
        assert ops.STATS.count == 3              # split pass, tensor-core convolution, GN+ReLU
        assert rel(got, want) < 2e-5
        with torch.no_grad():                    # GN+ReLU written as split planes only (between two convolutions)
            from scflow_b200 import convtc as C
            sp = m.forward_split(C.to_split(x.permute(0, 2, 3, 1)), want="split")
        assert rel(sp.float().permute(0, 3, 1, 2), want) < 3e-5


@pytest.mark.parametrize("n,feat", [(64, 32), (3, 32), (2, 64), (5, 8)])
def test_pose_head_fused_vs_module_path(n, feat):
    from scflow_b200.pose_head import MultiClassPoseHead
    g = torch.Generator().manual_seed(9)
    head = MultiClassPoseHead(21, 224, "Basic", dict(type="GN", num_groups=32), dict(type="ReLU"),
                              feat_size=(feat, feat), rotation_mode="ortho6d").to(DEV)
    with torch.no_grad():
        for p_ in head.parameters():
            p_.copy_(torch.randn(p_.shape, generator=g).to(DEV) * (0.05 if p_.dim() > 1 else 0.5))
    x = torch.randn(n, 224, feat, feat, generator=g).to(DEV)
    label = torch.randint(0, 21, (n,), generator=g).to(DEV)
    with torch.enable_grad():                                  # module path (cuDNN / cuBLAS fp32)
        dr_w, dt_w = head(x, label)
    with torch.no_grad():
        head(x.contiguous(memory_format=torch.channels_last), label)       # packs the filters
        ops.STATS.reset()
        dr, dt = head(x.contiguous(memory_format=torch.channels_last), label)
    assert ops.STATS.count == 10               # split pass + 3 x (tensor-core conv, GN+ReLU -> split planes) + 3 FC launches
    assert dr.shape == dr_w.shape and dt.shape == dt_w.shape
    assert rel(dr, dr_w) < 1e-4 and rel(dt, dt_w) < 1e-4


def test_pose_head_out_of_range_class_is_loud():
    """pose_head.py:209-210: index_select raises on a class outside [0, num_class).  The kernel path cannot raise
    without a host synchronisation; it returns NaN poses instead of silently clamping the class."""
    from scflow_b200.pose_head import MultiClassPoseHead
    head = MultiClassPoseHead(21, 224, "Basic", dict(type="GN", num_groups=32), dict(type="ReLU"),
                              feat_size=(32, 32), rotation_mode="ortho6d").to(DEV).eval()
    x = torch.randn(2, 224, 32, 32, device=DEV).contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        dr, dt = head(x, torch.tensor([3, 3], device=DEV))
        assert torch.isfinite(dr).all() and torch.isfinite(dt).all()
        for bad in (21, -1):
            dr, dt = head(x, torch.tensor([bad, 3], device=DEV))
            assert torch.isnan(dr).all() and torch.isnan(dt).all()


@pytest.mark.parametrize("tag,s,seed", [("cfg1", 256, 63)])
def test_refiner_has_one_precision_mode(golden_dir, tag, s, seed):
    """No cuDNN / cuBLAS call is left on the inference path: torch's TF32 switches (which moved round 1's
    result by 3-6e-4 rad) change nothing, and the one mode there is meets the 1e-4 rad / 0.1 mm budget
    against the unmodified reference's poses."""
    g = np.load(os.path.join(golden_dir, f"refiner_{tag}.npz"))
    es = json.load(open(os.path.join(golden_dir, "encoder_state_shapes.json")))
    core = SCFlowRefinerCore(s).eval()
    core.encoder.load_state_dict(syn.fill_state_dict(es["IN"], seed=61, gain=1.7))
    core.context.load_state_dict(syn.fill_state_dict(es["BN"], seed=62, gain=1.7))
    shp = {k: tuple(v.shape) for k, v in core.decoder.state_dict().items()}
    core.decoder.load_state_dict(syn.fill_state_dict(shp, seed=64))
    core.to(DEV)
    cr = _cuda(syn.make_crops(1, s, seed=seed))
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    outs = []
    try:
        for flag in (True, False):
            torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = flag
            with torch.no_grad():
                out = core.get_pose(cr["rendered_images"], cr["real_images"], cr["R0"], cr["t0"], cr["depth"],
                                    cr["K"], cr["labels"])
            outs.append((out[2][-1].cpu(), out[3][-1].cpu()))
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    # (InstanceNorm statistics are fp64 atomics: the summation order, hence the last bits, may differ)
    assert float(O.rotation_geodesic(outs[0][0], outs[1][0]).max()) < 2e-6
    assert float((outs[0][1] - outs[1][1]).abs().max()) < 2e-3
    Rg, tg = torch.from_numpy(g["rotations"]), torch.from_numpy(g["translations"])
    assert float(O.rotation_geodesic(outs[0][0], Rg[-1]).max()) < 1e-4
    assert float((outs[0][1] - tg[-1]).abs().max()) < 0.1
