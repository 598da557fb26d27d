"""GPU tier: the split-bf16 tensor-core convolution family (csrc/conv_tc.cu) through the C ABI against
ATen's convolution evaluated in fp64 on the CPU (what the reference's nn.Conv2d computes, models/
encoder/raft_encoder.py:286-314, models/decoder/raft_decoder.py:152-166,256-294, models/head/pose_head.py:
108-150).  Tolerance: 2e-5 of the output's max magnitude (three bf16 MMAs per product: ~2^-17 relative
per product, fp32 accumulation); TF32 sits at ~5e-4 on the same inputs.  The tensor core adds into its
fp32 accumulator with truncation, a drift of <= 2^-24 per MMA that grows with the number of accumulation
steps (3 K / 16): the bound is widened by that amount for the deep filters."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from scflow_b200 import convtc as C

DEV = "cuda"
TOL = 2e-5


def _ref(x, w, b, stride, pad, relu=False, res=None):
    y = F.conv2d(x.double().permute(0, 3, 1, 2), w.double(), None if b is None else b.double(), stride, pad)
    if res is not None:
        y = y + res.double().permute(0, 3, 1, 2)
    if relu:
        y = torch.relu(y)
    return y.permute(0, 2, 3, 1)


def _err(got, want):
    want = want.double()
    return float((got.double().cpu() - want).abs().max() / want.abs().max().clamp_min(1e-30))


CASES = [
    # n, h, w, cin, cout, k, stride, pad, bk
    (3, 32, 32, 64, 64, 3, 1, 1, 64),
    (3, 32, 32, 64, 64, 3, 1, 1, 32),
    (2, 32, 32, 64, 96, 3, 2, 1, 64),        # stride 2 through elementStrides
    (2, 32, 32, 64, 96, 1, 2, 0, 64),        # 1x1 stride-2 downsample
    (2, 16, 16, 96, 128, 3, 1, 1, 64),       # cin not a multiple of the k-block
    (2, 32, 32, 128, 256, 1, 1, 0, 64),
    (2, 32, 32, 128, 256, 1, 1, 0, 32),
    (2, 32, 32, 128, 512, 3, 1, 1, 32),      # two output-channel tiles
    (3, 32, 32, 224, 128, 3, 2, 1, 64),      # pose head: 32 -> 16
    (3, 16, 16, 128, 128, 3, 2, 1, 64),      # 16 -> 8: two images per tile, n not a multiple
    (11, 8, 8, 128, 128, 3, 2, 1, 64),       # 8 -> 4: eight images per tile
    (1, 8, 128, 16, 64, 3, 1, 1, 64),        # full-width tile rows
    (1, 4, 256, 32, 32, 3, 1, 1, 64),        # two tiles per row
    (2, 32, 32, 336, 256, 1, 1, 0, 32),      # lookup width (324 padded to 336)
    (2, 32, 32, 256, 126, 3, 1, 1, 64),      # 126 output channels
    (1, 64, 64, 128, 64, 3, 1, 1, 64),
    (2, 32, 32, 384, 128, 5, 1, 2, 64),      # wide taps
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride,pad,bk", CASES)
def test_conv_tc_vs_fp64(n, h, w, cin, cout, k, stride, pad, bk):
    g = torch.Generator().manual_seed(n * 1000 + cin + cout + k)
    x = torch.randn(n, h, w, cin, generator=g)
    wt = torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5
    b = torch.randn(cout, generator=g)
    pk = C.pack_weights(wt.to(DEV), b.to(DEV), stride, pad, bk=bk)
    xs = C.to_split(x.to(DEV))
    assert _err(xs.float(), x) < 2e-5
    got = C.conv(xs, pk, relu=True)
    torch.cuda.synchronize()
    want = _ref(x, wt, b, stride, pad, relu=True)
    assert tuple(got.shape) == tuple(want.shape[:3])
    e = _err(got.float(), want)
    assert e < TOL + 3 * (cin * k * k / 16) * 2.0 ** -24, e


PATCH_CASES = [
    # n, h, w, cin, cout, kh, kw, pad_h, pad_w      (stride 1, bk 64: the patch-resident kernel's domain)
    (2, 32, 32, 64, 64, 1, 1, 0, 0),             # patch = the tile itself
    (2, 32, 32, 64, 64, 3, 3, 1, 1),             # resident filter
    (3, 16, 8, 96, 96, 3, 3, 1, 1),              # one tile per image: an odd tile count (PAIR: one dead tile)
    (2, 32, 32, 384, 256, 1, 5, 0, 2),           # ConvGRU (1,5): streamed filter, 6 k-blocks
    (2, 32, 32, 128, 256, 5, 1, 2, 0),           # (5,1)
    (1, 64, 64, 128, 64, 3, 3, 1, 1),
    (2, 32, 32, 324, 256, 1, 1, 0, 0),           # lookup width: last k-block has one k-step
    (2, 32, 32, 256, 126, 3, 3, 1, 1),           # 126 output channels (tile of 128)
    (1, 128, 128, 64, 64, 3, 3, 1, 1),           # 128 tiles per image
    (1, 16, 8, 128, 512, 3, 3, 1, 1),            # one tile per image, two output-channel tiles
    (2, 32, 32, 128, 128, 7, 7, 3, 3),           # wide patch (22 x 14 rows), 49 taps > 32: per-tap kernel unless forced
]


@pytest.mark.parametrize("algo", [C.ALGO_TAP_BOXES, C.ALGO_PATCH, C.ALGO_PATCH_PAIR])
@pytest.mark.parametrize("n,h,w,cin,cout,kh,kw,ph,pw", PATCH_CASES)
def test_conv_tc_algorithms_vs_fp64(n, h, w, cin, cout, kh, kw, ph, pw, algo):
    """The three operand-movement variants of scflow_conv_tc_fwd (per-tap boxes, patch-resident, patch-resident
    on CTA pairs) against fp64, and against each other: same products, same accumulation order per output."""
    from scflow_b200._lib import ScflowError
    g = torch.Generator().manual_seed(n * 1000 + cin + cout + kh)
    x = torch.randn(n, h, w, cin, generator=g)
    wt = torch.randn(cout, cin, kh, kw, generator=g) / (cin * kh * kw) ** 0.5
    b = torch.randn(cout, generator=g)
    pk = C.pack_weights(wt.to(DEV), b.to(DEV), 1, (ph, pw))
    assert pk.bk == 64
    xs = C.to_split(x.to(DEV))
    if kh * kw > 32 and algo != C.ALGO_TAP_BOXES:
        with pytest.raises(ScflowError):
            C.conv(xs, pk, relu=True, algo=algo)
        return
    got = C.conv(xs, pk, relu=True, algo=algo)
    torch.cuda.synchronize()
    want = F.conv2d(x.double().permute(0, 3, 1, 2), wt.double(), b.double(), 1, (ph, pw)).relu().permute(0, 2, 3, 1)
    e = _err(got.float(), want)
    assert e < TOL + 3 * (cin * kh * kw / 16) * 2.0 ** -24, e


@pytest.mark.parametrize("algo", [C.ALGO_TAP_BOXES, C.ALGO_PATCH, C.ALGO_PATCH_PAIR])
def test_conv_tc_two_input_tensors_equal_concatenation(algo):
    """x2: the filter's input channels continue in a second tensor (raft_decoder.py:250, cat([r*h, x]) read from
    where the two lie): identical to the convolution of the concatenated operand."""
    g = torch.Generator().manual_seed(21)
    n, h, w, c1, c2, cout = 2, 32, 32, 128, 256, 128
    a = torch.randn(n, h, w, c1, generator=g)
    b = torch.randn(n, h, w, c2, generator=g)
    wt = torch.randn(cout, c1 + c2, 1, 5, generator=g) / (5 * (c1 + c2)) ** 0.5
    bias = torch.randn(cout, generator=g)
    pk = C.pack_weights(wt.to(DEV), bias.to(DEV), 1, (0, 2))
    wide = C.alloc_split(n, h, w, 512, DEV, zero=True)              # b lives in a channel slice of a wider buffer
    bs = C.to_split(b.to(DEV), out=wide.slice(64, c2))
    got = C.conv(C.to_split(a.to(DEV)), pk, x2=bs, relu=True, algo=algo)
    one = C.conv(C.to_split(torch.cat([a, b], -1).to(DEV)), pk, relu=True, algo=algo)
    torch.cuda.synchronize()
    assert torch.equal(got.buf, one.buf)
    want = _ref(torch.cat([a, b], -1), wt, bias, 1, (0, 2), relu=True)
    assert _err(got.float(), want) < TOL + 3 * (5 * (c1 + c2) / 16) * 2.0 ** -24


def test_conv_tc_outputs_residuals_stats():
    g = torch.Generator().manual_seed(7)
    n, h, w, cin, cout = 3, 16, 16, 64, 96
    x = torch.randn(n, h, w, cin, generator=g)
    wt = torch.randn(cout, cin, 3, 3, generator=g) / 24
    b = torch.randn(cout, generator=g)
    r = torch.randn(n, h, w, cout, generator=g)
    pk = C.pack_weights(wt.to(DEV), b.to(DEV), 1, 1)
    xs = C.to_split(x.to(DEV))
    want = _ref(x, wt, b, 1, 1, relu=True, res=r)
    # split residual -> all three outputs at once, into channel slices of wider buffers
    rs = C.to_split(r.to(DEV))
    wide = C.alloc_split(n, h, w, 224, DEV, zero=True)
    wide_f = torch.zeros(n, h, w, 160, device=DEV)
    planar = torch.empty(n, cout, h, w, device=DEV)
    C.conv(xs, pk, relu=True, res=rs, out=wide.slice(64, cout), out_f32=wide_f[..., 32:32 + cout],
           out_planar=planar)
    torch.cuda.synchronize()
    assert _err(wide.slice(64, cout).float(), want) < TOL
    assert float(wide.slice(0, 64).float().abs().max()) == 0 and float(wide.slice(160, 64).float().abs().max()) == 0
    assert _err(wide_f[..., 32:32 + cout], want) < TOL
    assert float(wide_f[..., :32].abs().max()) == 0 and float(wide_f[..., 32 + cout:].abs().max()) == 0
    assert _err(planar.permute(0, 2, 3, 1), want) < TOL
    # fp32 residual + InstanceNorm statistics of the pre-residual value
    stats = torch.zeros(n, C.STATS_REPLICAS, cout, 2, device=DEV, dtype=torch.float64)
    got = C.conv(xs, pk, relu=True, res_f32=r.to(DEV), stats=stats, want="f32")
    torch.cuda.synchronize()
    assert _err(got[..., :cout], want) < TOL
    pre = _ref(x, wt, b, 1, 1)
    assert int((stats.abs().sum((0, 2, 3)) > 0).sum()) > 1          # the atomics are spread over the replicas
    stats = stats.sum(1)
    assert _err(stats[:, :, 0], pre.sum((1, 2))) < 1e-5
    assert _err(stats[:, :, 1], (pre * pre).sum((1, 2))) < 1e-5
    # planar output rounded to TF32: low 13 mantissa bits are zero
    pl = C.conv(xs, pk, want="planar", planar_tf32=True)
    torch.cuda.synchronize()
    assert int((pl.view(torch.int32) & 0x1FFF).abs().max()) == 0
    assert _err(pl.permute(0, 2, 3, 1), pre) < 5e-4


def test_conv_tc_rejects():
    from scflow_b200._lib import ScflowError
    x = C.alloc_split(1, 24, 24, 64, DEV, zero=True)            # 24 does not tile 128 pixels
    pk = C.pack_weights(torch.zeros(64, 64, 3, 3, device=DEV), None, 1, 1)
    with pytest.raises(ScflowError):
        C.conv(x, pk)
    with pytest.raises(ScflowError):
        C.to_split(torch.zeros(1, 8, 8, 8))                    # CPU tensor


# ------------------------------------------------------------------ f3: the RAFT encoders, layer by layer
@pytest.mark.parametrize("norm", ["IN", "BN"])
@pytest.mark.parametrize("n,s", [(2, 64), (1, 256)])
def test_encoder_layers_vs_oracle(golden_dir, norm, n, s):
    """RAFTEncoder on the tensor-core convolution family against the oracle's restatement of
    raft_encoder.py:286-314 / resnet.py:12-92 (ATen fp32 on the CPU): the stem, every residual block and the
    final features, each within 1e-4 of the layer's max magnitude (cuDNN TF32 reads ~3e-3 here)."""
    import json, os
    from oracle import scflow_oracle as O
    from scflow_b200 import synthetic as syn
    from scflow_b200.encoder import RAFTEncoder
    es = json.load(open(os.path.join(golden_dir, "encoder_state_shapes.json")))
    sd = syn.fill_state_dict(es[norm], seed=61, gain=1.7)
    enc = RAFTEncoder(3, 256, net_type="Basic", norm_cfg=dict(type=norm)).eval()
    enc.load_state_dict(sd)
    enc.to(DEV)
    x = syn.make_crops(n, s, seed=5)["rendered_images"]
    want_taps = []
    want = O.raft_encoder(x, sd, norm, taps=want_taps)
    got_taps = []
    with torch.no_grad():
        got = enc(x.to(DEV), taps=got_taps)
        planar = enc(x.to(DEV), planar_tf32=True)
    torch.cuda.synchronize()
    assert len(got_taps) == len(want_taps) == 7
    for i, (g, w) in enumerate(zip(got_taps, want_taps)):
        e = _err(g, w.permute(0, 2, 3, 1))
        assert e < 1e-4, (i, e)
    assert got.shape == want.shape
    assert _err(got, want) < 1e-4
    # the correlation kernel's operand format: NCHW-contiguous, rounded to the nearest TF32
    assert planar.is_contiguous() and planar.shape == want.shape
    assert int((planar.view(torch.int32) & 0x1FFF).abs().max()) == 0
    assert _err(planar, want) < 6e-4


def test_norm_apply_split_vs_torch():
    """InstanceNorm + residual + ReLU from epilogue statistics (resnet.py:64-92) against F.instance_norm."""
    g = torch.Generator().manual_seed(11)
    n, h, w, c = 3, 16, 8, 96
    x = torch.randn(n, h, w, c, generator=g) * 3 + 1
    r = torch.randn(n, h, w, c, generator=g)
    xd, rd = x.to(DEV), r.to(DEV)
    def sums(t):
        d = t.double()
        full = torch.stack([d.sum((1, 2)), (d * d).sum((1, 2))], -1)              # (n, c, 2)
        part = torch.rand(C.STATS_REPLICAS, 1, 1, 1, device=t.device, dtype=torch.float64)
        return (full[None] * part / part.sum()).permute(1, 0, 2, 3).contiguous()    # replicas that add up to it
    inorm = lambda t: F.instance_norm(t.double().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
    got = C.norm_apply(xd, sums(xd), relu=True).float()
    assert _err(got, torch.relu(inorm(x))) < 2e-5
    got = C.norm_apply(xd, sums(xd), relu=True, res=C.to_split(rd)).float()
    assert _err(got, torch.relu(inorm(x) + r.double())) < 2e-5
    f32 = torch.empty(n, h, w, c, device=DEV)
    got = C.norm_apply(xd, sums(xd), relu=False, res_raw=rd, res_sums=sums(rd), out_f32=f32).float()
    assert _err(got, inorm(x) + inorm(r)) < 2e-5
    assert _err(f32, inorm(x) + inorm(r)) < 2e-6


@pytest.mark.parametrize("ks,co,n,hw,cmid", [(3, 2, 3, 32, 256), (1, 1, 3, 32, 256), (3, 2, 2, 64, 256), (3, 1, 5, 16, 96)])
def test_head_predict_vs_fp64(ks, co, n, hw, cmid):
    """A head's last 3x3 layer + ReLU + its 1-2 channel predictor (raft_decoder.py:256-294) with the predictor's tap
    products accumulated in the layer's epilogue (dot_out) and summed by scflow_tap_sum_fwd, against the two
    convolutions in fp64.  The (n, cmid, h, w) feature map is never written."""
    g = torch.Generator().manual_seed(100 * ks + co + n)
    cin = 128
    x = torch.randn(n, hw, hw, cin, generator=g)
    w1 = torch.randn(cmid, cin, 3, 3, generator=g) / (cin * 9) ** 0.5
    b1 = torch.randn(cmid, generator=g) * 0.1
    pred = torch.nn.Conv2d(cmid, co, ks, padding=ks // 2)
    with torch.no_grad():
        pred.weight.copy_(torch.randn(pred.weight.shape, generator=g) / (cmid * ks * ks) ** 0.5)
        pred.bias.copy_(torch.randn(co, generator=g))
    mid = _ref(x, w1, b1, 1, 1, relu=True)                                   # (n, h, w, cmid) fp64
    want = F.conv2d(mid.permute(0, 3, 1, 2), pred.weight.double(), pred.bias.double(), 1, ks // 2)
    pk = C.pack_weights(w1.to(DEV), b1.to(DEV), 1, (1, 1))
    got = C.head_predict(C.to_split(x.to(DEV)), pk, pred.to(DEV))
    assert got.shape == want.shape
    assert _err(got, want) < 3e-5
    got2 = C.head_predict(C.to_split(x.to(DEV)), pk, pred)                   # cached projection vectors, fresh scratch
    assert _err(got2, want) < 3e-5


@pytest.mark.parametrize("mode,tol", [("bf16x2", 3e-3), ("bf16", 6e-3)])
def test_reduced_precision_modes_are_opt_in_and_bounded(mode, tol):
    """scflow_conv_tc_desc.mma_terms: the default is three MMAs per product (2e-5); the opt-in modes drop the
    activation's low part (bf16x2) or both low parts (bf16) and stay within the tolerance DESIGN 2 states for them.
    Covers the pair, fold and plain issue paths (cout 256 / 64 / 192) and the per-tap kernel (1x1)."""
    assert C.MMA_TERMS == 3
    g = torch.Generator().manual_seed(5)
    try:
        for cin, cout, k in ((128, 256, 3), (64, 64, 3), (256, 192, 3), (128, 256, 1)):
            x = torch.randn(2, 32, 32, cin, generator=g)
            wt = torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5
            b = torch.randn(cout, generator=g)
            want = _ref(x, wt, b, 1, k // 2)
            pk = C.pack_weights(wt.to(DEV), b.to(DEV), 1, (k // 2, k // 2))
            xs = C.to_split(x.to(DEV))
            C.set_precision("bf16x3")
            e3 = _err(C.conv(xs, pk, want="f32")[..., :cout], want)
            C.set_precision(mode)
            er = _err(C.conv(xs, pk, want="f32")[..., :cout], want)
            assert e3 < TOL and TOL < er < tol, (cin, cout, k, e3, er)
    finally:
        C.set_precision("bf16x3")
