"""CPU restatement of SCFlow's recurrent shape-constrained refinement hot path.

TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline / ``--impl reference`` legs may import this module.
Nothing under ``scflow_b200/`` imports it; the product path has no CPU fallback.

Every function restates one reference function (cited file:line, relative to
GiaKhangLuu/SCFlow) in plain fp32 torch on the CPU.  Third-party arithmetic the
reference reaches through PyTorch (``matmul``, ``avg_pool2d``, ``grid_sample``,
``interpolate``, ``conv2d``, ``inverse``) is restated from ATen's published
semantics where the CUDA kernels need the exact formula (bilinear window
selection, align_corners resampling, pooling order) and called through torch
where it is a plain library primitive (conv2d, matmul, group_norm).

Parity pin: ``oracle/gen_golden.py`` runs the UNMODIFIED reference files (imported from
/root/reference through ``oracle/_ref_shims.py``) on seeded inputs and stores the outputs
under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks this restatement against
them.  The reference itself ships no golden vectors or tests for this path
(SURVEY.md section 4), so the pin is "reference code run here", torch 2.11.0 CPU.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------------------
# a1  CorrelationPyramid.forward            models/decoder/raft_decoder.py:35-58
# ----------------------------------------------------------------------------------------
def avg_pool_2x2(x: Tensor) -> Tensor:
    """nn.AvgPool2d(2, 2) restated: window sum in (kh outer, kw inner) order, then /4
    (ATen avg_pool2d; raft_decoder.py:32,54-56).  x: (B, 1, H, W), H and W even."""
    a = x[..., 0::2, 0::2]
    b = x[..., 0::2, 1::2]
    c = x[..., 1::2, 0::2]
    d = x[..., 1::2, 1::2]
    return (((a + b) + c) + d) / 4.0


def correlation_pyramid(feat1: Tensor, feat2: Tensor, num_levels: int = 4) -> List[Tensor]:
    """corr[n, q, t] = sum_c feat1[n, c, q] * feat2[n, c, t] / sqrt(C); then
    (num_levels-1) 2x2 means over the t plane (raft_decoder.py:47-56).
    Returns tensors of shape (N*H*W, 1, H>>l, W>>l)."""
    N, C, H, W = feat1.shape
    a = feat1.reshape(N, C, H * W).permute(0, 2, 1)
    b = feat2.reshape(N, C, H * W)
    corr = torch.matmul(a, b).reshape(N * H * W, 1, H, W)
    corr = corr / torch.sqrt(torch.tensor(float(C)))
    pyr = [corr]
    for _ in range(num_levels - 1):
        pyr.append(avg_pool_2x2(pyr[-1]))
    return pyr


# ----------------------------------------------------------------------------------------
# a2  CorrLookup.forward                     models/utils/corr_lookup.py:102-136
#     bilinear_sample                        models/utils/corr_lookup.py:31-67
#     (+ ATen grid_sampler_2d, bilinear / zeros / align_corners=True)
# ----------------------------------------------------------------------------------------
DIV_IEEE = 0  # `x * 2. / max(W-1, 1)` evaluated as a true fp32 division (ATen CPU)
DIV_RCP = 1   # ... evaluated as x*2 * (1/max(W-1,1)) (ATen CUDA div-by-scalar fast path)


def lookup_axis_coords(centre: Tensor, size: int, level: int, radius: int,
                       div_mode: int = DIV_IEEE) -> Tensor:
    """Un-normalised sample coordinate along one axis for the 2r+1 taps.

    Replays, one separately rounded fp32 op at a time:
      corr_lookup.py:127  centroid = grid / 2**i
      corr_lookup.py:128  coords   = centroid + delta           (delta = -r..r)
      corr_lookup.py:64   g = coords * 2. / max(size-1, 1) - 1.
      ATen grid_sampler_unnormalize(align_corners=True):  ((g + 1) / 2) * (size - 1)
    centre: (...,) fp32 pixel coordinate (x + flow_x or y + flow_y).  Returns (..., 2r+1)."""
    c = centre.to(torch.float32) / float(2 ** level)
    delta = torch.linspace(-radius, radius, 2 * radius + 1, dtype=torch.float32)
    c = c[..., None] + delta
    d = float(max(size - 1, 1))
    g = c * 2.0
    if div_mode == DIV_IEEE:
        g = g / d
    else:
        g = g * (torch.tensor(1.0, dtype=torch.float32) / torch.tensor(d, dtype=torch.float32))
    g = g - 1.0
    return ((g + 1.0) / 2.0) * float(size - 1)


def corr_lookup(pyramid: Sequence[Tensor], flow: Tensor, radius: int = 4,
                div_mode: int = DIV_IEEE, return_windows: bool = False):
    """(N, L*(2r+1)^2, H, W) lookup.  Tap (a, b) of level l samples row q of level l at
    x = (x_q + flow_x)/2^l + (a - r),  y = (y_q + flow_y)/2^l + (b - r): the FIRST delta
    component is added to x (corr_lookup.py:118-128), so channel = T*l + (2r+1)*a + b with
    a the x offset.  Zero padding applies per corner (ATen within_bounds_2d).

    With return_windows also returns int32 (N*H*W, L, 2, 2r+1): floor(ix) per a, floor(iy)
    per b -- the "window selection" that must match bit-exactly."""
    N, _, H, W = flow.shape
    D = 2 * radius + 1
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, W).expand(N, H, W)
    ys = torch.arange(H, dtype=torch.float32).view(1, H, 1).expand(N, H, W)
    cx = (xs + flow[:, 0]).reshape(-1)  # corr_lookup.py:115
    cy = (ys + flow[:, 1]).reshape(-1)
    Q = N * H * W
    outs, wins = [], []
    for l, corr in enumerate(pyramid):
        Hl, Wl = corr.shape[-2:]
        img = corr.reshape(Q, Hl * Wl)
        ix = lookup_axis_coords(cx, Wl, l, radius, div_mode)  # (Q, D) indexed by a
        iy = lookup_axis_coords(cy, Hl, l, radius, div_mode)  # (Q, D) indexed by b
        x0 = torch.floor(ix)
        y0 = torch.floor(iy)
        wins.append(torch.stack([x0, y0], 1).to(torch.int32))
        # ATen grid_sampler_2d (CUDA form): nw=(x1-ix)(y1-iy) ne=(ix-x0)(y1-iy)
        #                                   sw=(x1-ix)(iy-y0) se=(ix-x0)(iy-y0)
        wx1 = ix - x0
        wx0 = (x0 + 1.0) - ix
        wy1 = iy - y0
        wy0 = (y0 + 1.0) - iy
        x0i = x0.to(torch.int64)
        y0i = y0.to(torch.int64)
        acc = torch.zeros(Q, D, D, dtype=torch.float32)
        for dy, wy in ((0, wy0), (1, wy1)):
            for dx, wx in ((0, wx0), (1, wx1)):
                xi = (x0i + dx)[:, :, None]          # (Q, a, 1)
                yi = (y0i + dy)[:, None, :]          # (Q, 1, b)
                ok = (xi >= 0) & (xi < Wl) & (yi >= 0) & (yi < Hl)
                idx = (yi.clamp(0, Hl - 1) * Wl + xi.clamp(0, Wl - 1)).reshape(Q, D * D)
                v = torch.gather(img, 1, idx).reshape(Q, D, D)
                w = wx[:, :, None] * wy[:, None, :]
                acc = acc + torch.where(ok, v * w, torch.zeros((), dtype=torch.float32))
        outs.append(acc.reshape(N, H, W, D * D))     # channel within level = D*a + b
    out = torch.cat(outs, dim=-1).permute(0, 3, 1, 2).contiguous().float()
    if return_windows:
        return out, torch.stack(wins, 1)
    return out


# ----------------------------------------------------------------------------------------
# a7  align_corners=True bilinear resampling  models/decoder/scflow_decoder.py:189-198,223-228
#     (ATen upsample_bilinear2d, area_pixel_compute_source_index with align_corners)
# ----------------------------------------------------------------------------------------
def resize_bilinear_ac(x: Tensor, out_h: int, out_w: int) -> Tensor:
    """F.interpolate(x, mode='bilinear', align_corners=True) restated.
    src = dst * (in-1)/(out-1) in fp32; i0 = int(src); i1 = i0 + (i0 < in-1);
    l1 = src - i0; l0 = 1 - l1;  out = h0*(w0*v00 + w1*v01) + h1*(w0*v10 + w1*v11)."""
    _, _, H, W = x.shape

    def axis(n_in, n_out):
        scale = torch.tensor((n_in - 1) / (n_out - 1) if n_out > 1 else 0.0, dtype=torch.float32)
        src = torch.arange(n_out, dtype=torch.float32) * scale
        i0 = src.to(torch.int64)
        i1 = i0 + (i0 < n_in - 1).to(torch.int64)
        l1 = src - i0.to(torch.float32)
        l0 = 1.0 - l1
        return i0, i1, l0, l1

    y0, y1, hy0, hy1 = axis(H, out_h)
    x0, x1, wx0, wx1 = axis(W, out_w)
    r0 = x[:, :, y0]
    r1 = x[:, :, y1]
    top = wx0 * r0[..., x0] + wx1 * r0[..., x1]
    bot = wx0 * r1[..., x0] + wx1 * r1[..., x1]
    return hy0[:, None] * top + hy1[:, None] * bot


# ----------------------------------------------------------------------------------------
# a3  ConvGRU.forward (SeqConv)              models/decoder/raft_decoder.py:235-253
# ----------------------------------------------------------------------------------------
def conv_gru(h: Tensor, x: Tensor, sd: Dict[str, Tensor], prefix: str = "") -> Tensor:
    """Two passes, (1,5) pad (0,2) then (5,1) pad (2,0).  sd keys follow the reference
    state_dict: conv_{z,r,q}.{0,1}.conv.{weight,bias}."""
    for p, pad in ((0, (0, 2)), (1, (2, 0))):
        def g(name, inp):
            return F.conv2d(inp, sd[f"{prefix}conv_{name}.{p}.conv.weight"],
                            sd[f"{prefix}conv_{name}.{p}.conv.bias"], padding=pad)
        hx = torch.cat([h, x], dim=1)
        z = torch.sigmoid(g("z", hx))
        r = torch.sigmoid(g("r", hx))
        q = torch.tanh(g("q", torch.cat([r * h, x], dim=1)))
        h = (1 - z) * h + z * q
    return h


# ----------------------------------------------------------------------------------------
# a4  cal_3d_2d_corr / lift_2d_to_3d         models/utils/pose.py:26-64
# ----------------------------------------------------------------------------------------
def lift_points(depth: Tensor, K: Tensor, R0: Tensor, t0: Tensor) -> Tuple[Tensor, Tensor]:
    """Ragged form, one crop: foreground (depth>0) pixels in row-major order.
    X_cam = K^-1 [x*d, y*d, d]^T;  X_obj = R0^-1 (X_cam - t0).  Returns (P,2) xy, (P,3)."""
    ys, xs = torch.nonzero(depth > 0, as_tuple=True)
    d = depth[ys, xs]
    homo = torch.stack([xs.float(), ys.float(), torch.ones_like(d)], -1) * d[:, None]
    cam = torch.mm(torch.inverse(K), homo.t()).t()
    obj = torch.mm(torch.inverse(R0), (cam - t0[None]).t()).t()
    return torch.stack([xs, ys], -1).float(), obj


def lift_points_dense(depth: Tensor, K: Tensor, R0: Tensor, t0: Tensor) -> Tuple[Tensor, Tensor]:
    """Dense batched form of the same arithmetic: (N,3,S,S) object-frame points and the
    (N,S,S) foreground mask.  Background entries are 0."""
    N, H, W = depth.shape
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, W).expand(N, H, W)
    ys = torch.arange(H, dtype=torch.float32).view(1, H, 1).expand(N, H, W)
    homo = torch.stack([xs * depth, ys * depth, depth], 1).reshape(N, 3, H * W)
    cam = torch.bmm(torch.inverse(K), homo)
    obj = torch.bmm(torch.inverse(R0), cam - t0[:, :, None]).reshape(N, 3, H, W)
    mask = depth > 0
    return obj * mask[:, None], mask


# ----------------------------------------------------------------------------------------
# a5  get_flow_from_delta_pose_and_points    models/utils/pose.py:66-88
# ----------------------------------------------------------------------------------------
def pose_flow_dense(R: Tensor, t: Tensor, K: Tensor, pts: Tensor, mask: Tensor,
                    invalid_num: float = 0.0) -> Tensor:
    """p = K (R X_obj + t); flow = p.xy / p.z - (x, y) on foreground, invalid_num elsewhere.
    pts (N,3,H,W), mask (N,H,W) -> (N,2,H,W)."""
    N, _, H, W = pts.shape
    p = torch.bmm(K, torch.bmm(R, pts.reshape(N, 3, -1)) + t[:, :, None]).reshape(N, 3, H, W)
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, W)
    ys = torch.arange(H, dtype=torch.float32).view(1, H, 1)
    z = torch.where(mask, p[:, 2], torch.ones_like(p[:, 2]))
    fx = p[:, 0] / z - xs
    fy = p[:, 1] / z - ys
    inv = torch.full_like(fx, invalid_num)
    return torch.stack([torch.where(mask, fx, inv), torch.where(mask, fy, inv)], 1)


# ----------------------------------------------------------------------------------------
# a6  get_pose_from_delta_pose / get_rotation_matrix_from_ortho6d   models/utils/pose.py:124-169
# ----------------------------------------------------------------------------------------
def rotation_from_ortho6d(o: Tensor) -> Tensor:
    x = F.normalize(o[:, 0:3], p=2, dim=1)
    z = F.normalize(torch.cross(x, o[:, 3:6], dim=1), p=2, dim=1)
    y = torch.cross(z, x, dim=1)
    return torch.stack([x, y, z], dim=2)  # columns


def pose_compose(dR6: Tensor, dt: Tensor, R: Tensor, t: Tensor, weight: float = 10.0,
                 detach_depth_for_xy: bool = False):
    """R' = dR R;  tz' = tz / exp(dtz);  tx' = tz' (dtx/weight + tx/tz);  ty' likewise
    (depth_transform='exp'; detach_depth_for_xy only affects gradients, pose.py:142-144)."""
    Rn = torch.bmm(rotation_from_ortho6d(dR6), R)
    vz = t[:, 2] / torch.exp(dt[:, 2])
    vz_xy = vz.detach() if detach_depth_for_xy else vz
    vx = vz_xy * (dt[:, 0] / weight + t[:, 0] / t[:, 2])
    vy = vz_xy * (dt[:, 1] / weight + t[:, 1] / t[:, 2])
    return Rn, torch.stack([vx, vy, vz], -1)


# ----------------------------------------------------------------------------------------
# non-hot-path sub-modules of the loop (cuDNN in the product), restated functionally
#     MotionEncoder raft_decoder.py:152-166, XHead :288-294, MultiClassPoseHead pose_head.py:201-211
# ----------------------------------------------------------------------------------------
def _conv(x, sd, key, padding=0, stride=1, act=None):
    y = F.conv2d(x, sd[key + ".weight"], sd.get(key + ".bias"), stride=stride, padding=padding)
    if act == "relu":
        y = F.relu(y)
    return y


def motion_encoder(corr: Tensor, flow: Tensor, sd, p="encoder.") -> Tensor:
    c = _conv(corr, sd, p + "corr_net.0.conv", 0, act="relu")
    c = _conv(c, sd, p + "corr_net.1.conv", 1, act="relu")
    f = _conv(flow, sd, p + "flow_net.0.conv", 3, act="relu")
    f = _conv(f, sd, p + "flow_net.1.conv", 1, act="relu")
    o = _conv(torch.cat([c, f], 1), sd, p + "out_net.0.conv", 1, act="relu")
    return torch.cat([o, flow], 1)


def xhead(h: Tensor, sd, p: str, kind: str) -> Tensor:
    y = _conv(h, sd, p + "layers.0.conv", 1, act="relu")
    return _conv(y, sd, p + "predict_layer", 1 if kind == "flow" else 0)


def pose_head(x: Tensor, label: Tensor, sd, p="pose_pred.", num_class=21, groups=32):
    for i in range(3):
        x = F.conv2d(x, sd[f"{p}conv_layers.{i}.conv.weight"], None, stride=2, padding=1)
        x = F.group_norm(x, groups, sd[f"{p}conv_layers.{i}.gn.weight"], sd[f"{p}conv_layers.{i}.gn.bias"])
        x = F.relu(x)
    x = x.flatten(1)
    for i in range(2):
        x = F.relu(F.linear(x, sd[f"{p}fc_layers.{i}.0.weight"], sd[f"{p}fc_layers.{i}.0.bias"]))
    dt = F.linear(x, sd[p + "translation_pred.weight"], sd[p + "translation_pred.bias"]).view(-1, num_class, 3)
    dr = F.linear(x, sd[p + "rotation_pred.weight"], sd[p + "rotation_pred.bias"]).view(-1, num_class, 6)
    # pose_head.py:209-210: index_select(dim=1, index=label)[:, 0, :] -- every crop takes the
    # class of label[0] (reference behaviour, reproduced as is).
    return dr[:, label[0]], dt[:, label[0]]


# ----------------------------------------------------------------------------------------
# a8  SCFlowDecoder.forward                  models/decoder/scflow_decoder.py:151-252
# ----------------------------------------------------------------------------------------
def decoder_forward(sd: Dict[str, Tensor], feat_render, feat_real, h_feat, cxt_feat,
                    ref_rotation, ref_translation, depth, K, label, init_flow,
                    invalid_flow_num: float = 0.0, iters: int = 8, num_levels: int = 4,
                    radius: int = 4, div_mode: int = DIV_IEEE, train_detach: bool = False):
    """`train_detach` reproduces the shipped config's gradient cuts (scflow_ycbv_real.py:213-216:
    detach_flow / detach_pose / detach_depth_for_xy = True; scflow_decoder.py:193-194,231-232) --
    it changes no forward value."""
    pyr = correlation_pyramid(feat_render, feat_real, num_levels)
    scale = 2 ** (num_levels - 1)
    N, H, W = depth.shape
    pts, mask = lift_points_dense(depth, K, ref_rotation, ref_translation)
    R, t = ref_rotation, ref_translation
    flow = init_flow
    outs = dict(flow_from_pose=[], flow_from_pred=[], rotations=[], translations=[],
                masks=[], delta_rotations=[], delta_translations=[])
    h = h_feat
    for _ in range(iters):
        if train_detach:
            flow, R, t = flow.detach(), R.detach(), t.detach()
        flow_lr = (1.0 / scale) * resize_bilinear_ac(flow, H // scale, W // scale)
        corr = corr_lookup(pyr, flow_lr, radius, div_mode)
        motion = motion_encoder(corr, flow_lr, sd)
        h = conv_gru(h, torch.cat([cxt_feat, motion], 1), sd, "gru.")
        dflow = xhead(h, sd, "flow_pred.", "flow")
        m = torch.sigmoid(xhead(h, sd, "mask_pred.", "mask"))
        df = _conv(_conv(dflow, sd, "delta_flow_encoder.0.conv", 3, act="relu"), sd,
                   "delta_flow_encoder.1.conv", 1, act="relu")
        mf = _conv(_conv(m, sd, "mask_encoder.0.conv", 1, act="relu"), sd,
                   "mask_encoder.1.conv", 1, act="relu")
        dR6, dt = pose_head(torch.cat([h, df, mf], 1), label, sd)
        flow_pred = scale * resize_bilinear_ac(flow_lr + dflow, H, W)
        mask_up = resize_bilinear_ac(m, H, W)
        R, t = pose_compose(dR6, dt, R, t, detach_depth_for_xy=train_detach)
        flow = pose_flow_dense(R, t, K, pts, mask, invalid_flow_num)
        outs["flow_from_pose"].append(flow)
        outs["flow_from_pred"].append(flow_pred)
        outs["rotations"].append(R)
        outs["translations"].append(t)
        outs["masks"].append(mask_up)
        outs["delta_rotations"].append(dR6)
        outs["delta_translations"].append(dt)
    return outs


# ----------------------------------------------------------------------------------------
# RAFTEncoder 'Basic' (upstream of the hot path; inside the timed end-to-end metric)
#     models/encoder/raft_encoder.py:286-314, models/backbone/resnet.py:12-92,676-771
# ----------------------------------------------------------------------------------------
def raft_encoder(x: Tensor, sd: Dict[str, Tensor], norm: str, p: str = "", taps: list = None) -> Tensor:
    """`taps`, when given, collects the stem output and each residual block's output (per-layer parity)."""
    def nrm(y, key):
        if norm == "IN":
            return F.instance_norm(y)
        return F.batch_norm(y, sd[key + ".running_mean"], sd[key + ".running_var"],
                            sd[key + ".weight"], sd[key + ".bias"], False)
    n1, n2 = ("in1", "in2") if norm == "IN" else ("bn1", "bn2")
    y = F.relu(nrm(F.conv2d(x, sd[p + "conv1.weight"], sd[p + "conv1.bias"], stride=2, padding=3), p + n1))
    if taps is not None:
        taps.append(y)
    for li, stride in ((1, 1), (2, 2), (3, 2)):
        for bi in range(2):
            q = f"{p}res_layer{li}.{bi}."
            s = stride if bi == 0 else 1
            o = F.relu(nrm(F.conv2d(y, sd[q + "conv1.weight"], sd[q + "conv1.bias"], stride=s, padding=1), q + n1))
            o = nrm(F.conv2d(o, sd[q + "conv2.weight"], sd[q + "conv2.bias"], padding=1), q + n2)
            if (q + "downsample.0.weight") in sd:
                y = nrm(F.conv2d(y, sd[q + "downsample.0.weight"], sd[q + "downsample.0.bias"], stride=s),
                        q + "downsample.1")
            y = F.relu(o + y)
            if taps is not None:
                taps.append(y)
    return F.conv2d(y, sd[p + "conv2.weight"], sd[p + "conv2.bias"])


def refiner_forward(sd_enc, sd_cxt, sd_dec, render_images, real_images, R0, t0, depth, K, label,
                    iters: int = 8, div_mode: int = DIV_IEEE, train_detach: bool = False):
    """SCFlowRefiner.get_pose (models/refiner/scflow_refiner.py:84-138): shared IN encoder on
    real and rendered images, BN context encoder on the rendered image, tanh/relu split,
    zero init flow, invalid_flow_num = 0."""
    feat_real = raft_encoder(real_images, sd_enc, "IN")
    feat_render = raft_encoder(render_images, sd_enc, "IN")
    cxt = raft_encoder(render_images, sd_cxt, "BN")
    h, c = torch.split(cxt, [128, 128], dim=1)
    h, c = torch.tanh(h), torch.relu(c)
    N, _, H, W = real_images.shape
    init_flow = torch.zeros(N, 2, H, W, dtype=torch.float32)
    return decoder_forward(sd_dec, feat_render, feat_real, h, c, R0, t0, depth, K, label,
                           init_flow, 0.0, iters, div_mode=div_mode, train_detach=train_detach)


def rotation_geodesic(Ra: Tensor, Rb: Tensor) -> Tensor:
    """fp64 chordal form 2*asin(||Ra-Rb||_F / (2*sqrt(2))) (the fp32 acos form
    cannot resolve below ~1e-3 rad)."""
    d = (Ra.double() - Rb.double()).flatten(-2).norm(dim=-1)
    return 2.0 * torch.asin(torch.clamp(d / (2.0 * math.sqrt(2.0)), max=1.0))
