"""Deterministic synthetic YCB-V-shaped inputs and weights.

No dataset, mesh or checkpoint is available offline, so tests, ``bench.py`` and the golden
generator all draw from here.  Everything comes from ``numpy.random.RandomState`` (a frozen,
version-stable bit stream) in float64 and is cast to float32 once, so the same seed gives the
same bits in the dev container, on the GPU box and in the committed golden fixtures.
"""
from __future__ import annotations

import math
from typing import Dict

import numpy as np
import torch

# YCB-V camera (reference tools/eval.py:41)
YCBV_K = np.array([[1066.778, 0.0, 312.9869], [0.0, 1067.487, 241.3109], [0.0, 0.0, 1.0]])


def _random_rotation(rs: np.random.RandomState) -> np.ndarray:
    q, r = np.linalg.qr(rs.standard_normal((3, 3)))
    q = q * np.sign(np.diag(r))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q


def _axis_angle(axis: np.ndarray, ang: float) -> np.ndarray:
    a = axis / (np.linalg.norm(axis) + 1e-12)
    Kx = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + math.sin(ang) * Kx + (1 - math.cos(ang)) * (Kx @ Kx)


def make_crops(n: int, size: int = 256, seed: int = 1234) -> Dict[str, torch.Tensor]:
    """n object crops of size x size.

    depth: an ellipsoid (diameter U[90,270] mm) posed at (R0, t0), t0.z ~ U[500,1200] mm,
    ray-cast analytically, 0 outside; K: YCB-V focal length composed with the crop affine that
    maps a 1.1x bounding box to the crop (principal point = crop centre +- 10 px);
    (R_gt, t_gt): (R0, t0) jittered by sigma = 15 deg / 15,15,50 mm (scflow_ycbv_real.py:53-59);
    images: low-frequency texture inside the mask over 0.5 grey, real = shifted copy + noise.
    CPU float32 tensors."""
    rs = np.random.RandomState(seed)
    S = size
    depth = np.zeros((n, S, S))
    Ks = np.zeros((n, 3, 3))
    R0s = np.zeros((n, 3, 3))
    t0s = np.zeros((n, 3))
    Rgs = np.zeros((n, 3, 3))
    tgs = np.zeros((n, 3))
    rend = np.full((n, 3, S, S), 0.5)
    real = np.full((n, 3, S, S), 0.5)
    ys, xs = np.meshgrid(np.arange(S, dtype=np.float64), np.arange(S, dtype=np.float64), indexing="ij")
    for i in range(n):
        semi = rs.uniform(45.0, 135.0, size=3)
        R0 = _random_rotation(rs)
        tz = rs.uniform(500.0, 1200.0)
        t0 = np.array([rs.uniform(-60, 60), rs.uniform(-60, 60), tz])
        f = S * tz / (2.2 * semi.max())
        f = min(f, YCBV_K[0, 0] * 4.0)
        cx = S / 2.0 - f * t0[0] / tz + rs.uniform(-10, 10)
        cy = S / 2.0 - f * t0[1] / tz + rs.uniform(-10, 10)
        K = np.array([[f, 0, cx], [0, f * YCBV_K[1, 1] / YCBV_K[0, 0], cy], [0, 0, 1.0]])
        # ray d = K^-1 [x,y,1]; solve (s d - t)^T A (s d - t) = 1 for the nearest s (= depth, d.z = 1)
        A = R0 @ np.diag(1.0 / semi ** 2) @ R0.T
        d = np.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], np.ones_like(xs)], -1)
        Ad = d @ A.T
        a = (d * Ad).sum(-1)
        b = -2.0 * (Ad @ t0)
        c = t0 @ A @ t0 - 1.0
        disc = b * b - 4 * a * c
        hit = disc > 0
        s = (-b - np.sqrt(np.where(hit, disc, 0.0))) / (2 * a)
        depth[i] = np.where(hit & (s > 0), s, 0.0)
        ang = np.deg2rad(15.0) * rs.standard_normal()
        Rg = _axis_angle(rs.standard_normal(3), ang) @ R0
        tg = t0 + rs.standard_normal(3) * np.array([15.0, 15.0, 50.0])
        # texture: bilinear-upsampled 16x16 noise
        tex = rs.uniform(0, 1, size=(3, 17, 17))
        gy, gx = ys / S * 16, xs / S * 16
        y0, x0 = np.floor(gy).astype(int), np.floor(gx).astype(int)
        wy, wx = gy - y0, gx - x0
        up = ((1 - wy) * (1 - wx) * tex[:, y0, x0] + (1 - wy) * wx * tex[:, y0, x0 + 1]
              + wy * (1 - wx) * tex[:, y0 + 1, x0] + wy * wx * tex[:, y0 + 1, x0 + 1])
        m = depth[i] > 0
        rend[i] = np.where(m, up, 0.5)
        sh = rs.randint(-6, 7, size=2)
        real[i] = np.clip(np.roll(rend[i], (int(sh[0]), int(sh[1])), axis=(1, 2))
                          + 0.02 * rs.standard_normal((3, S, S)), 0, 1)
        Ks[i], R0s[i], t0s[i], Rgs[i], tgs[i] = K, R0, t0, Rg, tg
    labels = rs.randint(0, 21, size=n)
    f32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return dict(rendered_images=f32(rend), real_images=f32(real), depth=f32(depth), K=f32(Ks),
                R0=f32(R0s), t0=f32(t0s), R_gt=f32(Rgs), t_gt=f32(tgs),
                labels=torch.from_numpy(labels.astype(np.int64)))


def fill_state_dict(shapes: Dict[str, tuple], seed: int = 0, pose_head_std: float = 1e-2,
                    gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """Seeded weights for a {name: shape} spec, laid out like nn.Conv2d/Linear default init
    (uniform +-1/sqrt(fan_in)); norm weights 1, biases/running_mean 0, running_var 1.
    The pose head's output layers get N(0, pose_head_std) weights with the reference's
    biases ([1,0,0,0,1,0] per class for rotation, 0 for translation; pose_head.py:187-198) so
    that the delta pose is NOT the identity."""
    rs = np.random.RandomState(seed)
    out = {}
    for name in sorted(shapes):
        shp = tuple(shapes[name])
        if name.endswith("num_batches_tracked"):
            out[name] = torch.zeros((), dtype=torch.int64)
            continue
        if name.endswith("running_var"):
            v = np.ones(shp) + 0.1 * rs.uniform(0, 1, shp)
        elif name.endswith("running_mean"):
            v = 0.05 * rs.standard_normal(shp)
        elif ".gn." in name or ".bn" in name or "downsample.1" in name or name.split(".")[-2].startswith(("bn", "gn")):
            v = np.ones(shp) if name.endswith("weight") else np.zeros(shp)
            v = v + 0.05 * rs.standard_normal(shp)
        elif "rotation_pred" in name or "translation_pred" in name:
            if name.endswith("weight"):
                v = pose_head_std * rs.standard_normal(shp)
            elif "rotation_pred" in name:
                v = np.tile(np.array([1.0, 0, 0, 0, 1.0, 0]), shp[0] // 6)
            else:
                v = np.zeros(shp)
        elif name.endswith("weight"):
            fan_in = int(np.prod(shp[1:]))
            bound = gain / math.sqrt(fan_in)
            v = rs.uniform(-bound, bound, shp)
        else:  # bias
            v = rs.uniform(-0.05, 0.05, shp)
        out[name] = torch.from_numpy(np.asarray(v, dtype=np.float32).reshape(shp))
    return out


def make_features(n: int, h: int, w: int, seed: int = 7, c: int = 256, hc: int = 128) -> Dict[str, torch.Tensor]:
    """Seeded stand-ins for the encoder outputs the decoder consumes: feat_render, feat_real
    (n,c,h,w) ~ N(0,1) with feat_real a noisy copy (so the correlation has structure),
    h_feat = tanh(N(0,1)), cxt_feat = relu(N(0,1)) (scflow_refiner.py:101-104)."""
    rs = np.random.RandomState(seed)
    f1 = rs.standard_normal((n, c, h, w))
    f2 = 0.7 * f1 + 0.7 * rs.standard_normal((n, c, h, w))
    hf = np.tanh(rs.standard_normal((n, hc, h, w)))
    cf = np.maximum(rs.standard_normal((n, hc, h, w)), 0.0)
    f32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return dict(feat_render=f32(f1), feat_real=f32(f2), h_feat=f32(hf), cxt_feat=f32(cf))


def synthetic_loss(outs, gamma: float = 0.8, seed: int = 7) -> torch.Tensor:
    """Deterministic scalar over every differentiable output of the decoder (the 7 lists of
    scflow_decoder.py:252, or the oracle's dict): a SequenceLoss-style gamma-weighted sum of fixed
    random linear functionals of flow_from_pred, flow_from_pose, masks, rotations and translations.
    Used by the gradient fixtures, the gradient parity tests and the training bench, so that all
    three back-propagate through the same paths the reference's losses do
    (scflow_refiner.py:196-232)."""
    if isinstance(outs, dict):
        outs = (outs["flow_from_pose"], outs["flow_from_pred"], outs["rotations"], outs["translations"],
                outs["masks"])
    fpose, fpred, Rs, ts, masks = outs[:5]
    dev = fpred[0].device
    S = fpred[0].shape[-1]
    rs = np.random.RandomState(seed)
    to = lambda a: torch.from_numpy(a.astype(np.float32)).to(dev)
    Wf, Wp = to(rs.standard_normal((1, 2, S, S))), to(rs.standard_normal((1, 2, S, S)))
    Wm = to(rs.standard_normal((1, 1, S, S)))
    wr, wt = to(rs.standard_normal((1, 3, 3))), to(rs.standard_normal((1, 3)) * 1e-2)
    n = len(fpred)
    loss = fpred[0].new_zeros(())
    for i in range(n):
        g = gamma ** (n - 1 - i)
        loss = loss + g * ((fpred[i] * Wf).mean() + 0.5 * (fpose[i] * Wp).mean() + (masks[i] * Wm).mean()
                           + (Rs[i] * wr).sum(dim=(1, 2)).mean() + (ts[i] * wt).sum(dim=1).mean())
    return loss
