"""f4 host side: the renderer call of the refiner and the ADD / ADD-S evaluator as CUDA launches.

  Renderer.forward      models/utils/rendering.py:185-248 (what it asks of pytorch3d with the shipped configuration,
                        configs/refine_models/scflow_ycbv_real.py:261-274) -> dict(images, zbuf, masks): what
                        models/refiner/base_refiner.py:106-112 reads back (images[..., :3] as (N,3,H,W),
                        fragments.zbuf[..., 0], zbuf > 0)
  add_errors            ADD.eval_pose_error, metrics/add.py:354-394

Meshes are plain arrays (verts (V,3), faces (F,3), vertex colours (V,3)) packed once per Renderer / evaluator; the
reference reads them from .ply files through pytorch3d / trimesh, which stays the caller's business.  CUDA only.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import _lib
from .ops import _ptr, _stream, _timed


def _vertex_normals(verts: torch.Tensor, faces: torch.Tensor) -> torch.Tensor:
    """Area-weighted face normals summed onto the vertices, normalised (pytorch3d Meshes.verts_normals_packed)."""
    v = verts.double()
    f = faces.long()
    v0, v1, v2 = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    n = torch.zeros_like(v)
    n.index_add_(0, f[:, 1], torch.linalg.cross(v2 - v1, v0 - v1))
    n.index_add_(0, f[:, 2], torch.linalg.cross(v0 - v2, v1 - v2))
    n.index_add_(0, f[:, 0], torch.linalg.cross(v1 - v0, v2 - v0))
    return (n / n.norm(dim=1, keepdim=True).clamp_min(1e-6)).float()


class MeshSet:
    """Meshes of all classes packed into flat device arrays with per-class offsets (label -> mesh)."""

    def __init__(self, meshes: Dict[int, tuple], device):
        labels = sorted(meshes)
        if labels != list(range(len(labels))):
            raise _lib.ScflowError("MeshSet: labels must be 0..C-1")
        verts, faces, colors, normals, voff, foff = [], [], [], [], [0], [0]
        for l in labels:
            v, f = meshes[l][0], meshes[l][1]
            c = meshes[l][2] if len(meshes[l]) > 2 and meshes[l][2] is not None else np.full_like(np.asarray(v), 0.7)
            v = torch.as_tensor(np.asarray(v), dtype=torch.float32)
            f = torch.as_tensor(np.asarray(f), dtype=torch.int32)
            verts.append(v); faces.append(f); colors.append(torch.as_tensor(np.asarray(c), dtype=torch.float32))
            normals.append(_vertex_normals(v, f))
            voff.append(voff[-1] + v.shape[0]); foff.append(foff[-1] + f.shape[0])
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.ScflowError("MeshSet: expected a CUDA device (scflow_b200 has no CPU path)")
        self.verts = torch.cat(verts).contiguous().to(dev)
        self.faces = torch.cat(faces).contiguous().to(dev)
        self.colors = torch.cat(colors).contiguous().to(dev)
        self.normals = torch.cat(normals).contiguous().to(dev)
        self.voff = torch.tensor(voff, dtype=torch.int32, device=dev)
        self.foff = torch.tensor(foff, dtype=torch.int32, device=dev)
        self.vmax = max(b - a for a, b in zip(voff, voff[1:]))
        self.fmax = max(b - a for a, b in zip(foff, foff[1:]))
        self.num_classes = len(labels)


class Renderer(torch.nn.Module):
    """Drop-in for the call `self.renderer(rotations, translations, internel_k, labels)` of the refiner with the
    shipped configuration (hard Phong, one face per pixel, separate lights, background 0.5)."""

    def __init__(self, meshes: MeshSet, image_size, background_color=(0.5, 0.5, 0.5)):
        super().__init__()
        self.meshes, self.image_size, self.background = meshes, tuple(image_size), tuple(background_color)

    @torch.no_grad()
    def forward(self, rotations, translations, internel_k, labels):
        lib = _lib.load()
        n = rotations.shape[0]
        assert translations.shape[0] == n == internel_k.shape[0] == labels.shape[0]
        if not rotations.is_cuda:
            raise _lib.ScflowError("Renderer: expected CUDA tensors (scflow_b200 has no CPU path)")
        H, W = self.image_size
        m = self.meshes
        dev = rotations.device
        R = rotations.float().contiguous(); t = translations.float().reshape(n, 3).contiguous()
        K = internel_k.float().contiguous(); lab = labels.to(torch.int64).contiguous()
        image = torch.empty((n, 3, H, W), device=dev, dtype=torch.float32)
        zbuf = torch.empty((n, H, W), device=dev, dtype=torch.float32)
        mask = torch.empty((n, H, W), device=dev, dtype=torch.float32)
        ws = torch.empty(lib.scflow_render_workspace_bytes(n, m.vmax, H, W), device=dev, dtype=torch.uint8)
        with _timed("scflow_render_fwd", 5):
            rc = lib.scflow_render_fwd(_ptr(m.verts), _ptr(m.normals), _ptr(m.colors), _ptr(m.faces), _ptr(m.voff),
                                       _ptr(m.foff), m.vmax, m.fmax, _ptr(lab), _ptr(R), _ptr(t), _ptr(K), n, H, W,
                                       self.background[0], self.background[1], self.background[2], _ptr(image),
                                       _ptr(zbuf), _ptr(mask), _ptr(ws), _stream())
        _lib.check(rc, "scflow_render_fwd")
        return dict(images=image, zbuf=zbuf, masks=mask)


@torch.no_grad()
def add_errors(meshes: MeshSet, gt_r, gt_t, pred_r, pred_t, labels, k, symmetric: Sequence[bool],
               mesh_diameters: Optional[Sequence[float]] = None):
    """(error_3d_normalized, error_2d, error_3d), each (N,) float64 on the device: ADD.eval_pose_error
    (metrics/add.py:354-394).  symmetric[c]: class c uses ADD-S."""
    lib = _lib.load()
    n = gt_r.shape[0]
    dev = meshes.verts.device
    f = lambda x: torch.as_tensor(x).to(dev).float().reshape(n, -1).contiguous()
    gr, gt, pr, pt, kk = f(gt_r), f(gt_t), f(pred_r), f(pred_t), f(k)
    lab = torch.as_tensor(labels).to(dev).to(torch.int64).contiguous()
    sym = torch.tensor([1 if s else 0 for s in symmetric], dtype=torch.int32, device=dev)
    assert sym.numel() == meshes.num_classes
    e3 = torch.empty(n, device=dev, dtype=torch.float64)
    e2 = torch.empty(n, device=dev, dtype=torch.float64)
    with _timed("scflow_add_error_fwd", 1):
        rc = lib.scflow_add_error_fwd(_ptr(meshes.verts), _ptr(meshes.voff), meshes.vmax, _ptr(lab), _ptr(sym),
                                      _ptr(gr), _ptr(gt), _ptr(pr), _ptr(pt), _ptr(kk), _ptr(e3), _ptr(e2), n, _stream())
    _lib.check(rc, "scflow_add_error_fwd")
    if mesh_diameters is None:
        return None, e2, e3
    diam = torch.as_tensor(np.asarray(mesh_diameters, np.float64), device=dev)[lab]
    return e3 / diam, e2, e3
