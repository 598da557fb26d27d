"""Functional mirror of models/utils/pose.py for the hot path (same names and argument meaning).

The reference keeps ragged per-crop point lists built with ``nonzero`` (a host sync per crop).
Here ``cal_3d_2d_corr`` accepts a whole batch and returns a dense ``PointSet``; the two
``*_list`` arguments of ``get_flow_from_delta_pose_and_points`` accept that PointSet, or the reference's ragged
lists (scattered into a PointSet on the device)."""
from __future__ import annotations

from dataclasses import dataclass

import torch

from . import ops


@dataclass
class PointSet:
    """Dense stand-in for (points_2d_list, points_3d_list): object-frame points (N,3,H,W), zero
    on background, and the foreground mask (N,H,W) uint8.  points_2d is implicit (pixel grid)."""
    points3d: torch.Tensor
    mask: torch.Tensor

    def __len__(self):
        return self.points3d.shape[0]


def cal_3d_2d_corr(depth, internel_k, rotation, translation, occlusion=None):
    """pose.py:44-64.  depth (N,H,W) or (H,W); K (N,3,3); R (N,3,3); t (N,3)."""
    if occlusion is not None:
        raise NotImplementedError("occlusion mask is not used on the SCFlow path")
    if depth.dim() == 2:
        depth, internel_k = depth[None], internel_k[None]
        rotation, translation = rotation[None], translation[None]
    pts, mask = ops.lift_points(depth, internel_k, rotation, translation)
    ps = PointSet(pts, mask)
    return ps, ps


def get_flow_from_delta_pose_and_points(rotation_dst, translation_dst, k, points_2d_list,
                                        points_3d_list, height, width, invalid_num=400.):
    """pose.py:66-88 -> (n, 2, height, width)."""
    ps = points_3d_list
    if not isinstance(ps, PointSet):
        # the reference's ragged form (what a RAFTDecoder-style caller built with the reference's own
        # cal_3d_2d_corr): per image (Ni, 2) pixel coordinates (x, y) and (Ni, 3) object-frame points, scattered
        # into the dense PointSet the kernel reads (device-side indexing, no host synchronisation)
        n = len(points_3d_list)
        if len(points_2d_list) != n or n != rotation_dst.shape[0]:
            raise ValueError("points_2d_list / points_3d_list must hold one entry per image")
        dev = rotation_dst.device
        pts = torch.zeros((n, 3, height, width), device=dev, dtype=torch.float32)
        msk = torch.zeros((n, height, width), device=dev, dtype=torch.uint8)
        for i, (p2, p3) in enumerate(zip(points_2d_list, points_3d_list)):
            x, y = p2[:, 0].to(torch.int64), p2[:, 1].to(torch.int64)
            pts[i, :, y, x] = p3.t().float()
            msk[i, y, x] = 1
        ps = PointSet(pts, msk)
    if torch.is_grad_enabled() and (rotation_dst.requires_grad or translation_dst.requires_grad):
        raise NotImplementedError("differentiable pose-induced flow: use scflow_b200.autograd.pose_flow "
                                  "(depth path), as SCFlowDecoder does when training")
    _, _, full, _ = ops.pose_flow(rotation_dst, translation_dst, k, points3d=ps.points3d,
                                  mask=ps.mask, invalid_num=invalid_num, want_lr=False)
    return full


def get_pose_from_delta_pose(rotation_delta, translation_delta, rotation_src, translation_src,
                             weight=10., depth_transform='exp', detach_depth_for_xy=False):
    """pose.py:124-149 (ortho6d + 'exp' only).  K is not needed: only the 3x3 algebra runs."""
    if rotation_delta.size(1) != 6:
        raise NotImplementedError("quaternion deltas (kornia path) are not implemented")
    if depth_transform != 'exp':
        raise NotImplementedError("depth_transform != 'exp'")
    if torch.is_grad_enabled() and any(t.requires_grad for t in (rotation_delta, translation_delta,
                                                                 rotation_src, translation_src)):
        from . import autograd
        return autograd.compose_pose(rotation_delta, translation_delta, rotation_src, translation_src,
                                     weight=weight, detach_depth_for_xy=detach_depth_for_xy)
    n = rotation_src.shape[0]
    eye = torch.eye(3, device=rotation_src.device).expand(n, 3, 3).contiguous()
    R, t, _, _ = ops.pose_flow(rotation_src, translation_src, eye, dR6=rotation_delta,
                               dt=translation_delta, depth=torch.zeros(n, 4, 4, device=eye.device),
                               R0=eye, t0=translation_src, weight=weight, want_full=False,
                               want_lr=False)
    return R, t
