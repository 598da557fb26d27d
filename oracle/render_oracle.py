"""CPU (numpy) restatement of the two f4 pieces of SURVEY 8f around the hot path.  TEST INFRASTRUCTURE ONLY: imported
by tests/ and tools/, never by scflow_b200/.

1. `add_errors` -- ADD / ADD-S / 2-D projection error of `metrics/add.py:354-394` (`ADD.eval_pose_error`) with
   `datasets/pose.py:18-70` (`project_3d_point`).  PINNED: tests/golden/add_eval.npz holds the outputs of the
   UNMODIFIED reference method (oracle/gen_golden_add.py imports metrics/add.py through stubs for its uninstalled
   imports) and tests/test_oracle_golden.py holds this restatement to them.

2. `render` -- what `models/utils/rendering.py:185-248` (`Renderer.forward`) asks of pytorch3d with the shipped
   configuration (`configs/refine_models/scflow_ycbv_real.py:261-274`: HardPhongShader, faces_per_pixel=1,
   blur_radius=0, seperate_lights=True, background 0.5) and what `models/refiner/base_refiner.py:106-112` reads back
   (images[..., :3], fragments.zbuf[..., 0], mask = zbuf > 0).  PARITY UNPINNED: pytorch3d is a third-party
   dependency of the reference (requirements: pytorch3d, version not pinned by the repository) and is absent from
   this image, and the reference ships no rendered fixture; this restates pytorch3d's published conventions:
     * cameras_from_opencv_projection (rendering.py:17-47): an OpenCV camera (K, R, t) mapped to NDC with
       scale (min(W,H)-1)/2 and c0 = (size-1)/2, while the rasteriser samples pixel (i, j) at NDC
       1 - (2j+1)/W, 1 - (2i+1)/H: together pixel j samples the OpenCV image coordinate u = (j + 0.5) (W-1)/W;
     * rasterize_meshes, blur_radius 0, one face per pixel, no back-face culling: a pixel belongs to a face when
       its sample lies inside the projected triangle; nearest view-space z wins; zbuf = perspective-correct
       interpolated view z, -1 where no face;
     * HardPhongShader with TexturesVertex: perspective-correct barycentric interpolation of world position,
       vertex normal (area-weighted face normals, normalised) and vertex colour; PointLights(location) with the
       default colours ambient 0.5, diffuse 0.3, specular 0.2, Materials() defaults (all 1, shininess 64):
       colour = (ambient + diffuse) * texel + specular; background where no face.
       Light location: rendering.py:203-206,214-219: R @ (0, 0, max(min_z - 400, 0)) with min_z the smallest
       camera-frame z of the object's vertices.
"""
import numpy as np


# --------------------------------------------------------------------------------------- ADD / ADD-S
def project_3d_point(pt3d, K, rotation, translation):
    """datasets/pose.py:18-70, multi-image branch: (N, n, 2) pixel coordinates and (N, n, 3) camera points."""
    pts_cam = np.matmul(rotation, pt3d.transpose()) + translation          # (N, 3, n)
    pts_2d = np.matmul(K, pts_cam).transpose((0, 2, 1))                    # torch.bmm in the reference
    pts_2d = pts_2d.copy()
    pts_2d[..., 0] = pts_2d[..., 0] / (pts_2d[..., -1] + 1e-8)
    pts_2d[..., 1] = pts_2d[..., 1] / (pts_2d[..., -1] + 1e-8)
    return pts_2d[..., :-1], pts_cam.transpose((0, 2, 1))


def add_errors(verts_list, gt_t, gt_r, pred_t, pred_r, labels, k, symmetry_types, mesh_diameters):
    """metrics/add.py:354-394.  gt_t / pred_t (N, 3, 1); returns (error_3d_normalized, error_2d, error_3d)."""
    n = len(gt_t)
    e3n, e2, e3 = np.zeros(n), np.zeros(n), np.zeros(n)
    for i in np.unique(labels):
        idx = labels == i
        verts = verts_list[i]
        gt_2d, gt_3d = project_3d_point(verts, k[idx], gt_r[idx], gt_t[idx])
        pr_2d, pr_3d = project_3d_point(verts, k[idx], pred_r[idx], pred_t[idx])
        if symmetry_types.get(f"cls_{i + 1}", False):
            out = []
            for g, p in zip(gt_3d, pr_3d):
                d = np.linalg.norm(g[:, None, :] - p[None, :, :], axis=-1)
                out.append(p[np.argmin(d, axis=-1)])
            pr_3d = np.stack(out, axis=0)
        err = np.linalg.norm(gt_3d - pr_3d, axis=-1).mean(axis=-1)
        e3n[idx] = err / mesh_diameters[i]
        e2[idx] = np.linalg.norm(gt_2d - pr_2d, axis=-1).mean(axis=-1)
        e3[idx] = err
    return e3n, e2, e3


# --------------------------------------------------------------------------------------- renderer
def vertex_normals(verts, faces):
    """pytorch3d Meshes.verts_normals_packed: face normals (cross product, i.e. area-weighted) summed onto the
    face's vertices, normalised (eps 1e-6)."""
    v0, v1, v2 = verts[faces[:, 0]], verts[faces[:, 1]], verts[faces[:, 2]]
    n = np.zeros_like(verts, dtype=np.float64)
    np.add.at(n, faces[:, 1], np.cross(v2 - v1, v0 - v1))
    np.add.at(n, faces[:, 2], np.cross(v0 - v2, v1 - v2))
    np.add.at(n, faces[:, 0], np.cross(v1 - v0, v2 - v0))
    return n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-6)


def light_location(verts, R, t):
    """rendering.py:203-206 / 214-219 (seperate_lights=True)."""
    z = (verts @ R.T + t)[:, 2]
    return R @ np.array([0.0, 0.0, max(float(z.min()) - 400.0, 0.0)])


def render(verts, faces, colors, R, t, K, H, W, background=(0.5, 0.5, 0.5), ambient=0.5, diffuse=0.3,
           specular=0.2, shininess=64.0):
    """One object at one pose -> image (3, H, W), zbuf (H, W) (-1 = background), face index (H, W)."""
    verts = np.asarray(verts, np.float64); colors = np.asarray(colors, np.float64)
    R = np.asarray(R, np.float64); t = np.asarray(t, np.float64); K = np.asarray(K, np.float64)
    normals = vertex_normals(verts, faces)
    cam = verts @ R.T + t                                            # camera frame (OpenCV: z forward)
    u = K[0, 0] * cam[:, 0] / cam[:, 2] + K[0, 2]
    v = K[1, 1] * cam[:, 1] / cam[:, 2] + K[1, 2]
    su = (np.arange(W) + 0.5) * (W - 1) / W                           # where pixel j samples the OpenCV image
    sv = (np.arange(H) + 0.5) * (H - 1) / H
    zbuf = np.full((H, W), np.inf)
    fidx = np.full((H, W), -1, np.int64)
    bary = np.zeros((H, W, 3))
    for f, (a, b, c) in enumerate(faces):
        if cam[a, 2] <= 1e-6 or cam[b, 2] <= 1e-6 or cam[c, 2] <= 1e-6:
            continue
        area = (u[b] - u[a]) * (v[c] - v[a]) - (v[b] - v[a]) * (u[c] - u[a])
        if abs(area) < 1e-12:
            continue
        j0, j1 = np.searchsorted(su, min(u[a], u[b], u[c])), np.searchsorted(su, max(u[a], u[b], u[c]), "right")
        i0, i1 = np.searchsorted(sv, min(v[a], v[b], v[c])), np.searchsorted(sv, max(v[a], v[b], v[c]), "right")
        if j0 >= j1 or i0 >= i1:
            continue
        X, Y = np.meshgrid(su[j0:j1], sv[i0:i1])
        w0 = ((u[b] - X) * (v[c] - Y) - (v[b] - Y) * (u[c] - X)) / area
        w1 = ((u[c] - X) * (v[a] - Y) - (v[c] - Y) * (u[a] - X)) / area
        w2 = 1.0 - w0 - w1
        inside = (w0 >= 0) & (w1 >= 0) & (w2 >= 0)
        iz = w0 / cam[a, 2] + w1 / cam[b, 2] + w2 / cam[c, 2]
        z = 1.0 / np.where(iz > 0, iz, 1.0)
        zb = zbuf[i0:i1, j0:j1]
        win = inside & (z < zb)
        if not win.any():
            continue
        zb[win] = z[win]
        fidx[i0:i1, j0:j1][win] = f
        pb = np.stack([w0 / cam[a, 2], w1 / cam[b, 2], w2 / cam[c, 2]], -1) * z[..., None]   # perspective-correct
        bary[i0:i1, j0:j1][win] = pb[win]
    img = np.empty((3, H, W))
    img[:] = np.asarray(background, np.float64)[:, None, None]
    hit = fidx >= 0
    fa = faces[fidx[hit]]
    bw = bary[hit]
    pos = (verts[fa] * bw[..., None]).sum(1)
    nrm = (normals[fa] * bw[..., None]).sum(1)
    tex = (colors[fa] * bw[..., None]).sum(1)
    nrm = nrm / np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-6)
    loc = light_location(verts, R, t)
    ldir = loc[None] - pos
    ldir = ldir / np.maximum(np.linalg.norm(ldir, axis=1, keepdims=True), 1e-6)
    cosang = (nrm * ldir).sum(1)
    dif = diffuse * np.maximum(cosang, 0.0)
    center = -R.T @ t                                                 # camera centre in the object frame
    vdir = center[None] - pos
    vdir = vdir / np.maximum(np.linalg.norm(vdir, axis=1, keepdims=True), 1e-6)
    refl = -ldir + 2.0 * cosang[:, None] * nrm
    spec = specular * (cosang > 0) * np.maximum((vdir * refl).sum(1), 0.0) ** shininess
    col = (ambient + dif)[:, None] * tex + spec[:, None]
    img[:, hit] = col.T
    zout = np.where(hit, zbuf, -1.0)
    return img, zout, fidx


def icosphere(subdiv=2, radius=60.0, squash=(1.0, 0.8, 0.6)):
    """A closed test mesh with vertex colours (no mesh files ship with the reference)."""
    g = (1 + 5 ** 0.5) / 2
    v = [(-1, g, 0), (1, g, 0), (-1, -g, 0), (1, -g, 0), (0, -1, g), (0, 1, g), (0, -1, -g), (0, 1, -g),
         (g, 0, -1), (g, 0, 1), (-g, 0, -1), (-g, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
         (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
         (8, 6, 7), (9, 8, 1)]
    v = [np.array(p, np.float64) / np.linalg.norm(p) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []
        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    verts = np.array(v) * radius * np.array(squash)
    faces = np.array(f, np.int32)
    colors = 0.5 + 0.5 * np.array(v)                                   # smooth, position-dependent colours
    return verts.astype(np.float32), faces, colors.astype(np.float32)
