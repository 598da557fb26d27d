"""GPU tier, f4: the renderer call and the ADD / ADD-S evaluator (csrc/render.cu) through the C ABI against
oracle/render_oracle.py and the golden vectors of the unmodified reference evaluator (tests/golden/add_eval.npz)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import render_oracle as RO
from scflow_b200 import render as RD

DEV = "cuda"


def _meshes():
    return {0: RO.icosphere(2, 60.0, (1.0, 0.8, 0.6)), 1: RO.icosphere(3, 45.0, (0.7, 1.0, 1.0)),
            2: RO.icosphere(1, 80.0, (1.0, 1.0, 0.5))}


def _pose(g, n):
    q, _ = np.linalg.qr(g.normal(size=(n, 3, 3)))
    q[np.linalg.det(q) < 0, :, 0] *= -1
    t = np.stack([g.uniform(-40, 40, n), g.uniform(-40, 40, n), g.uniform(650, 950, n)], -1)
    return q.astype(np.float32), t.astype(np.float32)


def test_add_errors_equal_reference_golden(golden_dir):
    """ADD, ADD-S (class 1) and the 2-D projection error against the outputs of the UNMODIFIED ADD.eval_pose_error."""
    g = np.load(os.path.join(golden_dir, "add_eval.npz"))
    ms = RD.MeshSet({i: (g[f"verts{i}"], np.zeros((1, 3), np.int32)) for i in range(3)}, DEV)
    e3n, e2, e3 = RD.add_errors(ms, g["gt_r"], g["gt_t"], g["pred_r"], g["pred_t"], g["labels"], g["k"],
                                [bool(s) for s in g["sym"]], list(g["diam"]))
    for got, want in ((e3, g["e3"]), (e2, g["e2"]), (e3n, g["e3n"])):
        assert float((got.cpu() - torch.from_numpy(want)).abs().max() / np.abs(want).max()) < 2e-5


def test_add_errors_identity_and_translation():
    ms = RD.MeshSet(_meshes(), DEV)
    g = np.random.default_rng(3)
    R, t = _pose(g, 5)
    K = np.tile(np.array([[1066.8, 0, 128.0], [0, 1067.5, 128.0], [0, 0, 1.0]], np.float32), (5, 1, 1))
    lab = np.array([0, 1, 2, 1, 0])
    _, e2, e3 = RD.add_errors(ms, R, t, R, t, lab, K, [False, True, False])
    assert float(e3.abs().max()) == 0 and float(e2.abs().max()) == 0
    shift = t + np.array([0, 0, 10.0], np.float32)
    _, e2, e3 = RD.add_errors(ms, R, t, R, shift, lab, K, [False, False, False])
    assert float((e3 - 10.0).abs().max()) < 1e-4               # a pure translation moves every vertex by its length


@pytest.mark.parametrize("size", [(64, 64), (128, 96)])
def test_render_vs_oracle(size):
    """Image, z-buffer and mask of a batch against the numpy restatement: the coverage may differ on samples that
    fall exactly on an edge in one arithmetic (fp32 here, fp64 there); everywhere else depth to 1e-4 relative and
    colour to 2e-3."""
    H, W = size
    meshes = _meshes()
    ms = RD.MeshSet(meshes, DEV)
    g = np.random.default_rng(11)
    n = 4
    R, t = _pose(g, n)
    lab = np.array([0, 1, 2, 1])
    K = np.tile(np.array([[900.0, 0, W / 2], [0, 900.0, H / 2], [0, 0, 1.0]], np.float32), (n, 1, 1)) * 0.3
    K[:, 2, 2] = 1.0
    K[:, 0, 2], K[:, 1, 2] = W / 2 + g.uniform(-4, 4, n), H / 2 + g.uniform(-4, 4, n)
    out = RD.Renderer(ms, (H, W))(torch.from_numpy(R).to(DEV), torch.from_numpy(t).to(DEV), torch.from_numpy(K).to(DEV),
                                  torch.from_numpy(lab).to(DEV))
    torch.cuda.synchronize()
    for i in range(n):
        v, f, c = meshes[int(lab[i])]
        img, z, _ = RO.render(v, f, c, R[i], t[i], K[i], H, W)
        gz, gi, gm = out["zbuf"][i].cpu().numpy(), out["images"][i].cpu().numpy(), out["masks"][i].cpu().numpy()
        hit_w, hit_g = z > 0, gz > 0
        assert hit_w.sum() > 0.02 * H * W                        # the object is in view
        assert np.array_equal(hit_g, gm > 0)
        assert (hit_w != hit_g).sum() <= max(2, 0.01 * hit_w.sum())
        both = hit_w & hit_g
        assert np.abs(gz[both] - z[both]).max() / z[both].max() < 1e-4
        assert np.abs(gi[:, both] - img[:, both]).max() < 2e-3
        assert np.allclose(gi[:, ~hit_g], 0.5) and np.all(gz[~hit_g] == -1)
