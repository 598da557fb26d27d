"""Golden vectors of the ADD / ADD-S evaluator from the UNMODIFIED reference method
`ADD.eval_pose_error` (/root/reference/metrics/add.py:354-394) with `project_3d_point`
(/root/reference/datasets/pose.py:18-70).  Run in the dev container (where /root/reference is mounted):

    python oracle/gen_golden_add.py        # writes tests/golden/add_eval.npz

The two reference files are imported as they are; only their uninstalled third-party imports (mmengine,
terminaltables, trimesh, registry, the datasets package's other modules) are stubbed -- none of them is touched
by eval_pose_error.  Test infrastructure only."""
import importlib.util
import os
import sys
import types

import numpy as np

REF = os.environ.get("SCFLOW_REFERENCE", "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def reference_eval_pose_error():
    class _Reg:
        def register_module(self, *a, **k):
            return lambda c: c
    _mod("trimesh")
    _mod("terminaltables", AsciiTable=object)
    me = _mod("mmengine")
    _mod("mmengine.evaluator", BaseMetric=object)
    _mod("mmengine.logging", print_log=print)
    me.evaluator, me.logging = sys.modules["mmengine.evaluator"], sys.modules["mmengine.logging"]
    _mod("registry", METRICS=_Reg())
    ds = _mod("datasets", Compose=object)
    ds.__path__ = []
    _mod("datasets.utils", dumps_json=None)
    pose = _load("datasets.pose", "datasets/pose.py")               # the real project_3d_point
    ds.pose = pose
    add = _load("ref_metrics_add", "metrics/add.py")
    return add.ADD.eval_pose_error


def make_case(seed=0, n=12, nv=(200, 333, 64)):
    g = np.random.default_rng(seed)
    verts_list = [g.normal(size=(v, 3)) * np.array([60.0, 40.0, 25.0]) for v in nv]

    def rot(k):
        q, _ = np.linalg.qr(g.normal(size=(k, 3, 3)))
        q[np.linalg.det(q) < 0, :, 0] *= -1
        return q
    gt_r = rot(n)
    d = rot(n)
    pred_r = np.stack([r @ (np.eye(3) + 0.05 * (dd - dd.T)) for r, dd in zip(gt_r, g.normal(size=(n, 3, 3)))])
    pred_r = np.stack([np.linalg.qr(p)[0] * np.sign(np.diag(np.linalg.qr(p)[1])) for p in pred_r])
    gt_t = np.stack([g.uniform(-80, 80, n), g.uniform(-80, 80, n), g.uniform(600, 1100, n)], -1)[..., None]
    pred_t = gt_t + g.normal(size=(n, 3, 1)) * np.array([2.0, 2.0, 8.0])[None, :, None]
    labels = g.integers(0, len(nv), n)
    k = np.tile(np.array([[1066.778, 0, 312.9869], [0, 1067.487, 241.3109], [0, 0, 1.0]]), (n, 1, 1))
    k[:, 0, 2] += g.uniform(-50, 50, n)
    sym = {"cls_2": True}
    diam = [172.0, 269.5, 95.1]
    return dict(verts_list=verts_list, gt_t=gt_t, gt_r=gt_r, pred_t=pred_t, pred_r=pred_r, labels=labels, k=k,
                symmetry_types=sym, mesh_diameters=diam)


if __name__ == "__main__":
    fn = reference_eval_pose_error()
    c = make_case()
    # fp32-representable inputs (what the CUDA evaluator takes), evaluated by the reference in float64
    for key in ("gt_t", "gt_r", "pred_t", "pred_r", "k"):
        c[key] = c[key].astype(np.float32).astype(np.float64)
    c["verts_list"] = [v.astype(np.float32).astype(np.float64) for v in c["verts_list"]]
    e3n, e2, e3 = fn(None, c["verts_list"], c["gt_t"][..., 0], c["gt_r"], c["pred_t"][..., 0], c["pred_r"],
                     c["labels"], c["k"], c["symmetry_types"], c["mesh_diameters"])
    out = os.path.join(ROOT, "tests", "golden", "add_eval.npz")
    np.savez_compressed(out, **{f"verts{i}": v.astype(np.float32) for i, v in enumerate(c["verts_list"])},
                        gt_t=c["gt_t"].astype(np.float32), gt_r=c["gt_r"].astype(np.float32),
                        pred_t=c["pred_t"].astype(np.float32), pred_r=c["pred_r"].astype(np.float32),
                        labels=c["labels"], k=c["k"].astype(np.float32), sym=np.array([0, 1, 0]),
                        diam=np.array(c["mesh_diameters"]), e3n=e3n, e2=e2, e3=e3)
    print("wrote", out, e3[:4], e2[:4])
