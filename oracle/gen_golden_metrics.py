"""Golden vectors for the evaluation metrics, produced by the REFERENCE's own functions in the build container.
    python oracle/gen_golden_metrics.py    (needs /root/reference; writes tests/golden/metrics.npz)
src/video_mocap/evaluation/metrics.py does not import here (igl, pytorch3d, seaborn), so the source segments of the
pure-torch functions are compiled on their own, unmodified."""
import ast
import os
import sys
from typing import List

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/src/video_mocap/evaluation/metrics.py"
WANT = ["compute_MPJPE", "compute_MPJPE_joints", "compute_MPJVE", "compute_MPJVE_joints", "compute_PA_MPJPE",
        "compute_PA_MPJPE_joints", "compute_PA_MPJVE", "compute_PA_MPJVE_joints", "compute_V2V",
        "compute_similarity_transform"]
tree = ast.parse(open(SRC).read())
fns = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in WANT]
ns = {"torch": torch, "np": np, "List": List}
exec(compile(ast.Module(body=fns, type_ignores=[]), SRC, "exec"), ns)

rng = np.random.default_rng(7)
F, J, V = 33, 45, 500
gt = rng.normal(0, 0.4, size=(F, J, 3)).astype(np.float32)
gt += np.cumsum(rng.normal(0, 0.01, size=(F, 1, 3)), axis=0).astype(np.float32)
# prediction = a rotated, scaled, shifted, noisy copy: the Procrustes step has something to undo
ang = 0.7
R = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], dtype=np.float32)
pred = (1.1 * gt @ R.T + np.array([0.3, -0.2, 0.1], dtype=np.float32) + rng.normal(0, 0.02, size=gt.shape)).astype(np.float32)
pred[5] = gt[5] * np.array([1, 1, -1], dtype=np.float32)          # a mirrored frame: the det(R) = +1 fix matters
gv = rng.normal(0, 0.5, size=(F, V, 3)).astype(np.float32)
pv = (gv + rng.normal(0, 0.01, size=gv.shape)).astype(np.float32)
ids = [1, 2, 4, 5, 7, 8, 20, 21]
T = torch.from_numpy
out = {"gt": gt, "pred": pred, "gv": gv, "pv": pv, "ids": np.array(ids), "freq": np.float32(120.0)}
out["MPJPE"] = ns["compute_MPJPE"](T(pred), T(gt)).numpy()
out["MPJPE_joints"] = ns["compute_MPJPE_joints"](T(pred), T(gt), ids).numpy()
out["MPJVE"] = ns["compute_MPJVE"](T(pred), T(gt), 120.0).numpy()
out["MPJVE_joints"] = ns["compute_MPJVE_joints"](T(pred), T(gt), 120.0, ids).numpy()
out["PA_MPJPE"] = ns["compute_PA_MPJPE"](T(pred), T(gt)).numpy()
out["PA_MPJPE_joints"] = ns["compute_PA_MPJPE_joints"](T(pred), T(gt), ids).numpy()
out["PA_MPJVE"] = ns["compute_PA_MPJVE"](T(pred), T(gt), 120.0).numpy()
out["PA_MPJVE_joints"] = ns["compute_PA_MPJVE_joints"](T(pred), T(gt), 120.0, ids).numpy()
out["V2V"] = ns["compute_V2V"](T(pv), T(gv)).numpy()
out["similarity"] = ns["compute_similarity_transform"](T(pred), T(gt)).numpy()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "metrics.npz"), **out)
print({k: (float(v) if np.ndim(v) == 0 else v.shape) for k, v in out.items()})
