"""Evaluation metrics against golden vectors produced by the reference's own functions
(oracle/gen_golden_metrics.py executes src/video_mocap/evaluation/metrics.py's pure-torch functions unmodified), and
the marker-to-surface distance against the oracle's independent float64 restatement."""
import numpy as np
import pytest
import torch

from conftest import rel_err


def test_mesh_oracle_known_answers():
    """CPU: the oracle's point-to-triangle distance on cases with analytic answers (face, edge, vertex regions)."""
    from oracle.mesh_oracle import point_mesh_distance
    v = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], dtype=np.float64)
    f = np.array([[0, 1, 2]])
    p = np.array([[[0.25, 0.25, 2.0], [2, 0, 0], [-1, -1, 0], [0.5, 0.5, 0], [1, 1, 0], [0.5, -3, 4]]], dtype=np.float64)
    ref = np.array([[2.0, 1.0, np.sqrt(2.0), 0.0, np.sqrt(0.5), 5.0]])
    assert np.allclose(point_mesh_distance(p, v, f), ref, atol=1e-12)


@pytest.mark.gpu
def test_metrics_match_reference_golden(golden):
    from uuo_mocap_b200 import metrics
    g = golden("metrics.npz")
    C = lambda k: torch.from_numpy(g[k]).cuda()
    pred, gt, pv, gv = C("pred"), C("gt"), C("pv"), C("gv")
    ids, freq = g["ids"].tolist(), float(g["freq"])
    got = {
        "MPJPE": metrics.compute_MPJPE(pred, gt), "MPJPE_joints": metrics.compute_MPJPE_joints(pred, gt, ids),
        "MPJVE": metrics.compute_MPJVE(pred, gt, freq), "MPJVE_joints": metrics.compute_MPJVE_joints(pred, gt, freq, ids),
        "PA_MPJPE": metrics.compute_PA_MPJPE(pred, gt), "PA_MPJPE_joints": metrics.compute_PA_MPJPE_joints(pred, gt, ids),
        "PA_MPJVE": metrics.compute_PA_MPJVE(pred, gt, freq),
        "PA_MPJVE_joints": metrics.compute_PA_MPJVE_joints(pred, gt, freq, ids), "V2V": metrics.compute_V2V(pv, gv),
    }
    for k, v in got.items():
        assert v.dim() == 0 and v.is_cuda
        tol = 2e-5 if k.startswith("PA_") else 2e-6      # the PA- variants go through a 3x3 SVD (fp32 in the reference)
        assert abs(float(v) - float(g[k])) <= tol * abs(float(g[k])), (k, float(v), float(g[k]))
    sim = metrics.compute_similarity_transform(pred, gt)
    assert rel_err(sim.cpu().numpy(), g["similarity"]) < 1e-5
    with pytest.raises(ValueError):
        metrics.compute_MPJPE(pred, gt[:, :5])
    with pytest.raises(IndexError):
        metrics.compute_MPJPE_joints(pred, gt, [0, 99])
    with pytest.raises(Exception):
        metrics.compute_V2V(pv.cpu(), gv.cpu())


@pytest.mark.gpu
@pytest.mark.parametrize("F,M,V,NF", [(1, 1, 3, 1), (3, 41, 300, 596), (2, 45, 6890, 13776), (2, 9, 50, 7)])
def test_point_mesh_distance_matches_oracle(F, M, V, NF, get_model):
    from uuo_mocap_b200 import metrics
    from oracle.mesh_oracle import point_mesh_distance
    rng = np.random.default_rng(F * 10 + M)
    verts = rng.normal(0, 0.5, size=(F, V, 3)).astype(np.float32)
    faces = np.stack([rng.permutation(V)[:3] for _ in range(NF)]).astype(np.int64) if V < 100 else \
        np.stack([np.arange(NF) % V, (np.arange(NF) * 7 + 1) % V, (np.arange(NF) * 13 + 5) % V], axis=1)
    faces = faces[(faces[:, 0] != faces[:, 1]) & (faces[:, 1] != faces[:, 2]) & (faces[:, 0] != faces[:, 2])]
    pts = rng.normal(0, 0.6, size=(F, M, 3)).astype(np.float32)
    pts[:, 0] = verts[:, faces[0, 0]]                         # a point ON the mesh: distance 0
    ref = point_mesh_distance(pts, verts, faces)
    got = metrics.point_mesh_distance(torch.from_numpy(pts).cuda(), torch.from_numpy(verts).cuda(),
                                      torch.from_numpy(faces).cuda())
    assert got.shape == (F, M)
    assert np.abs(got.cpu().numpy() - ref).max() < 2e-5
    mean = metrics.compute_marker_to_surface_distance(torch.from_numpy(verts).cuda(),
                                                      torch.from_numpy(faces).cuda().unsqueeze(0).expand(F, -1, -1),
                                                      torch.from_numpy(pts).cuda())
    assert abs(float(mean) - ref.mean()) < 2e-5
