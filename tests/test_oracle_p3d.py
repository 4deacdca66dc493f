"""Cross-checks standing in for the missing pytorch3d pin (parity unpinned, see oracle/p3d_oracle.py)."""
import os
import sys

import numpy as np
import pytest
import torch
from scipy.spatial.transform import Rotation

from oracle import closures, p3d_oracle as p3d


def test_axis_angle_to_matrix_vs_scipy():
    rng = np.random.default_rng(0)
    aa = rng.normal(0, 1.0, size=(64, 3))
    aa[0] = 0.0
    aa[1] = [0, 0, 1e-7]
    got = p3d.axis_angle_to_matrix(torch.from_numpy(aa)).numpy()
    ref = Rotation.from_rotvec(aa).as_matrix()
    assert np.abs(got - ref).max() < 1e-12


def test_yaw_gradient_at_zero_is_finite():
    z = torch.zeros(4, 1, 1, requires_grad=True)
    R = p3d.compute_root_orient_z(z)
    R[..., 1, 0].sum().backward()
    assert torch.allclose(z.grad, torch.ones_like(z.grad))   # d sin(z)/dz at 0


def test_6d_roundtrip_is_gram_schmidt():
    rng = np.random.default_rng(1)
    m = torch.from_numpy(rng.normal(size=(32, 3, 3)))
    r = p3d.normalize_rot(m)
    eye = torch.eye(3, dtype=r.dtype).expand(32, 3, 3)
    assert torch.allclose(r @ r.transpose(-1, -2), eye, atol=1e-12)
    assert torch.allclose(torch.linalg.det(r), torch.ones(32, dtype=r.dtype), atol=1e-12)
    # first row direction preserved, a proper rotation is a fixed point
    assert torch.allclose(r[:, 0], m[:, 0] / m[:, 0].norm(dim=-1, keepdim=True))
    q = torch.from_numpy(Rotation.random(8, random_state=3).as_matrix())
    assert torch.allclose(p3d.normalize_rot(q), q, atol=1e-12)


def test_knn_first_min_and_reference_closest_point():
    rng = np.random.default_rng(2)
    pts = rng.normal(size=(1, 500, 3)).astype(np.float32)
    pts[0, 300] = pts[0, 20]      # exact duplicate -> tie, lowest index must win
    q = rng.normal(size=(1, 40, 3)).astype(np.float32)
    q[0, 0] = pts[0, 20] + np.float32(1e-3)
    d, idx = p3d.knn_k1(torch.from_numpy(q), torch.from_numpy(pts))
    assert int(idx[0, 0]) == 20
    brute = ((q[0][:, None, :].astype(np.float64) - pts[0][None].astype(np.float64)) ** 2).sum(-1)
    assert np.array_equal(idx[0].numpy(), brute.argmin(-1)) or \
        np.allclose(np.take_along_axis(brute, idx[0].numpy()[:, None], 1)[:, 0], brute.min(-1), rtol=1e-6)
    # the reference's own numpy nearest-vertex helper, restated inline from
    # src/video_mocap/utils/points.py:19-42: norm of differences, argmin over the cloud axis
    dist = np.linalg.norm(q[0][None, :, :] - pts[0][:, None, :], axis=-1)
    ref_idx = np.argmin(dist, axis=0)
    same = ref_idx == idx[0].numpy()
    # sqrt is monotone; the only permitted differences are fp ties created by the sqrt rounding
    assert same.mean() > 0.95
    for m in np.where(~same)[0]:
        assert abs(dist[ref_idx[m], m] - dist[idx[0, m], m]) <= 1e-6 * dist[ref_idx[m], m]


def test_knn_backward_matches_dense_autograd():
    rng = np.random.default_rng(3)
    x = torch.from_numpy(rng.normal(size=(3, 7, 3)).astype(np.float32)).requires_grad_(True)
    y = torch.from_numpy(rng.normal(size=(3, 50, 3)).astype(np.float32)).requires_grad_(True)
    w = torch.tensor([[1, 0, 1, 1, 0, 1, 1]] * 3, dtype=torch.bool)
    loss, _ = p3d.weighted_chamfer_distance(x, y, w)
    loss.backward()
    gx, gy = x.grad.clone(), y.grad.clone()
    x.grad = None
    y.grad = None
    d = ((x[:, :, None, :] - y[:, None, :, :]) ** 2).sum(-1).min(-1)[0]
    ref = (d * w).sum() / w.sum()
    ref.backward()
    assert torch.allclose(loss, ref, rtol=1e-6)
    assert torch.allclose(gx, x.grad, rtol=1e-5, atol=1e-7)
    assert torch.allclose(gy, y.grad, rtol=1e-5, atol=1e-7)


def test_chamfer_smoke_seed_value(golden):
    """The one deterministic input the reference ships: losses/chamfer_distance.py:24-36."""
    x = torch.arange(0, 16 * 7 * 3, dtype=torch.float32).reshape(16, 7, 3)
    y = torch.arange(0, 16 * 19 * 3, dtype=torch.float32).reshape(16, 19, 3)
    w = torch.ones(16, 7)
    w[:, ::2] = 0
    val, _ = p3d.weighted_chamfer_distance(x, y, w)
    # closed form: x[b,p] = 21b+3p+(0,1,2), y[b,q] = 57b+3q+(0,1,2): nearest q minimises |21b+3p-57b-3q|
    b = np.arange(16)[:, None]
    p = np.arange(7)[None, :]
    q = np.arange(19)[None, None, :]
    delta = (21 * b + 3 * p)[..., None] - (57 * b[..., None] + 3 * q)
    d = 3.0 * (delta.astype(np.float64) ** 2).min(-1)
    wn = w.numpy()
    ref = (d * wn).sum() / wn.sum()
    assert abs(val.item() - ref) <= 1e-6 * ref
    assert abs(val.item() - float(golden("closure_top4.npz")["smoke_weighted_chamfer"])) <= 1e-6 * ref


def test_all_dropped_markers_give_zero_loss():
    x = torch.zeros(2, 5, 3, requires_grad=True)
    y = torch.randn(2, 30, 3)
    loss, _ = p3d.weighted_chamfer_distance(x, y, closures.get_marker_mask(x))
    assert loss.item() == 0.0
    loss.backward()
    assert torch.count_nonzero(x.grad) == 0


def test_one_hot_gather_equals_index(get_model):
    rng = np.random.default_rng(4)
    verts = torch.from_numpy(rng.normal(size=(3, 200, 3)).astype(np.float32))
    idx = torch.tensor([5, 199, 0, 17])
    oh = torch.zeros(4, 200)
    oh[torch.arange(4), idx] = 1.0
    assert torch.equal(closures.one_hot_gather_dense(verts, oh), verts[:, idx])


def test_closure_golden_reproduced(get_model, golden):
    """oracle drift check: closures reproduce the committed loss/grad/idx fixtures."""
    from oracle.smpl_oracle import OracleSMPL
    g = golden("closure_top4.npz")
    model = get_model("top4")
    smpl = OracleSMPL.from_model(model)
    T = lambda k: torch.from_numpy(g[k]).clone()
    markers = T("markers")
    pose, betas, trans, z = (T("ch_pose").requires_grad_(True), T("ch_betas").requires_grad_(True),
                             T("ch_trans").requires_grad_(True), T("ch_z").requires_grad_(True))
    losses = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0, "trans_vel": 1.0}
    loss, idx = closures.chamfer_stage_loss(smpl, markers, pose, T("ch_o_pose"), betas, T("ch_o_betas"),
                                            T("ch_root"), trans, z, losses, return_idx=True)
    loss.backward()
    assert np.array_equal(idx.numpy(), g["ch_idx"])
    assert abs(loss.item() - float(g["ch_loss"])) <= 1e-5 * abs(float(g["ch_loss"]))
    for got, key in ((pose.grad, "ch_g_pose"), (betas.grad, "ch_g_betas"), (trans.grad, "ch_g_trans"),
                     (z.grad, "ch_g_z")):
        assert np.abs(got.numpy() - g[key]).max() <= 1e-4 * np.abs(g[key]).max(), key


def test_so3_relative_angle_vs_scipy_and_extrapolation():
    """so3_relative_angle (pytorch3d defaults): scipy's geodesic angle inside the bound, the first-order extension of
    arccos at cos = 1 - 1e-4 above it (finite gradient for identical rotations)."""
    r1 = Rotation.random(64, random_state=5)
    r2 = Rotation.random(64, random_state=6)
    R1, R2 = torch.from_numpy(r1.as_matrix()), torch.from_numpy(r2.as_matrix())
    got = p3d.so3_relative_angle(R1, R2).numpy()
    ref = (r1 * r2.inv()).magnitude()
    inside = np.abs(np.cos(ref)) < 1 - 1e-4
    assert inside.sum() > 50 and np.abs(got[inside] - ref[inside]).max() < 1e-9
    same = R1.clone().requires_grad_(True)
    ang = p3d.so3_relative_angle(same, R1)
    b = 1.0 - 1e-4
    expect = np.arccos(b) + (1.0 - b) * (-1.0 / np.sqrt(1.0 - b * b))
    assert torch.allclose(ang, torch.full_like(ang, expect), atol=1e-9)
    ang.sum().backward()
    assert torch.isfinite(same.grad).all()
    with pytest.raises(ValueError):
        p3d.so3_relative_angle(2.0 * R1, R1)
