"""Root stage (optim_root, optimization.py:21-144): the closure on the CUDA path
against the oracle's restatement in all three yaw modes with every working term on, and an end-to-end fit that must
take out a yaw + translation error in place."""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.fixture(scope="module")
def ctx(get_model):
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    from oracle.smpl_oracle import OracleSMPL
    m = get_model("top4")
    return m, SMPL(_as_struct(m), device="cuda:0"), OracleSMPL.from_model(m)


def _case(model, F, seed, stride=8):
    """Every ``stride``-th frame of a longer smooth capture: frame-to-frame root rotations of ~0.1 rad, so the relative
    angles sit inside arccos' well-conditioned range."""
    from uuo_mocap_b200 import synthetic
    cap = synthetic.make_capture(model, F * stride, num_markers=41, seed=seed, noise_std=0.001, permute=False)
    sl = slice(0, F * stride, stride)
    rot = synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa[sl].astype(np.float64)).astype(np.float32)
    labels = np.argmax(model.weights, axis=-1)[cap.marker_vertex_ids]
    return cap.markers[sl].astype(np.float32), rot, cap.gt_trans[sl].astype(np.float32), cap.gt_betas.astype(np.float32), \
        np.repeat(labels[None], F, 0)


@pytest.mark.parametrize("mode", ["constrained", "yaw_lock", "free"])
def test_root_closure_matches_oracle(ctx, mode):
    """loss 1e-5, d trans / d z_angle / d betas 1e-4 vs oracle.closures.root_stage_loss."""
    from uuo_mocap_b200 import root
    from oracle import closures
    model, smpl, orc = ctx
    F = 12
    mk, rot, tr0, b0, lab = _case(model, F, seed=31)
    mk[2, 5] = 0.0                                        # a dropped marker: masked out of full_chamfer
    stage = {"constrained_rotation": mode == "constrained", "yaw_lock": mode == "yaw_lock", "single_directional": False,
             "losses": {"part_chamfer": 3.0, "full_chamfer": 10.0, "root_orient_vel": 2.0, "trans_vel": 1.0}}
    rng = np.random.default_rng(7)
    if mode == "constrained":
        z0 = np.full((1, 1, 1), 0.25, np.float32)
    elif mode == "yaw_lock":
        z0 = rng.uniform(-0.3, 0.3, size=(F, 1, 1)).astype(np.float32)
    else:
        z0 = (rot[:, :1] + rng.normal(0, 0.05, size=(F, 1, 3, 3))).astype(np.float32)   # not orthonormal on purpose
    out = []
    for dev in ("cpu", "cuda:0"):
        z = _T(z0).to(dev).requires_grad_(True)
        tr = (_T(tr0) + 0.02).to(dev).requires_grad_(True)
        betas = (_T(b0) * 0.7).to(dev).requires_grad_(True)
        r = _T(rot).to(dev)
        if dev == "cpu":
            loss = closures.root_stage_loss(orc, _T(mk), r[:, 1:], betas, r[:, :1], tr, z, _T(lab), stage)
        else:
            loss = root.root_closure_loss(smpl, _T(mk).cuda(), r[:, 1:], betas, r[:, :1], tr, z, _T(lab).cuda(), stage)
        loss.backward()
        out.append((float(loss), tr.grad.cpu().numpy(), z.grad.cpu().numpy(), betas.grad.cpu().numpy()))
    assert abs(out[1][0] - out[0][0]) <= 1e-5 * abs(out[0][0])
    for k in (1, 2, 3):
        assert rel_err(out[1][k], out[0][k]) < 1e-4, (mode, k)


def test_so3_relative_angle_matches_oracle():
    from uuo_mocap_b200 import root
    from oracle import p3d_oracle as p3d
    from scipy.spatial.transform import Rotation
    R1 = torch.from_numpy(Rotation.random(256, random_state=1).as_matrix()).float()
    R2 = torch.from_numpy(Rotation.random(256, random_state=2).as_matrix()).float()
    R2[:8] = R1[:8]                                       # identical pairs: the extrapolated branch
    a = R1.clone().cuda().requires_grad_(True)
    got = root.so3_relative_angle(a, R2.cuda())
    b = R1.clone().requires_grad_(True)
    ref = p3d.so3_relative_angle(b, R2)
    assert torch.allclose(got.cpu(), ref, atol=2e-5)
    got.sum().backward()
    ref.sum().backward()
    assert torch.isfinite(a.grad).all() and torch.allclose(a.grad.cpu(), b.grad, atol=1e-3, rtol=1e-3)
    with pytest.raises(ValueError):
        root.so3_relative_angle(2.0 * a.detach(), R2.cuda())


@pytest.mark.parametrize("optimizer", ["lbfgs", "adam"])
def test_optim_root_removes_yaw_and_offset(ctx, optimizer):
    """One yaw for the take (constrained_rotation) + per-frame translation: starting 0.3 rad / 4 cm off, the fit must
    bring the loss down and write root_orient / trans / betas back in place (optimization.py:137-144)."""
    from uuo_mocap_b200 import root
    from uuo_mocap_b200.fit import rot_z
    model, smpl, orc = ctx
    F = 16
    mk, rot, tr0, b0, lab = _case(model, F, seed=33, stride=4)
    yaw = 0.3
    root0 = (rot_z(torch.full((F, 1, 1), yaw)) @ _T(rot[:, :1])).contiguous()
    root_orient = root0.clone()
    trans = (_T(tr0) + torch.tensor([0.04, -0.03, 0.0])).contiguous()
    betas = _T(b0).clone()
    cfg = {"optimizer": {"name": optimizer, "tolerance_grad": 1e-9, "tolerance_change": 1e-12},
           "stages": {"root": {"constrained_rotation": True, "yaw_lock": True, "num_iters": 60 if optimizer == "lbfgs" else 150,
                               "lr": 1.0, "lr_adam": 0.02, "single_directional": True,
                               "losses": {"full_chamfer": 10.0, "trans_vel": 1.0}}}}
    hist = []

    def iter_fn(stage, iteration, pose_body, betas, trans, root_orient):
        assert stage == "root" and root_orient.shape == (F, 1, 3, 3) and trans.shape == (F, 3)
        hist.append(iteration)
    losses = []
    import builtins
    real_print = builtins.print
    builtins.print = lambda *a, **k: losses.append(a[2]) if a and a[0] == "Root" else real_print(*a, **k)
    try:
        res = root.optim_root(_T(mk), _T(rot[:, 1:]), betas, root_orient, trans, _T(lab), smpl, cfg, verbose=True, iter_fn=iter_fn)
    finally:
        builtins.print = real_print
    assert res["iterations"] == len(hist) == len(losses) and len(losses) > 5
    assert min(losses) < 0.2 * losses[0], (losses[0], min(losses))
    z = float(res["z_angle"].reshape(-1)[0])
    assert abs(z + yaw) < 0.1, z
    assert torch.allclose(root_orient, (rot_z(torch.full((F, 1, 1), z)) @ root0), atol=1e-5)
    assert not torch.equal(trans, _T(tr0) + torch.tensor([0.04, -0.03, 0.0]))
    # zero iterations (the shipped yaml): nothing moves
    cfg["stages"]["root"]["num_iters"] = 0
    r2, t2, b2 = root_orient.clone(), trans.clone(), betas.clone()
    root.optim_root(_T(mk), _T(rot[:, 1:]), b2, r2, t2, _T(lab), smpl, cfg)
    assert torch.allclose(r2, root_orient, atol=1e-6) and torch.equal(t2, trans) and torch.equal(b2, betas)


def test_optim_root_rejects_broken_terms(ctx):
    from uuo_mocap_b200 import root
    model, smpl, orc = ctx
    mk, rot, tr0, b0, lab = _case(model, 4, seed=34, stride=2)
    cfg = {"optimizer": {"name": "lbfgs"}, "stages": {"root": {"constrained_rotation": True, "yaw_lock": True, "num_iters": 2,
                                                               "single_directional": True, "losses": {"reg_betas": 1.0}}}}
    with pytest.raises(NotImplementedError):
        root.optim_root(_T(mk), _T(rot[:, 1:]), _T(b0), _T(rot[:, :1]).clone(), _T(tr0), _T(lab), smpl, cfg)
