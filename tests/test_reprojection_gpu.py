"""Reprojection stage (hmr_utils.py:14-425): camera recovery, set-up and the closure on
the CUDA path against oracle/reproj_oracle.py, and a short optim_reprojection run (output dict of the reference)."""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

CORR = np.array([[1, 0, 0], [0, 0, 1], [0, -1, 0]], np.float32)


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.fixture(scope="module")
def ctx(get_model):
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    from oracle.smpl_oracle import OracleSMPL
    m = get_model("top4")
    return m, SMPL(_as_struct(m), device="cuda:0"), OracleSMPL.from_model(m)


def _case(model, F=10, seed=41):
    """A capture seen by an HMR-like tracker: root in HMR axes, PHALP box parameters, two frames without detection."""
    from uuo_mocap_b200 import synthetic
    cap = synthetic.make_capture(model, F, num_markers=41, seed=seed, noise_std=0.001, permute=False)
    rot = synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)
    rng = np.random.default_rng(seed)
    d = {
        "markers": cap.markers.astype(np.float32), "pose_body": rot[:, 1:], "hmr_root": CORR.T @ rot[:, :1],
        "betas": cap.gt_betas.astype(np.float32), "hmr_betas": (cap.gt_betas + rng.normal(0, 0.1, size=(F, 10))).astype(np.float32),
        "trans": (cap.gt_trans + rng.normal(0, 0.05, size=(F, 3))).astype(np.float32),
        "pred_cam": np.stack([1.0 + rng.normal(0, 0.002, F), rng.normal(0, 0.05, F), rng.normal(0, 0.05, F)], 1).astype(np.float32),
        "center": (np.array([320.0, 240.0]) + rng.normal(0, 20, size=(F, 2))).astype(np.float32),
        "size": np.tile(np.array([[480.0, 640.0]], np.float32), (F, 1)),
        "scale": np.full((F, 1), 0.7, np.float32),
    }
    d["pred_cam"][3] = np.nan
    d["pred_cam"][7] = np.nan
    return d


def test_camera_recovery_and_setup_match_oracle(ctx):
    from uuo_mocap_b200 import reprojection as rp
    from oracle import reproj_oracle as ro
    model, smpl, orc = ctx
    d = _case(model)
    C = lambda k: _T(d[k]).cuda()
    a = ro.get_3d_parameters(orc, _T(d["hmr_betas"]), _T(d["pose_body"]), _T(d["hmr_root"]), _T(d["pred_cam"]), _T(d["center"]),
                             _T(d["size"]), _T(d["scale"]))
    b = rp.get_3d_parameters(smpl, C("hmr_betas"), C("pose_body"), C("hmr_root"), C("pred_cam"), C("center"), C("size"), C("scale"))
    ok = ~np.isnan(d["pred_cam"][:, 0])
    for k in ("pred_cam_t", "pred_keypoints_2d_smpl", "pred_joints", "focal_length", "camera_center"):
        x, y = b[k].cpu().numpy(), a[k].numpy()
        assert np.array_equal(np.isnan(x), np.isnan(y)), k
        assert rel_err(x[ok], y[ok]) < 1e-5, k
    sa = ro.reprojection_setup(orc, _T(d["markers"]), _T(d["pose_body"]), _T(d["hmr_betas"]), _T(d["hmr_root"]), _T(d["trans"]),
                               _T(d["pred_cam"]), _T(d["center"]), _T(d["size"]), _T(d["scale"]))
    sb = rp.reprojection_setup(smpl, C("markers"), C("pose_body"), C("hmr_betas"), C("hmr_root"), C("trans"), C("pred_cam"),
                               C("center"), C("size"), C("scale"))
    assert sb["mask"].cpu().tolist() == sa["mask"].tolist() and sa["mask"].sum() == 8
    for k in ("target", "smpl_trans0", "cam_single0", "focal_length"):
        assert rel_err(sb[k].cpu().numpy(), sa[k].numpy()) < 1e-5, k
    # projection helper with a rotation and a principal point
    rng = np.random.default_rng(1)
    pts, t = rng.normal(size=(4, 9, 3)).astype(np.float32), np.tile(np.array([[0, 0, 6.0]], np.float32), (4, 1))
    f, c = rng.uniform(1, 3, size=(4, 2)).astype(np.float32), rng.normal(size=(4, 2)).astype(np.float32)
    from scipy.spatial.transform import Rotation
    R = Rotation.random(4, random_state=2).as_matrix().astype(np.float32) * 0.2 + np.eye(3, dtype=np.float32) * 0.8
    x = rp.perspective_projection(_T(pts).cuda(), _T(t).cuda(), _T(f).cuda(), _T(c).cuda(), _T(R).cuda()).cpu().numpy()
    y = ro.perspective_projection(_T(pts), _T(t), _T(f), _T(c), _T(R)).numpy()
    assert rel_err(x, y) < 1e-5


def test_reprojection_closure_matches_oracle(ctx):
    """loss and both terms 1e-5; d y_angle / d body translations / d camera position 1e-4."""
    from uuo_mocap_b200 import reprojection as rp
    from oracle import reproj_oracle as ro
    model, smpl, orc = ctx
    d = _case(model, F=8, seed=42)
    F = 8
    w = {"reprojection": 1.0, "chamfer": 10.0}
    out = []
    for dev in ("cpu", "cuda:0"):
        X = lambda k: _T(d[k]).to(dev)
        args = (X("markers"), X("pose_body"), X("hmr_betas"), X("hmr_root"), X("trans"), X("pred_cam"), X("center"), X("size"), X("scale"))
        st = (ro.reprojection_setup(orc, *args) if dev == "cpu" else rp.reprojection_setup(smpl, *args))
        y = torch.full((1,), 0.35, device=dev, requires_grad=True)
        tr = (st["smpl_trans0"] + 0.03).clone().requires_grad_(True)
        cam = (st["cam_single0"] - 0.02).clone().requires_grad_(True)
        brow = X("betas").expand(F, -1).contiguous()
        if dev == "cpu":
            loss, l_rep, l_ch = ro.reprojection_loss(orc, st, X("markers"), X("pose_body"), brow, X("hmr_root"),
                                                     y.reshape(1, 1, 1, 1), tr[None], cam, w)
        else:
            loss, l_rep, l_ch, kp = rp.reprojection_loss(smpl, st, X("markers"), X("pose_body"), brow, X("hmr_root"), y, tr, cam, w)
            assert kp.shape == (F, 45, 2)
        loss.backward()
        out.append((float(loss), float(l_rep), float(l_ch), y.grad.cpu().numpy(), tr.grad.cpu().numpy(), cam.grad.cpu().numpy()))
    for k in (0, 1, 2):
        assert abs(out[1][k] - out[0][k]) <= 1e-5 * abs(out[0][k]), k
    assert out[0][1] > 0 and out[0][2] > 0
    for k in (3, 4, 5):
        assert rel_err(out[1][k], out[0][k]) < 1e-4, k


@pytest.mark.parametrize("optimizer", ["lbfgs", "adam"])
def test_optim_reprojection_runs_and_reports_like_the_reference(ctx, optimizer):
    from uuo_mocap_b200 import reprojection as rp
    model, smpl, orc = ctx
    F = 8
    d = _case(model, F=F, seed=43)
    cfg = {"optimizer": {"name": optimizer, "tolerance_grad": 1e-9, "tolerance_change": 1e-12, "lr": 0.01},
           "stages": {"reprojection_part": {"losses": {"reprojection": 1.0, "chamfer": 10.0}}}}
    seen = []

    def iter_fn(**kw):
        assert kw["stage"] == "reprojection" and kw["pred_2d_joints"].shape == (F, 45, 2) and kw["trans"].shape == (F, 3)
        seen.append(kw["iteration"])
    hist = []
    import builtins
    real_print = builtins.print
    builtins.print = lambda *a, **k: hist.append(a[2]) if a and a[0] == "Reprojection" else real_print(*a, **k)
    try:
        out = rp.optim_reprojection(_T(d["markers"]), _T(d["pose_body"]), _T(d["betas"]), _T(d["hmr_betas"]), _T(d["hmr_root"]),
                                    _T(d["trans"]), _T(d["pred_cam"]), _T(d["center"]), _T(d["size"]), _T(d["scale"]),
                                    torch.tensor(0.4), torch.ones(F, dtype=torch.bool), smpl, 25, cfg, verbose=True, iter_fn=iter_fn)
    finally:
        builtins.print = real_print
    assert len(seen) == len(hist) > 5 and min(hist) < hist[0]
    shapes = {"pose_body": (1, F, 23, 3, 3), "betas": (1, F, 10), "root_orient": (1, F, 1, 3, 3), "trans": (1, F, 3),
              "joints_2d": (1, F, 45, 2), "joints_2d_gt": (1, F, 45, 2), "cam_trans": (1, F, 3), "camera_center": (F, 2),
              "focal_length": (1, 2), "reproject_mask": (F,)}
    for k, s in shapes.items():
        assert tuple(out[k].shape) == s, k
    assert abs(out["input_angle"] - 0.4) < 1e-6 and isinstance(out["output_angle"], float)
    assert out["metrics"]["chamfer"] > 0 and out["metrics"]["reproject"] >= 0
    assert float(out["reproject_mask"].sum()) == F - 2
    R = out["root_orient"][0, :, 0]
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3, device=R.device).expand(F, 3, 3), atol=1e-5)
