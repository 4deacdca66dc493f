"""pipeline.fit_capture: the stage order of multimodal_video_mocap (multimodal.py:184-708) on tensors -- output dict of
the reference, the fit moves the body onto the markers, and the result goes through the stage-II writer."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.parametrize("with_root", [False, True])
def test_fit_capture_end_to_end(get_model, tmp_path, with_root):
    from uuo_mocap_b200 import fit, ops, pipeline, stageii, synthetic
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    model = get_model("top4")
    smpl = SMPL(_as_struct(model), device="cuda:0")
    F = 24
    cap = synthetic.make_capture(model, F, num_markers=41, seed=51, noise_std=0.001, n_ghost=2, dropout_p=0.05, permute=False,
                                 prior_yaw_err=0.1, prior_trans_err=0.05)
    cfg = fit.default_config(num_iters=150, optimizer="adam", lr=0.01)
    cfg["num_root_orient_angles"] = 2
    cfg["recompute_marker_labels"] = True
    if with_root:
        cfg["stages"]["root"] = {"constrained_rotation": True, "yaw_lock": True, "num_iters": 10, "lr": 1.0, "lr_adam": 0.01,
                                 "single_directional": True, "losses": {"full_chamfer": 10.0, "trans_vel": 1.0}}
    mk = _T(cap.markers)
    prior = (_T(cap.prior_pose_body), _T(cap.prior_betas), _T(cap.prior_root_orient), _T(cap.prior_trans))

    def residual(pose, betas, root, trans):
        o = smpl(betas=betas.cuda().reshape(-1, 10)[:1].expand(F, -1), body_pose=pose.cuda(), global_orient=root.cuda(),
                 transl=trans.cuda(), pose2rot=False)
        d, _ = ops.nn_argmin(mk.cuda(), o.vertices.contiguous())
        w = ops.get_marker_mask(mk.cuda())
        w[:, -2:] = 0                                     # ghosts have no vertex to sit on
        return float(torch.sqrt((d * w).sum() / w.sum()))
    before = residual(*prior)
    out = pipeline.fit_capture(mk, *prior, smpl, cfg, mocap_frame_rate=100.0)
    after = residual(out["pose_body"], out["betas"], out["root_orient"], out["trans"])
    assert after < 0.75 * before, (before, after)
    assert out["trans"].shape == (F, 3) and out["root_orient"].shape == (F, 1, 3, 3)
    assert out["pose_body"].shape == (F, 23, 3, 3) and out["betas"].shape == (F, 10)
    assert out["markers_labels"].shape == (F, mk.shape[1]) and len(out["scores"]) == 2
    assert out["mocap_markers"].shape == tuple(mk.shape) and out["mocap_frame_rate"] == 100.0
    want = {"chamfer", "marker", "marker_final"} | ({"root"} if with_root else set())      # test/test.py:131-143
    assert set(out["stages"]) == want
    R = out["pose_body"].reshape(-1, 3, 3)
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand_as(R), atol=1e-5)
    name = str(tmp_path / "take_stageii.npz")
    stageii.save_stageii(name, out["betas"][:1], out["trans"], out["root_orient"], out["pose_body"], out["mocap_frame_rate"],
                         out["mocap_markers"], stages=out["stages"])
    z = stageii.load_stageii(name)
    assert z["poses"].shape == (F, 72) and z["betas"].shape == (10,)
    zs = stageii.load_stageii(str(tmp_path / "take_stageii.marker_final.npz"))
    assert zs["poses"].shape == (F, 72)


def test_fit_capture_resamples_half_rate_prior(get_model):
    """multimodal.py:144-182: a 50 Hz video prior under a 100 Hz capture is brought to the marker rate first."""
    from uuo_mocap_b200 import fit, pipeline, synthetic
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    model = get_model("top4")
    smpl = SMPL(_as_struct(model), device="cuda:0")
    F = 24
    cap = synthetic.make_capture(model, F, num_markers=41, seed=52, noise_std=0.001, permute=False)
    cfg = fit.default_config(num_iters=20, optimizer="adam", lr=0.01)
    mk = _T(cap.markers)
    half = (_T(cap.prior_pose_body)[::2], _T(cap.prior_betas), _T(cap.prior_root_orient)[::2], _T(cap.prior_trans)[::2])
    out = pipeline.fit_capture(mk, *half, smpl, cfg, mocap_frame_rate=100.0, img_frame_rate=50.0)
    assert out["trans"].shape == (F, 3) and out["pose_body"].shape == (F, 23, 3, 3)
    assert torch.isfinite(out["trans"]).all() and torch.isfinite(out["pose_body"]).all()
    o2 = pipeline.fit_capture(mk, *half, smpl, cfg, mocap_frame_rate=100.0)       # no rate given: truncated (:191-199)
    assert o2["trans"].shape == (F // 2, 3)
