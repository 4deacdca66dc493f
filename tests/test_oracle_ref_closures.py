"""oracle/closures.py against the REFERENCE'S OWN SOURCE: tests/golden/closure_ref.npz holds the outputs of
optim_chamfer / optim_markers / compute_nearest_points / closest_point / ... compiled unmodified from
/root/reference/src/video_mocap (oracle/gen_golden_ref_closures.py), on the inputs of closure_top4.npz.
CPU only; the CUDA closures are checked against the same file in tests/test_fit_gpu.py."""
import numpy as np
import pytest
import torch

from conftest import rel_err


@pytest.fixture(scope="module")
def ctx(get_model):
    from oracle.smpl_oracle import OracleSMPL
    m = get_model("top4")
    return m, OracleSMPL.from_model(m)


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).clone()


def test_golden_files_share_inputs(ctx, golden):
    m, _ = ctx
    assert float(golden("closure_ref.npz")["model_checksum"]) == float(golden("closure_top4.npz")["model_checksum"]) == m.checksum()


@pytest.mark.parametrize("trans_vel", [True, False])
def test_chamfer_closure_matches_reference_source(ctx, golden, trans_vel):
    from oracle import closures
    _, orc = ctx
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    tag = "ch" if trans_vel else "ch0"
    pose, betas, trans, z = (_T(g[k]).requires_grad_(True) for k in ("ch_pose", "ch_betas", "ch_trans", "ch_z"))
    losses = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0}
    if trans_vel:
        losses["trans_vel"] = 1.0
    loss = closures.chamfer_stage_loss(orc, _T(g["markers"]), pose, _T(g["ch_o_pose"]), betas, _T(g["ch_o_betas"]),
                                       _T(g["ch_root"]), trans, z, losses)
    loss.backward()
    assert abs(float(loss) - float(r[tag + "_loss"])) <= 1e-6 * abs(float(r[tag + "_loss"]))
    for t, k in ((pose, "g_pose"), (betas, "g_betas"), (trans, "g_trans"), (z, "g_z")):
        assert rel_err(t.grad.numpy(), r[tag + "_" + k]) < 1e-5, k


def test_marker_closure_matches_reference_source(ctx, golden):
    from oracle import closures
    _, orc = ctx
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    pose, betas, root, trans = (_T(g[k]).requires_grad_(True) for k in ("ch_pose", "ch_betas", "mk_root", "ch_trans"))
    loss = closures.marker_stage_loss(orc, _T(g["markers"]), pose, _T(g["ch_o_pose"]), betas, _T(g["ch_o_betas"]), root, trans,
                                      torch.from_numpy(g["mk_vidx"]), {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0,
                                                                       "temporal": 1.0})
    loss.backward()
    assert abs(float(loss) - float(r["mk_loss"])) <= 1e-6 * abs(float(r["mk_loss"]))
    for t, k in ((pose, "mk_g_pose"), (betas, "mk_g_betas"), (root, "mk_g_root"), (trans, "mk_g_trans")):
        assert rel_err(t.grad.numpy(), r[k]) < 1e-5, k


def test_compute_nearest_points_matches_reference_source(ctx, golden):
    from oracle import closures
    _, orc = ctx
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    idx, _ = closures.mean_distance_argmin(orc, _T(g["markers"]), _T(g["ch_pose"]), _T(g["ch_betas"]), _T(g["mk_root"]),
                                           _T(g["ch_trans"]), _T(r["np_img_mask"]))
    assert np.array_equal(idx, r["np_vertex_idx"])


def test_helpers_match_reference_source(ctx, golden):
    from oracle import closures, p3d_oracle as p3d
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    mk = _T(g["markers"])
    assert np.array_equal(closures.get_marker_mask(mk).numpy(), r["marker_mask"])
    assert np.array_equal(p3d.compute_root_orient_z(_T(r["rot_z_angles"])).numpy(), r["rot_z"])
    # closest_point (utils/points.py:19-42): un-squared distances, first minimum
    v = _T(r["cp_vertices"])
    d2, idx = p3d.knn_k1(mk[:1], v[None])
    same = idx[0].numpy() == r["cp_idx"]
    assert same.mean() > 0.95            # sqrt can merge two neighbouring squared distances into a tie
    assert rel_err(np.sqrt(d2[0].numpy()), r["cp_dist"]) < 1e-6
    x = _T(r["wmse_x"])
    w = torch.arange(4.0).view(4, 1, 1)
    assert abs(float(torch.mean(torch.nn.functional.mse_loss(x, x * 0.5, reduction="none") * w)) - float(r["wmse"])) < 1e-7


def test_part_fit_closure_matches_reference_source(ctx, golden):
    """closure_fit_subtree of find_best_part_fits (markers_utils.py:454-562), all 26 three-joint sub-trees at z = 0.2."""
    from oracle import closures
    m, orc = ctx
    r = golden("closure_ref.npz")
    vlab = np.argmax(m.weights, axis=-1)
    rot, mk = _T(r["pf_rot"]), _T(r["pf_markers"])
    Fp = mk.shape[0]
    for s, subtree in enumerate(r["pf_subtrees"]):
        vidx = torch.from_numpy(np.concatenate([np.where(vlab == j)[0] for j in subtree]))
        z = torch.full((1, 1, 1), 0.2, requires_grad=True)
        tr = torch.median(mk, dim=1)[0].clone().requires_grad_(True)
        betas = _T(r["pf_betas"]).requires_grad_(True)
        loss = closures.part_fit_loss(orc, mk, rot[:, 1:], betas, _T(r["pf_betas"]), rot[:, :1], tr, z, vidx,
                                      {"chamfer": 10.0, "reg_betas": 0.1})
        loss.backward()
        assert abs(float(loss) - float(r["pfA_loss"][s])) <= 1e-6 * abs(float(r["pfA_loss"][s])), subtree
        assert rel_err(z.grad.numpy().reshape(-1), r["pfA_g_z"][s]) < 1e-5
        assert rel_err(tr.grad.numpy(), r["pfA_g_trans"][s]) < 1e-5
        assert rel_err(betas.grad.numpy(), r["pfA_g_betas"][s]) < 1e-5
        if s >= 7:
            break          # the GPU test walks all of them; a handful is enough for the CPU suite's time budget
