"""Fitting-loop parity: CUDA closures (loss, correspondences, every gradient) against the oracle closures
and their golden vectors; fused Adam against torch.optim.Adam on the oracle closure (end-point RMSE)."""
import numpy as np
import pytest
import torch

from conftest import frame_rel_err, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(get_model):
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    from oracle.smpl_oracle import OracleSMPL
    m = get_model("top4")
    return m, SMPL(_as_struct(m), device="cuda:0"), OracleSMPL.from_model(m)


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def test_chamfer_closure_golden(ctx, golden):
    """closure_stage_chamfer incl. the optional trans_vel term: loss 1e-5, idx exact, grads 1e-4."""
    from uuo_mocap_b200.fit import FitProblem, ParamBlock, chamfer_closure
    _, smpl, _ = ctx
    g = golden("closure_top4.npz")
    F = g["markers"].shape[0]
    prob = FitProblem(smpl, _T(g["markers"]), _T(g["ch_o_pose"]), _T(g["ch_o_betas"]))
    pb = ParamBlock.build({"trans": _T(g["ch_trans"]), "z_angle": _T(g["ch_z"]).reshape(F), "betas": _T(g["ch_betas"]),
                           "pose_body": _T(g["ch_pose"])}, "cuda:0")
    losses = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0, "trans_vel": 1.0}
    loss = chamfer_closure(prob, pb, _T(g["ch_root"]).cuda(), losses)
    assert abs(float(loss) - float(g["ch_loss"])) <= 1e-5 * abs(float(g["ch_loss"]))
    on = prob.mask.cpu().numpy() != 0           # dropped markers weigh 0: the fit's kernel does not search them
    assert on.sum() > 0.8 * on.size and np.array_equal(prob.idx.cpu().numpy().astype(np.int64)[on], g["ch_idx"][on])
    loss2 = chamfer_closure(prob, pb, _T(g["ch_root"]).cuda(), losses)      # second call starts from the first one's matches
    assert abs(float(loss2) - float(loss)) <= 1e-6 * abs(float(loss))
    assert np.array_equal(prob.idx.cpu().numpy().astype(np.int64)[on], g["ch_idx"][on])
    r = golden("closure_ref.npz")      # the same closure evaluated by the REFERENCE'S OWN SOURCE (optimization.py:147-285)
    assert abs(float(loss) - float(r["ch_loss"])) <= 1e-5 * abs(float(r["ch_loss"]))
    for k, ref in (("pose_body", "ch_g_pose"), ("betas", "ch_g_betas"), ("trans", "ch_g_trans"), ("z_angle", "ch_g_z")):
        got = pb.gviews[k].cpu().numpy().reshape(g[ref].shape)
        assert rel_err(got, g[ref]) < 1e-4, k
        assert rel_err(got, r[ref]) < 1e-4, k
        if k != "betas":
            assert frame_rel_err(got, r[ref]) < 1e-4, k      # normalised frame by frame


def test_marker_closure_golden(ctx, golden):
    from uuo_mocap_b200.fit import FitProblem, ParamBlock, marker_closure
    _, smpl, _ = ctx
    g = golden("closure_top4.npz")
    prob = FitProblem(smpl, _T(g["markers"]), _T(g["ch_o_pose"]), _T(g["ch_o_betas"]))
    pb = ParamBlock.build({"pose_body": _T(g["ch_pose"]), "betas": _T(g["ch_betas"]), "root_orient": _T(g["mk_root"]),
                           "trans": _T(g["ch_trans"])}, "cuda:0")
    losses = {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0, "temporal": 1.0}
    loss = marker_closure(prob, pb, _T(g["mk_vidx"]).to(torch.int32).cuda(), losses)
    assert abs(float(loss) - float(g["mk_loss"])) <= 1e-5 * abs(float(g["mk_loss"]))
    r = golden("closure_ref.npz")      # optim_markers compiled from the reference's source (optimization.py:288-399)
    assert abs(float(loss) - float(r["mk_loss"])) <= 1e-5 * abs(float(r["mk_loss"]))
    for k, ref in (("pose_body", "mk_g_pose"), ("betas", "mk_g_betas"), ("trans", "mk_g_trans"), ("root_orient", "mk_g_root")):
        got = pb.gviews[k].cpu().numpy().reshape(g[ref].shape)
        assert rel_err(got, g[ref]) < 1e-4, k
        assert rel_err(got, r[ref]) < 1e-4, k
        if k != "betas":
            assert frame_rel_err(got, r[ref]) < 1e-4, k


def test_compute_nearest_points(ctx, golden):
    from uuo_mocap_b200.fit import compute_nearest_points
    from oracle import closures
    _, smpl, orc = ctx
    g = golden("closure_top4.npz")
    mk, pose, betas, root, trans = (_T(g[k]) for k in ("markers", "ch_pose", "ch_betas", "mk_root", "ch_trans"))
    F = mk.shape[0]
    img_mask = torch.ones(F)
    img_mask[1] = 0
    idx_ref, mean_ref = closures.mean_distance_argmin(orc, mk, pose, betas, root, trans, img_mask)
    one_hot = compute_nearest_points(mk, pose.cuda(), betas.cuda(), root.cuda(), trans.cuda(), smpl, img_mask=img_mask)
    assert one_hot.shape == (mk.shape[1], 6890) and float(one_hot.sum()) == mk.shape[1]
    idx = torch.argmax(one_hot, dim=1).cpu().numpy()
    # vertices differ from the oracle's by ~1e-7: allow a different pick only on a near-tie
    for m in np.where(idx != idx_ref)[0]:
        assert abs(mean_ref[m, idx[m]] - mean_ref[m, idx_ref[m]]) <= 1e-6 * mean_ref[m, idx_ref[m]]
    assert (idx == idx_ref).mean() > 0.9
    r = golden("closure_ref.npz")          # compute_nearest_points compiled from the reference's source (:402-642)
    assert np.array_equal(r["np_img_mask"], img_mask.numpy())
    for m in np.where(idx != r["np_vertex_idx"])[0]:
        assert abs(mean_ref[m, idx[m]] - mean_ref[m, r["np_vertex_idx"][m]]) <= 1e-6 * mean_ref[m, idx_ref[m]]


@pytest.mark.parametrize("permute", [False, True])
def test_compute_nearest_points_bounded_path(get_model, permute, monkeypatch):
    """compute_nearest_points(use_mean) picks the bounded mean-distance kernel when a marker column follows one point of
    the body (trajectories) and the plain kernel for shuffled columns; forcing either path gives the same indices."""
    from uuo_mocap_b200 import fit, ops, synthetic
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    model = get_model("top4")
    smpl = SMPL(_as_struct(model), device="cuda:0")
    F = 220
    cap = synthetic.make_capture(model, F, num_markers=41, seed=9, n_ghost=4, dropout_p=0.05, noise_std=0.002,
                                 permute=permute)
    mk = torch.from_numpy(cap.markers).cuda()
    rot = synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)
    args = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (rot[:, 1:], cap.gt_betas, rot[:, :1], cap.gt_trans)]
    img_mask = torch.ones(F)
    img_mask[[0, 7]] = 0
    used = []
    real = ops.mean_distance_accumulate_bounded
    monkeypatch.setattr(ops, "mean_distance_accumulate_bounded", lambda *a, **k: (used.append(1), real(*a, **k))[1])
    run = lambda: fit.compute_nearest_points(mk, args[0], args[1], args[2], args[3], smpl, img_mask=img_mask,
                                             as_one_hot=False)
    idx_auto = run()
    assert bool(used) == (not permute)
    monkeypatch.setattr(fit, "BOUNDED_MIN_RATIO", 0.0)
    idx_bounded = run()
    monkeypatch.setattr(fit, "BOUNDED_MIN_RATIO", float("inf"))
    idx_plain = run()
    assert torch.equal(idx_auto, idx_plain) and torch.equal(idx_bounded, idx_plain)
    if not permute:      # the columns are trajectories: the pick is the attachment vertex or a neighbour of it
        own = torch.from_numpy(cap.marker_vertex_ids).cuda()
        with torch.no_grad():
            v0 = smpl(betas=args[1], body_pose=args[0][:1], global_orient=args[2][:1], transl=args[3][:1],
                      pose2rot=False).vertices[0]
        d = (v0[idx_plain[:41].long()] - v0[own]).norm(dim=-1)
        assert float(d.median()) < 0.05


def test_optim_chamfer_bakes_yaw_like_the_reference(ctx, golden):
    """optim_chamfer with zero iterations of change: root_orient[:] = Rz(z) @ root_orient (optimization.py:280-285),
    against the tensor the reference's own function left behind."""
    from uuo_mocap_b200.fit import rot_z
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    z = _T(g["ch_z"]).cuda()
    baked = (rot_z(z.view(-1, 1, 1)) @ _T(g["ch_root"]).cuda()).cpu().numpy()
    assert rel_err(baked, r["ch_root_baked"]) < 1e-6
    assert rel_err(rot_z(_T(r["rot_z_angles"]).cuda()).cpu().numpy(), r["rot_z"]) < 1e-6


def _setup_fit(ctx, F, seed, n_ghost=2, dropout=0.1):
    from uuo_mocap_b200 import synthetic
    model, smpl, orc = ctx
    cap = synthetic.make_capture(model, F, num_markers=41, seed=seed, n_ghost=n_ghost, dropout_p=dropout, dropout_block=2,
                                 prior_pose_noise=0.05, prior_trans_err=0.03, prior_yaw_err=0.1)
    return cap


def test_adam_fit_matches_oracle_adam(ctx):
    """Fixed iteration count, same init / lr: fused Adam on the CUDA closure vs torch.optim.Adam on the oracle
    closure (CPU).  Final fitted marker RMSE within 0.5 mm (BASELINE.json north_star)."""
    from uuo_mocap_b200 import fit
    from oracle import closures
    model, smpl, orc = ctx
    F, iters, lr = 10, 25, 0.01
    cap = _setup_fit(ctx, F, seed=31)
    mk = _T(cap.markers)
    cfg = fit.default_config(num_iters=iters, optimizer="adam", lr=lr)

    def init():
        return (_T(cap.prior_pose_body).clone(), _T(cap.prior_betas).clone(), _T(cap.prior_root_orient).clone(),
                _T(cap.prior_trans).clone())
    o_pose, o_betas = _T(cap.prior_pose_body), _T(cap.prior_betas)

    # ---- oracle: chamfer stage -> correspondences -> marker stage
    pose, betas, root, trans = [t.requires_grad_(True) for t in init()]
    root.requires_grad_(False)
    z = torch.zeros(F, 1, 1, requires_grad=True)
    lc = cfg["stages"]["chamfer"]["losses"]
    closures.run_adam([trans, z, betas, pose],
                      lambda: closures.chamfer_stage_loss(orc, mk, pose, o_pose, betas, o_betas, root, trans, z, lc), iters, lr)
    from oracle import p3d_oracle as p3d
    with torch.no_grad():
        root = (p3d.compute_root_orient_z(z) @ root).detach()
    vidx_ref, _ = closures.mean_distance_argmin(orc, mk, pose.detach(), betas.detach(), root, trans.detach(), torch.ones(F))
    root.requires_grad_(True)
    lm = cfg["stages"]["marker"]["losses"]
    closures.run_adam([pose, betas, root, trans],
                      lambda: closures.marker_stage_loss(orc, mk, pose, o_pose, betas, o_betas, root, trans,
                                                         torch.from_numpy(vidx_ref), lm), iters, lr)

    def rmse(orc_pose, orc_betas, orc_root, orc_trans, vidx):
        with torch.no_grad():
            v = orc.forward(orc_betas, p3d.normalize_rot(orc_pose), p3d.normalize_rot(orc_root), orc_trans, pose2rot=False).vertices
        d = torch.norm(mk - v[:, vidx], dim=-1)
        w = closures.get_marker_mask(mk).float()
        return float(torch.sqrt((w * (d - closures.MARKER_DISTANCE) ** 2).sum() / w.sum()))
    rmse_ref = rmse(pose.detach(), betas.detach(), root.detach(), trans.detach(), torch.from_numpy(vidx_ref))

    # ---- CUDA path through the reference's stage functions
    c_pose, c_betas, c_root, c_trans = [t.cuda() for t in init()]
    fit.optim_chamfer(mk, c_pose, o_pose, c_betas, o_betas, c_root, c_trans, torch.ones(F), None, smpl, cfg)
    one_hot = fit.compute_nearest_points(mk, c_pose, c_betas, c_root, c_trans, smpl, img_mask=torch.ones(F), config=cfg)
    fit.optim_markers(mk, c_pose, o_pose, c_betas, o_betas, c_root, c_trans, one_hot, torch.ones(F), smpl, cfg)
    vidx = torch.argmax(one_hot, dim=1).cpu()
    rmse_gpu = rmse(c_pose.cpu(), c_betas.cpu(), c_root.cpu(), c_trans.cpu(), vidx)
    assert abs(rmse_gpu - rmse_ref) < 0.5e-3, (rmse_gpu, rmse_ref)
    assert (vidx.numpy() == vidx_ref).mean() > 0.9
    assert rel_err(c_trans.cpu().numpy(), trans.detach().numpy()) < 1e-3
    assert rel_err(c_betas.cpu().numpy(), betas.detach().numpy()) < 1e-2


def test_lbfgs_drives_the_cuda_closure(ctx):
    """The reference's optimiser (torch LBFGS, strong Wolfe) over the CUDA closure: the loss must go down."""
    from uuo_mocap_b200 import fit
    model, smpl, orc = ctx
    F = 8
    cap = _setup_fit(ctx, F, seed=33)
    cfg = fit.default_config(num_iters=8, optimizer="lbfgs")
    mk = _T(cap.markers)
    pose, betas, root, trans = (_T(cap.prior_pose_body).cuda(), _T(cap.prior_betas).cuda(),
                                _T(cap.prior_root_orient).cuda(), _T(cap.prior_trans).cuda())
    o_pose, o_betas = _T(cap.prior_pose_body), _T(cap.prior_betas)
    prob = fit.FitProblem(smpl, mk, o_pose, o_betas)
    pb = fit.ParamBlock.build({"trans": trans, "z_angle": torch.zeros(F), "betas": betas, "pose_body": pose}, "cuda:0")
    l0 = float(fit.chamfer_closure(prob, pb, root, cfg["stages"]["chamfer"]["losses"]))
    fit.optim_chamfer(mk, pose, o_pose, betas, o_betas, root, trans, None, None, smpl, cfg)
    pb = fit.ParamBlock.build({"trans": trans, "z_angle": torch.zeros(F), "betas": betas, "pose_body": pose}, "cuda:0")
    l1 = float(fit.chamfer_closure(prob, pb, root, cfg["stages"]["chamfer"]["losses"]))
    assert l1 < l0


def test_stacked_hypotheses_equal_separate_fits(ctx):
    """MHT batched along the frame axis == the same hypotheses fitted one after the other (multimodal.py:462-574)."""
    from uuo_mocap_b200 import fit
    model, smpl, orc = ctx
    F, H = 6, 3
    cap = _setup_fit(ctx, F, seed=35, n_ghost=0, dropout=0.0)
    cfg = fit.default_config(num_iters=6, optimizer="adam", lr=0.01)
    mk = _T(cap.markers)
    o_pose, o_betas = _T(cap.prior_pose_body), _T(cap.prior_betas)
    args = (mk, _T(cap.prior_pose_body), o_pose, _T(cap.prior_betas), o_betas, _T(cap.prior_root_orient), _T(cap.prior_trans),
            torch.ones(F), smpl, cfg)
    angles = [0.0, 2 * np.pi / 3, 4 * np.pi / 3]
    best, scores, fitted = fit.fit_hypotheses(*args, angles=angles)
    for h, a in enumerate(angles):
        b1, s1, f1 = fit.fit_hypotheses(*args, angles=[a])
        assert abs(s1[0] - scores[h]) <= 1e-4 * abs(s1[0])
        assert rel_err(fitted[h]["trans"].cpu().numpy(), f1[0]["trans"].cpu().numpy()) < 1e-4
    assert best == int(np.argmin(scores))


@pytest.mark.parametrize("kind", ["top4", "dense"])
@pytest.mark.parametrize("shared,R,F", [(1, 45, 130), (1, 7, 3), (0, 45, 70), (0, 64, 5), (1, 80, 9),
                                        (2, 45, 140), (3, 7, 9), (4, 41, 512)])   # > 1: that many row groups
def test_rows_paths_match_dense_path(get_model, kind, shared, R, F):
    """uuo_lbs_forward_rows == gather of the dense forward; uuo_lbs_backward_rows (re-evaluating v_posed) ==
    uuo_lbs_backward_rows_cached (v_posed from the forward / from uuo_lbs_cache_vposed_rows) == the dense backward
    fed a gradient that is non-zero on those rows only."""
    from uuo_mocap_b200 import _lib, synthetic
    from uuo_mocap_b200._lib import check, ptr, stream_ptr
    from uuo_mocap_b200.smpl import SMPL, _as_struct, POSE_RAW6D
    model = get_model(kind)
    smpl = SMPL(_as_struct(model), device="cuda:0")
    mh, dev, L = smpl.handle, torch.device("cuda:0"), _lib.lib()
    V = mh.V
    rng = np.random.default_rng(R * 100 + F)
    mb = synthetic.make_lbs_microbench(F, seed=R)
    rot = torch.from_numpy(mb["rot"] + 0.05 * rng.normal(size=mb["rot"].shape).astype(np.float32)).to(dev)
    go, bp = rot[:, :1].contiguous(), rot[:, 1:].contiguous()
    betas, transl = torch.from_numpy(mb["betas"][:1]).to(dev), torch.from_numpy(mb["transl"]).to(dev)
    if shared == 1:
        rows = torch.from_numpy(rng.choice(V, size=R, replace=False).astype(np.int32)).to(dev)
        rows_f = rows.long().unsqueeze(0).expand(F, R)
    elif shared > 1:      # rows[G,R]: frame f uses set f // (F // G)
        rows = torch.from_numpy(rng.integers(0, V, size=(shared, R)).astype(np.int32)).to(dev)
        rows_f = rows.long().repeat_interleave(F // shared, dim=0)
    else:
        rows = torch.from_numpy(rng.integers(0, V, size=(F, R)).astype(np.int32)).to(dev)
        rows_f = rows.long()
    g_rows = torch.from_numpy(rng.normal(size=(F, R, 3)).astype(np.float32)).to(dev)
    g_rows[::2, ::5] = 0.0                                   # masked rows are skipped
    ws, st = mh.workspace(F), stream_ptr(dev)
    verts = torch.empty((F, V, 3), device=dev)
    check(L.uuo_lbs_forward(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, ptr(transl), F, ptr(verts), None,
                            ptr(ws), st))
    # dense backward with the sparse gradient scattered (duplicates add up)
    gv = torch.zeros((F, V, 3), device=dev)
    gv.view(F * V, 3).index_add_(0, (rows_f + torch.arange(F, device=dev).unsqueeze(1) * V).reshape(-1), g_rows.view(-1, 3))
    outs = {}

    def grads():
        return [torch.zeros_like(betas), torch.zeros_like(go), torch.zeros_like(bp), torch.zeros_like(transl)]
    gd = grads()
    check(L.uuo_lbs_backward(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, F, ptr(gv), None, ptr(ws),
                             ptr(gd[0]), ptr(gd[1]), ptr(gd[2]), None, ptr(gd[3]), st))
    outs["dense"] = gd
    out_rows = torch.empty((F, R, 3), device=dev)
    check(L.uuo_lbs_forward_rows(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, ptr(transl), F, ptr(rows),
                                 int(shared), R, ptr(out_rows), None, ptr(ws), st))
    ref_rows = torch.gather(verts, 1, rows_f.unsqueeze(-1).expand(F, R, 3))
    assert rel_err(out_rows.cpu().numpy(), ref_rows.cpu().numpy()) < 1e-5
    variants = [("recompute", L.uuo_lbs_backward_rows)]
    if R <= 64:
        variants.append(("cached", L.uuo_lbs_backward_rows_cached))
    for name, fn in variants:
        g = grads()
        check(fn(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, F, ptr(rows), int(shared), R,
                 ptr(g_rows), None, ptr(ws), ptr(g[0]), ptr(g[1]), ptr(g[2]), None, ptr(g[3]), st))
        outs[name] = g
    if R <= 64:   # v_posed rows picked out of a dense forward's workspace by position (= vertex id on the model itself)
        check(L.uuo_lbs_forward(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, ptr(transl), F, ptr(verts),
                                None, ptr(ws), st))
        pos = rows_f.to(torch.int32).contiguous()
        check(L.uuo_lbs_cache_vposed_rows(mh.handle, F, ptr(pos), R, ptr(ws), st))
        g = grads()
        check(L.uuo_lbs_backward_rows_cached(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, F, ptr(rows),
                                             int(shared), R, ptr(g_rows), None, ptr(ws), ptr(g[0]), ptr(g[1]),
                                             ptr(g[2]), None, ptr(g[3]), st))
        outs["cached_from_dense"] = g
        if shared:   # shared rows on the tensor cores (rows_tc.cu, 3xTF32 GEMMs) against the per-frame fp32 kernel
            g2 = grads()
            per_frame = rows_f.to(torch.int32).contiguous()
            check(L.uuo_lbs_backward_rows_cached(mh.handle, ptr(betas), 1, ptr(go), ptr(bp), POSE_RAW6D, None, F,
                                                 ptr(per_frame), 0, R, ptr(g_rows), None, ptr(ws), ptr(g2[0]), ptr(g2[1]),
                                                 ptr(g2[2]), None, ptr(g2[3]), st))
            for a, b in zip(g, g2):
                assert rel_err(a.cpu().numpy(), b.cpu().numpy()) < 1e-5
    for name, g in outs.items():
        if name == "dense":
            continue
        for a, b, what in zip(g, outs["dense"], ("betas", "global_orient", "body_pose", "transl")):
            assert rel_err(a.cpu().numpy(), b.cpu().numpy()) < 1e-4, (name, what)


def test_adam_graph_replay_equals_eager(ctx):
    """run_adam: 3 eager iterations + CUDA-graph replays (step number in device memory) == all-eager iterations."""
    from uuo_mocap_b200 import fit
    model, smpl, _ = ctx
    F, iters = 12, 17
    cap = _setup_fit(ctx, F, seed=33)
    mk = _T(cap.markers)
    res = {}
    for use_graph in (False, True):
        prob = fit.FitProblem(smpl, mk, _T(cap.prior_pose_body), _T(cap.prior_betas))
        pb = fit.ParamBlock.build({"trans": _T(cap.prior_trans), "z_angle": torch.zeros(F), "betas": _T(cap.prior_betas),
                                   "pose_body": _T(cap.prior_pose_body)}, "cuda:0")
        root = _T(cap.prior_root_orient).cuda()
        losses = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0, "trans_vel": 1.0}
        loss = fit.run_adam(pb, lambda: fit.chamfer_closure(prob, pb, root, losses), iters, 0.01, graph=use_graph)
        torch.cuda.synchronize()
        res[use_graph] = (pb.flat.clone(), float(loss), prob.idx.clone())
    assert torch.equal(res[True][2], res[False][2])
    assert rel_err(res[True][0].cpu().numpy(), res[False][0].cpu().numpy()) < 1e-6
    assert abs(res[True][1] - res[False][1]) <= 1e-6 * abs(res[False][1])
    # and against the host-step entry point
    from uuo_mocap_b200 import ops
    p = torch.randn(1000, device="cuda:0"); g = torch.randn(1000, device="cuda:0")
    s1 = [p.clone(), torch.zeros_like(p), torch.zeros_like(p)]
    s2 = [p.clone(), torch.zeros_like(p), torch.zeros_like(p)]
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    for it in range(1, 6):
        ops.adam_step(s1[0], g, s1[1], s1[2], 0.01, it)
        ops.adam_step_dev(s2[0], g, s2[1], s2[2], 0.01, cnt)
    assert int(cnt.item()) == 5
    assert torch.allclose(s1[0], s2[0], rtol=1e-6, atol=1e-7)


def test_marker_labels_from_coords_and_weighted_mse(ctx):
    """optimization.py:645-659, :718-724."""
    from uuo_mocap_b200.fit import compute_marker_labels_from_coords, weighted_mse_loss
    model, smpl, _ = ctx
    vids = torch.tensor([5, 900, 3000, 6889, 5])
    one_hot = torch.zeros(5, 6890)
    one_hot[torch.arange(5), vids] = 1.0
    want = torch.from_numpy(np.argmax(model.weights, axis=-1))[vids]
    for arg in (one_hot, vids):
        lab = compute_marker_labels_from_coords(smpl, arg, 7)
        assert lab.shape == (7, 5) and torch.equal(lab.cpu(), want.unsqueeze(0).expand(7, 5))
    a, b, w = torch.randn(4, 3, device="cuda:0"), torch.randn(4, 3, device="cuda:0"), torch.rand(4, 3, device="cuda:0")
    assert torch.allclose(weighted_mse_loss(a, b, w), (((a - b) ** 2) * w).mean())


def test_chamfer_closure_free_root_matches_reference_source(ctx, golden):
    """``yaw_lock: False`` (optimization.py:167-172, 194, 282-283): the free 3x3 per frame IS the root orientation inside
    the closure; loss, every gradient and the baked root against the reference's own function run from source."""
    from uuo_mocap_b200 import fit
    from uuo_mocap_b200.fit import FitProblem, ParamBlock, chamfer_closure
    _, smpl, _ = ctx
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    prob = FitProblem(smpl, _T(g["markers"]), _T(g["ch_o_pose"]), _T(g["ch_o_betas"]))
    pb = ParamBlock.build({"trans": _T(g["ch_trans"]), "z_angle": _T(r["ch3_z"]), "betas": _T(g["ch_betas"]),
                           "pose_body": _T(g["ch_pose"])}, "cuda:0")
    loss = chamfer_closure(prob, pb, _T(g["ch_root"]).cuda(), {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0})
    assert abs(float(loss) - float(r["ch3_loss"])) <= 1e-5 * abs(float(r["ch3_loss"]))
    for k, ref in (("pose_body", "ch3_g_pose"), ("betas", "ch3_g_betas"), ("trans", "ch3_g_trans"), ("z_angle", "ch3_g_z")):
        assert rel_err(pb.gviews[k].cpu().numpy().reshape(r[ref].shape), r[ref]) < 1e-4, k
    baked = (fit.normalize_rot(_T(r["ch3_z"]).cuda()) @ _T(g["ch_root"]).cuda()).cpu().numpy()
    assert rel_err(baked, r["ch3_root_baked"]) < 1e-6
    # through the stage function: zero iterations leave the identity offset, so root_orient comes back unchanged
    cfg = fit.default_config(num_iters=0, optimizer="adam")
    cfg["stages"]["chamfer"]["yaw_lock"] = False
    root = _T(g["ch_root"]).cuda()
    fit.optim_chamfer(_T(g["markers"]), _T(g["ch_pose"]).cuda(), _T(g["ch_o_pose"]), _T(g["ch_betas"]).cuda(),
                      _T(g["ch_o_betas"]), root, _T(g["ch_trans"]).cuda(), None, None, smpl, cfg)
    assert rel_err(root.cpu().numpy(), g["ch_root"]) < 1e-6


@pytest.mark.parametrize("gran", ["full", "marker", "part"])
def test_compute_nearest_points_closest_point_branch(ctx, golden, gran):
    """compute_locations.use_mean: False (optimization.py:503-510, 538-590): per-frame closest_point; the last frame that
    updates a marker's row wins.  Expected indices = the reference's own closest_point (utils/points.py:19-42) on that
    frame (the reference's matrix itself comes out all zero there: its scatter_ calls overwrite the 1, see the generator)."""
    from uuo_mocap_b200 import fit
    _, smpl, _ = ctx
    g, r = golden("closure_top4.npz"), golden("closure_ref.npz")
    cfg = fit.default_config()
    cfg["stages"]["compute_locations"] = {"use_velocity": False, "use_barycentric": False, "use_mean": False}
    mk, pose, betas, root, trans = (_T(g[k]) for k in ("markers", "ch_pose", "ch_betas", "mk_root", "ch_trans"))
    idx = fit.compute_nearest_points(mk, pose.cuda(), betas.cuda(), root.cuda(), trans.cuda(), smpl,
                                     marker_labels=_T(r["cp_labels"]), granularity=gran, img_mask=_T(r["cp_img_mask"]),
                                     config=cfg, as_one_hot=False).cpu().numpy()
    want = r["cp_idx_" + gran]
    assert (idx == want).mean() > 0.95          # vertices differ from the reference's by ~1e-7: a near-tie may flip
