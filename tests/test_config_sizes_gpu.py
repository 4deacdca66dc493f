"""Parity AT THE SIZES OF BASELINE.json's configs (the persistent-CTA / multi-wave / split-K regimes the bench runs in):
  configs[0]: F = 300, M = 41   both closures in full against the oracle (loss, indices, every gradient)
  configs[1]: F = 4096          lbs() forward + backward, 64 sampled frames against the oracle run on just those frames
  configs[2]: F = 5000, M = 45  chamfer + marker closures, 64 sampled frames (per-frame terms only) against the oracle
plus the L-BFGS end point (reference optimiser) of the CUDA closure against the oracle closure under the same L-BFGS.
Per-frame work is independent of the other frames in the sampled terms, so a sample is a fair check of the whole batch;
errors are normalised FRAME BY FRAME (conftest.frame_rel_err), not by one maximum over the tensor."""
import numpy as np
import pytest
import torch

from conftest import frame_rel_err, rel_err

pytestmark = pytest.mark.gpu


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.fixture(scope="module")
def ctx(get_model):
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    from oracle.smpl_oracle import OracleSMPL
    out = {}
    for kind in ("top4", "dense"):
        m = get_model(kind)
        out[kind] = (m, SMPL(_as_struct(m), device="cuda:0"), OracleSMPL.from_model(m))
    return out


def test_config0_closures_F300(ctx):
    """video_mocap.yaml fit of one 300-frame sequence, 41 markers: chamfer and marker closures in full."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.fit import FitProblem, ParamBlock, chamfer_closure, marker_closure
    from oracle import closures
    model, smpl, orc = ctx["top4"]
    F = 300
    cap = synthetic.make_capture(model, F, num_markers=41, seed=1, dropout_p=0.02)
    mk, opose, obetas = _T(cap.markers), _T(cap.prior_pose_body), _T(cap.prior_betas)
    root, trans = _T(cap.prior_root_orient), _T(cap.prior_trans)
    rng = np.random.default_rng(1)
    pose0 = opose + _T(rng.normal(0, 0.02, size=opose.shape).astype(np.float32))
    betas0 = obetas + 0.1
    z0 = _T(rng.normal(0, 0.1, size=(F, 1, 1)).astype(np.float32))
    lc = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0}
    lm = {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0}
    torch.set_num_threads(max(1, torch.get_num_threads()))
    # oracle
    pose, betas, tr, z = (t.clone().requires_grad_(True) for t in (pose0, betas0, trans, z0))
    loss_c, idx_ref = closures.chamfer_stage_loss(orc, mk, pose, opose, betas, obetas, root, tr, z, lc, return_idx=True)
    loss_c.backward()
    ref_c = {"pose_body": pose.grad.numpy(), "betas": betas.grad.numpy(), "trans": tr.grad.numpy(), "z_angle": z.grad.numpy()}
    vidx = torch.from_numpy(np.resize(cap.marker_vertex_ids, mk.shape[1]).astype(np.int64))
    pose, betas, rt, tr = (t.clone().requires_grad_(True) for t in (pose0, betas0, root, trans))
    loss_m = closures.marker_stage_loss(orc, mk, pose, opose, betas, obetas, rt, tr, vidx, lm)
    loss_m.backward()
    ref_m = {"pose_body": pose.grad.numpy(), "betas": betas.grad.numpy(), "root_orient": rt.grad.numpy(), "trans": tr.grad.numpy()}
    # CUDA
    prob = FitProblem(smpl, mk, opose, obetas)
    pb = ParamBlock.build({"trans": trans, "z_angle": z0.reshape(F), "betas": betas0, "pose_body": pose0}, "cuda:0")
    got = float(chamfer_closure(prob, pb, root.cuda(), lc))
    assert abs(got - float(loss_c)) <= 1e-5 * abs(float(loss_c))
    on = prob.mask.cpu().numpy() != 0
    same = prob.idx.cpu().numpy().astype(np.int64)[on] == idx_ref.numpy()[on]
    assert same.mean() > 0.999          # vertices differ from the oracle's by ~1e-7: a near-tie may flip
    for k, ref in ref_c.items():
        g = pb.gviews[k].cpu().numpy().reshape(ref.shape)
        assert rel_err(g, ref) < 1e-4, k
        if k in ("pose_body", "trans") and same.all():      # (a frame's z_angle is ONE number: nothing to normalise by)
            assert frame_rel_err(g, ref) < 2e-4, k
    pb = ParamBlock.build({"pose_body": pose0, "betas": betas0, "root_orient": root, "trans": trans}, "cuda:0")
    got = float(marker_closure(prob, pb, vidx.to(torch.int32).cuda(), lm))
    assert abs(got - float(loss_m)) <= 1e-5 * abs(float(loss_m))
    for k, ref in ref_m.items():
        g = pb.gviews[k].cpu().numpy().reshape(ref.shape)
        assert rel_err(g, ref) < 1e-4, k
        if k != "betas":
            assert frame_rel_err(g, ref) < 2e-4, k


def _sample(F, n=64):
    return np.unique(np.linspace(0, F - 1, n).round().astype(np.int64))


def test_config1_lbs_F4096_sampled(ctx):
    """lbs() forward + backward microbench at 4096 frames (32 m-tiles, 9 K-splits, > 148 work items)."""
    from uuo_mocap_b200 import synthetic
    model, smpl, orc = ctx["dense"]
    F = 4096
    mb = synthetic.make_lbs_microbench(F, seed=2)
    sel = _sample(F)
    rot = _T(mb["rot"]).cuda().requires_grad_(True)
    betas = _T(mb["betas"]).cuda().requires_grad_(True)
    transl = _T(mb["transl"]).cuda().requires_grad_(True)
    o = smpl(betas=betas, body_pose=rot[:, 1:], global_orient=rot[:, :1], transl=transl, pose2rot=False)
    torch.autograd.backward([o.vertices, o.joints], [_T(mb["g_vertices"]).cuda(), _T(mb["g_joints"]).cuda()])
    r_rot = _T(mb["rot"][sel]).requires_grad_(True)
    r_betas = _T(mb["betas"][sel]).requires_grad_(True)
    r_tr = _T(mb["transl"][sel]).requires_grad_(True)
    ro = orc.forward(r_betas, r_rot[:, 1:], r_rot[:, :1], r_tr, pose2rot=False)
    torch.autograd.backward([ro.vertices, ro.joints], [_T(mb["g_vertices"][sel]), _T(mb["g_joints"][sel])])
    st = torch.from_numpy(sel).cuda()
    assert frame_rel_err(o.vertices[st].detach().cpu().numpy(), ro.vertices.detach().numpy()) < 1e-5
    assert frame_rel_err(o.joints[st].detach().cpu().numpy(), ro.joints.detach().numpy()) < 1e-5
    for g, ref in ((rot.grad[st], r_rot.grad), (betas.grad[st], r_betas.grad), (transl.grad[st], r_tr.grad)):
        assert frame_rel_err(g.cpu().numpy(), ref.numpy()) < 1e-4
    # the frames that were NOT sampled are at least finite and of the same scale
    assert torch.isfinite(o.vertices).all() and torch.isfinite(rot.grad).all()


def test_config2_fit_F5000_sampled(ctx):
    """5k-frame fit (noise, ghosts, dropout): chamfer closure (vertices, indices, gradients) and marker closure on 64
    sampled frames; only per-frame terms are switched on so the sample's gradients do not depend on the other frames."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.fit import FitProblem, ParamBlock, chamfer_closure, marker_closure
    from oracle import closures, p3d_oracle as p3d
    model, smpl, orc = ctx["top4"]
    F = 5000
    cap = synthetic.make_capture(model, F, num_markers=41, seed=3, n_ghost=4, dropout_p=0.05, noise_std=0.002)
    mk, opose, obetas = _T(cap.markers), _T(cap.prior_pose_body), _T(cap.prior_betas)
    root, trans = _T(cap.prior_root_orient), _T(cap.prior_trans)
    sel = _sample(F)
    st = torch.from_numpy(sel).cuda()
    prob = FitProblem(smpl, mk, opose, obetas)
    pb = ParamBlock.build({"trans": trans, "z_angle": torch.zeros(F), "betas": obetas, "pose_body": opose}, "cuda:0")
    for rep in range(2):                  # the second evaluation starts from the first one's matches (hinted search)
        chamfer_closure(prob, pb, root.cuda(), {"full_chamfer": 10.0})
    ids = prob.twin.visit_ids.long()
    v_gpu = torch.empty((len(sel), 6890, 3), device="cuda")
    v_gpu[:, ids] = prob.vertices[st]
    v_gpu = v_gpu.cpu()
    pose, tr = opose[sel].clone().requires_grad_(True), trans[sel].clone().requires_grad_(True)
    z = torch.zeros(len(sel), 1, 1, requires_grad=True)
    loss = closures.chamfer_stage_loss(orc, mk[sel], pose, opose[sel], obetas, obetas, root[sel], tr, z, {"full_chamfer": 10.0})
    loss.backward()
    with torch.no_grad():
        v_ref = orc.smpl_inference(p3d.normalize_rot(pose), obetas.repeat(len(sel), 1), p3d.normalize_rot(root[sel]), tr)["vertices"]
    assert frame_rel_err(v_gpu.numpy(), v_ref.numpy()) < 1e-5
    _, idx_ref = p3d.knn_k1(mk[sel], v_gpu)                       # the oracle's argmin on the SAME vertices: bit-exact
    on = prob.mask[st].cpu() != 0
    assert torch.equal(prob.idx[st].cpu().long()[on], idx_ref[on])
    scale = float(on.sum()) / prob.wsum_global                   # sample's weight sum vs the whole capture's
    for k, ref in (("pose_body", pose.grad), ("trans", tr.grad), ("z_angle", z.grad.reshape(-1))):
        assert rel_err(pb.gviews[k][st].cpu().numpy(), ref.numpy() * scale) < 1e-4, k
    # marker stage: row-sparse forward / backward, "marker" term only (mean over ALL F*M)
    vidx = torch.from_numpy(np.resize(cap.marker_vertex_ids, mk.shape[1]).astype(np.int64))
    pbm = ParamBlock.build({"pose_body": opose, "betas": obetas, "root_orient": root, "trans": trans}, "cuda:0")
    marker_closure(prob, pbm, vidx.to(torch.int32).cuda(), {"marker": 1.0})
    pose, rt, tr = (t[sel].clone().requires_grad_(True) for t in (opose, root, trans))
    loss = closures.marker_stage_loss(orc, mk[sel], pose, opose[sel], obetas, obetas, rt, tr, vidx, {"marker": 1.0})
    loss.backward()
    scale = len(sel) / F
    for k, ref in (("pose_body", pose.grad), ("root_orient", rt.grad), ("trans", tr.grad)):
        g = pbm.gviews[k][st].cpu().numpy()
        assert rel_err(g, ref.numpy() * scale) < 1e-4, k
        assert frame_rel_err(g, ref.numpy() * scale) < 5e-4, k


def test_lbfgs_end_point_matches_oracle_lbfgs(ctx):
    """the CUDA closure under torch.optim.LBFGS (the reference's optimiser call, optimization.py:176-183)
    against the oracle closure under the same L-BFGS.  Trajectories are chaotic in the last bits, so the END POINT is
    compared: fitted marker RMSE within 0.5 mm, translations within 1 mm."""
    from uuo_mocap_b200 import fit, synthetic
    from oracle import closures, p3d_oracle as p3d
    model, smpl, orc = ctx["top4"]
    F, iters = 8, 30
    cap = synthetic.make_capture(model, F, num_markers=41, seed=33, n_ghost=0, dropout_p=0.0, prior_pose_noise=0.05,
                                 prior_trans_err=0.03, prior_yaw_err=0.1, permute=False)    # column m = marker m
    mk, opose, obetas = _T(cap.markers), _T(cap.prior_pose_body), _T(cap.prior_betas)
    cfg = fit.default_config(num_iters=iters, optimizer="lbfgs")
    lm = cfg["stages"]["marker"]["losses"]
    vidx = torch.from_numpy(np.resize(cap.marker_vertex_ids, mk.shape[1]).astype(np.int64))

    def rmse(pose, betas, root, trans):
        with torch.no_grad():
            v = orc.forward(betas, p3d.normalize_rot(pose), p3d.normalize_rot(root), trans, pose2rot=False).vertices
        d = torch.norm(mk - v[:, vidx], dim=-1)
        return float(torch.sqrt(((d - closures.MARKER_DISTANCE) ** 2).mean()))
    pose, betas, root, trans = (_T(a).clone().requires_grad_(True) for a in
                                (cap.prior_pose_body, cap.prior_betas, cap.prior_root_orient, cap.prior_trans))
    r0 = rmse(pose.detach(), betas.detach(), root.detach(), trans.detach())
    closures.run_lbfgs([pose, betas, root, trans],
                       lambda: closures.marker_stage_loss(orc, mk, pose, opose, betas, obetas, root, trans, vidx, lm), iters, 1.0)
    r_ref = rmse(pose.detach(), betas.detach(), root.detach(), trans.detach())
    c_pose, c_betas, c_root, c_trans = (_T(a).clone().cuda() for a in
                                        (cap.prior_pose_body, cap.prior_betas, cap.prior_root_orient, cap.prior_trans))
    fit.optim_markers(mk, c_pose, opose, c_betas, obetas, c_root, c_trans, vidx.to(torch.int32), torch.ones(F), smpl, cfg)
    r_gpu = rmse(c_pose.cpu(), c_betas.cpu(), c_root.cpu(), c_trans.cpu())
    assert r_ref < 0.5 * r0 and r_gpu < 0.5 * r0
    assert abs(r_gpu - r_ref) < 0.5e-3, (r_gpu, r_ref)
    assert float((c_trans.cpu() - trans.detach()).abs().max()) < 1e-3
