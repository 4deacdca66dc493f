"""Part fits: differentiable chamfer on the CUDA correspondence kernel against
the oracle's pytorch3d restatement, chamfer_distance_by_part, the closure of closure_fit_subtree and a small
find_best_part_fits run that must rank the true sub-tree first."""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.parametrize("N,P1,P2", [(5, 7, 19), (3, 41, 700), (2, 300, 45), (1, 1, 1)])
@pytest.mark.parametrize("single", [True, False])
@pytest.mark.parametrize("weighted", [False, True])
def test_chamfer_distance_matches_oracle(N, P1, P2, single, weighted):
    from uuo_mocap_b200 import losses
    from oracle import p3d_oracle as p3d
    rng = np.random.default_rng(N * 100 + P1)
    x0 = rng.normal(0, 0.5, size=(N, P1, 3)).astype(np.float32)
    y0 = rng.normal(0, 0.5, size=(N, P2, 3)).astype(np.float32)
    w0 = rng.uniform(0, 1, size=(N,)).astype(np.float32) if weighted else None
    res = []
    for dev, fn in (("cpu", p3d.chamfer_distance), ("cuda:0", losses.chamfer_distance)):
        x = _T(x0).to(dev).requires_grad_(True)
        y = _T(y0).to(dev).requires_grad_(True)
        w = _T(w0).to(dev) if weighted else None
        val, _ = fn(x, y, weights=w, single_directional=single)
        val.backward()
        res.append((float(val), x.grad.cpu().numpy(), y.grad.cpu().numpy()))
    assert abs(res[1][0] - res[0][0]) <= 1e-6 * abs(res[0][0])
    assert rel_err(res[1][1], res[0][1]) < 1e-5
    assert rel_err(res[1][2], res[0][2]) < 1e-5


def test_chamfer_distance_zero_weights_and_errors():
    from uuo_mocap_b200 import losses
    x = torch.randn(3, 4, 3, device="cuda:0", requires_grad=True)
    y = torch.randn(3, 9, 3, device="cuda:0")
    val, _ = losses.chamfer_distance(x, y, weights=torch.zeros(3, device="cuda:0"))
    assert float(val) == 0.0
    with pytest.raises(ValueError):
        losses.chamfer_distance(x, y, weights=torch.tensor([1.0, -1.0, 1.0], device="cuda:0"))
    with pytest.raises(ValueError):
        losses.chamfer_distance(x, y, weights=torch.ones(2, device="cuda:0"))
    with pytest.raises(Exception):
        losses.chamfer_distance(x.cpu(), y.cpu())


@pytest.fixture(scope="module")
def ctx(get_model):
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    from oracle.smpl_oracle import OracleSMPL
    m = get_model("top4")
    return m, SMPL(_as_struct(m), device="cuda:0"), OracleSMPL.from_model(m)


def _case(model, F=6, seed=21):
    from uuo_mocap_b200 import synthetic
    cap = synthetic.make_capture(model, F, num_markers=41, seed=seed, noise_std=0.001)
    rot = synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)
    return cap, rot


def test_chamfer_distance_by_part(ctx):
    """optimization.py:682-700 on the dense CUDA path: value 1e-5, gradients 1e-4 vs the oracle."""
    from uuo_mocap_b200 import losses
    from oracle import closures
    model, smpl, orc = ctx
    cap, rot = _case(model)
    F = rot.shape[0]
    labels = np.argmax(model.weights, axis=-1)[cap.marker_vertex_ids]          # joint owning each marker's vertex
    lab = np.repeat(labels[None], F, 0)
    # same capture without the per-frame permutation: column m is marker m in every frame (labels are per column)
    from uuo_mocap_b200 import synthetic
    cap2 = synthetic.make_capture(model, F, num_markers=41, seed=21, noise_std=0.001, permute=False)
    mk = cap2.markers
    out = []
    for dev in ("cpu", "cuda:0"):
        betas = _T(cap.gt_betas).to(dev).requires_grad_(True)
        tr = _T(cap.gt_trans).to(dev).requires_grad_(True)
        r = _T(rot).to(dev)
        if dev == "cpu":
            verts = orc.forward(betas.expand(F, -1), r[:, 1:], r[:, :1], tr, pose2rot=False).vertices
            val = closures.chamfer_distance_by_part(_T(mk), verts, _T(lab), _T(model.weights))
        else:
            verts = smpl(betas=betas.expand(F, -1), body_pose=r[:, 1:], global_orient=r[:, :1], transl=tr,
                         pose2rot=False).vertices
            val = losses.chamfer_distance_by_part(_T(mk).cuda(), verts, _T(lab).cuda(), smpl.lbs_weights)
        val.backward()
        out.append((float(val), betas.grad.cpu().numpy(), tr.grad.cpu().numpy()))
    assert abs(out[1][0] - out[0][0]) <= 1e-5 * abs(out[0][0])
    assert rel_err(out[1][1], out[0][1]) < 1e-4
    assert rel_err(out[1][2], out[0][2]) < 1e-4


def test_part_fit_closure_matches_oracle(ctx):
    """closure_fit_subtree (markers_utils.py:454-522): loss 1e-5, d z_angle / d trans / d betas 1e-4."""
    from uuo_mocap_b200 import losses
    from uuo_mocap_b200.fit import rot_z
    from oracle import closures
    model, smpl, orc = ctx
    cap, rot = _case(model, F=5, seed=22)
    F = rot.shape[0]
    subtree = [16, 18, 20]
    vlab = np.argmax(model.weights, axis=-1)
    vidx = np.concatenate([np.where(vlab == j)[0] for j in subtree])
    sel = np.where(np.isin(vlab[cap.marker_vertex_ids], subtree))[0]
    from uuo_mocap_b200 import synthetic
    mk = synthetic.make_capture(model, F, num_markers=41, seed=22, noise_std=0.001, permute=False).markers[:, sel]
    w = {"chamfer": 10.0, "reg_betas": 0.1}
    out = []
    for dev in ("cpu", "cuda:0"):
        z = torch.full((1, 1, 1), 0.2, device=dev, requires_grad=True)
        tr = (_T(cap.gt_trans) + 0.03).to(dev).requires_grad_(True)
        betas = (_T(cap.gt_betas) * 0.5).to(dev).requires_grad_(True)
        ob = _T(cap.prior_betas).to(dev)
        r = _T(rot).to(dev)
        if dev == "cpu":
            loss = closures.part_fit_loss(orc, _T(mk), r[:, 1:], betas, ob, r[:, :1], tr, z, _T(vidx), w)
        else:
            z_root = rot_z(z.expand(F, 1, 1)) @ r[:, :1]
            o = smpl(betas=betas.expand(F, -1), body_pose=r[:, 1:], global_orient=z_root, transl=tr, pose2rot=False)
            loss = losses.chamfer_distance(_T(mk).cuda(), o.vertices[:, _T(vidx).cuda()], single_directional=True)[0] * w["chamfer"]
            loss = loss + torch.nn.functional.mse_loss(betas, ob) * w["reg_betas"]
        loss.backward()
        out.append((float(loss), z.grad.cpu().numpy(), tr.grad.cpu().numpy(), betas.grad.cpu().numpy()))
    assert abs(out[1][0] - out[0][0]) <= 1e-5 * abs(out[0][0])
    for k in (1, 2, 3):
        assert rel_err(out[1][k], out[0][k]) < 1e-4, k


def test_find_best_part_fits(ctx):
    """markers_utils.py:413-620 end to end on a small case (L-BFGS like the reference): every candidate's fit lowers
    its loss, its score equals the oracle's bidirectional chamfer evaluated at the fitted parameters, the winner is
    the best score, labels / weights / shapes follow the reference's output dict."""
    from uuo_mocap_b200 import ops, parts, synthetic
    from oracle import p3d_oracle as p3d
    model, smpl, orc = ctx
    F = 8
    cap, rot = _case(model, F=F, seed=23)
    vlab = np.argmax(model.weights, axis=-1)
    true = [16, 18, 20]                                  # left shoulder, elbow, wrist
    rng = np.random.default_rng(5)
    vids = np.concatenate([rng.choice(np.where(vlab == j)[0], size=6, replace=False) for j in true])
    verts = synthetic.lbs_numpy(model, cap.gt_betas, rot.astype(np.float64), cap.gt_trans.astype(np.float64), vertex_ids=vids)
    markers = _T(verts.astype(np.float32))
    labels = torch.from_numpy(np.repeat(vlab[vids][None], F, 0).astype(np.int64))
    cfg = {"optimizer": {"name": "lbfgs", "tolerance_grad": 1e-7, "tolerance_change": 1e-9},
           "stages": {"part": {"mode": "cluster", "num_iters": 40, "losses": {"chamfer": 10.0, "reg_betas": 0.1}}}}
    cands = [[16, 18, 20], [17, 19, 21], [1, 4, 7]]
    r = _T(rot)
    seen = {}

    def iter_fn(stage, iteration, part_joints, loss):
        seen.setdefault(tuple(part_joints.tolist()), []).append(loss)
    out = parts.find_best_part_fits(markers, r[:, 1:], _T(cap.gt_betas), r[:, :1], labels, smpl, model.parents, cfg,
                                    subtrees=cands, iter_fn=iter_fn)
    fits = out["fits"]
    for fit in fits:
        hist = seen[tuple(fit["subtree"])]
        assert len(hist) > 3 and min(hist) < 0.5 * hist[0]
        o = orc.forward(fit["betas"].cpu().expand(F, -1), r[:, 1:], fit["root_orient"].cpu(), fit["trans"].cpu(), pose2rot=False)
        ref = float(p3d.chamfer_distance(markers, o.vertices[:, fit["vertex_indices"].cpu()], single_directional=False)[0])
        assert abs(fit["distance"] - ref) <= 1e-4 * abs(ref), fit["subtree"]
    rank = out["ranking"]
    assert [d for _, d in rank] == sorted(d for _, d in rank)
    assert out["chain"].tolist() == rank[0][0]
    best = [f for f in fits if f["subtree"] == rank[0][0]][0]
    vi, _ = ops.mean_distance_argmin(markers.cuda(), best["vertices"])
    assert torch.equal(out["marker_labels"][0].cpu(), torch.from_numpy(vlab)[vi.cpu().long()])
    assert torch.all(out["marker_weights"] == 1.0)       # every marker is in the chain: rank[1]/rank[0], rescaled by its max
    assert out["trans"].shape == (F, 3) and out["betas"].shape == (1, 10) and out["root_orient"].shape == (F, 1, 3, 3)
    assert out["aabb_volume_ratio"].shape == (F,) and torch.allclose(out["aabb_volume_ratio"].cpu(), torch.ones(F))
    assert out["markers_subset"].shape == markers.shape


def test_part_fits_match_reference_source(ctx, golden):
    """find_best_part_fits / closure_fit_subtree against the outputs of the REFERENCE'S OWN SOURCE
    (markers_utils.py:274-638 compiled unmodified, oracle/gen_golden_ref_closures.py): (A) every candidate's closure at
    z = 0.2 -- loss 1e-5, gradients 1e-4; (B) the function run with no optimiser step: same candidate order, same
    three best scores (printed by the reference with 6 decimals), same winner, labels, weights and box ratio."""
    from uuo_mocap_b200 import losses, parts
    from uuo_mocap_b200.fit import rot_z
    model, smpl, _ = ctx
    r = golden("closure_ref.npz")
    vlab = np.argmax(model.weights, axis=-1)
    rot, mk = _T(r["pf_rot"]).cuda(), _T(r["pf_markers"]).cuda()
    F = mk.shape[0]
    assert parts.sub_hierarchies(np.array(model.parents), 3) == r["pf_subtrees"].tolist()
    ob = _T(r["pf_betas"]).cuda()
    for s, subtree in enumerate(r["pf_subtrees"]):
        vidx = torch.from_numpy(np.concatenate([np.where(vlab == j)[0] for j in subtree])).cuda()
        z = torch.full((1, 1, 1), 0.2, device="cuda", requires_grad=True)
        tr = torch.median(mk, dim=1)[0].clone().requires_grad_(True)
        betas = ob.clone().requires_grad_(True)
        o = smpl(betas=betas.expand(F, -1), body_pose=rot[:, 1:], global_orient=rot_z(z.expand(F, 1, 1)) @ rot[:, :1], transl=tr,
                 pose2rot=False)
        loss = losses.chamfer_distance(mk, o.vertices[:, vidx], single_directional=True)[0] * 10.0
        loss = loss + torch.nn.functional.mse_loss(betas, ob) * 0.1
        loss.backward()
        assert abs(float(loss) - float(r["pfA_loss"][s])) <= 1e-5 * abs(float(r["pfA_loss"][s])), subtree
        assert rel_err(z.grad.cpu().numpy().reshape(-1), r["pfA_g_z"][s]) < 1e-4
        assert rel_err(tr.grad.cpu().numpy(), r["pfA_g_trans"][s]) < 1e-4
        assert rel_err(betas.grad.cpu().numpy(), r["pfA_g_betas"][s]) < 1e-4
    cfg = {"optimizer": {"name": "lbfgs", "tolerance_grad": 1e-7, "tolerance_change": 1e-9},
           "stages": {"part": {"mode": "cluster", "num_iters": 0, "losses": {"chamfer": 10.0, "reg_betas": 0.1}}}}
    out = parts.find_best_part_fits(mk, rot[:, 1:], ob, rot[:, :1], torch.from_numpy(r["pf_labels"]), smpl, model.parents, cfg)
    rank = out["ranking"]
    assert [list(j) for j, _ in rank[:3]] == r["pfB_top3_joints"].tolist()
    for (_, d), ref in zip(rank[:3], r["pfB_top3_dist"]):
        assert abs(d - ref) <= 1e-6 + 1e-5 * ref
    assert out["chain"].tolist() == r["pfB_chain"].tolist()
    assert np.array_equal(out["marker_labels"].cpu().numpy(), r["pfB_marker_labels"])
    assert rel_err(out["marker_weights"].cpu().numpy(), r["pfB_marker_weights"]) < 1e-6
    assert rel_err(out["aabb_volume_ratio"].cpu().numpy(), r["pfB_aabb"]) < 1e-6
    assert rel_err(out["trans"].cpu().numpy(), r["pfB_trans"]) < 1e-6
    assert rel_err(out["root_orient"].cpu().numpy(), r["pfB_root"]) < 1e-6


def test_batched_part_fits_equal_sequential_adam(ctx):
    """the candidates of find_best_part_fits as hypothesis groups of ONE stacked problem
    (fit_subtrees_batched: fused Adam, CUDA-graph replay) against the same candidates fitted one after the other with
    torch.optim.Adam over the autograd path (fit_subtrees): same parameters after a fixed iteration count, same scores,
    same ranking, same output dict."""
    from uuo_mocap_b200 import parts, synthetic
    model, smpl, _ = ctx
    F, iters = 8, 25
    cap, rot = _case(model, F=F, seed=23)
    vlab = np.argmax(model.weights, axis=-1)
    rng = np.random.default_rng(5)
    vids = np.concatenate([rng.choice(np.where(vlab == j)[0], size=6, replace=False) for j in (16, 18, 20)])
    verts = synthetic.lbs_numpy(model, cap.gt_betas, rot.astype(np.float64), cap.gt_trans.astype(np.float64), vertex_ids=vids)
    markers = _T(verts.astype(np.float32))
    markers[2, 3] = 0.0                                   # a dropped marker counts like any other here (unweighted chamfer)
    labels = torch.from_numpy(np.repeat(vlab[vids][None], F, 0).astype(np.int64))
    cfg = {"optimizer": {"name": "adam", "lr": 0.01},
           "stages": {"part": {"mode": "cluster", "num_iters": iters, "losses": {"chamfer": 10.0, "reg_betas": 0.1}}}}
    cands = [[16, 18, 20], [17, 19, 21], [1, 4, 7], [0, 1, 2]]
    r = _T(rot)
    b0 = _T(cap.gt_betas * 0.5)
    seq = parts.fit_subtrees(markers, r[:, 1:], b0, r[:, :1], smpl, cands, cfg)
    bat = parts.fit_subtrees_batched(markers, r[:, 1:], b0, r[:, :1], smpl, cands, cfg)
    for a, b in zip(seq, bat):
        assert a["subtree"] == b["subtree"]
        assert rel_err(b["trans"].cpu().numpy(), a["trans"].cpu().numpy()) < 1e-4
        assert rel_err(b["betas"].cpu().numpy(), a["betas"].cpu().numpy()) < 1e-3
        assert abs(float(b["z_angle"]) - float(a["z_angle"])) < 1e-4
        assert abs(b["distance"] - a["distance"]) <= 1e-3 * abs(a["distance"])
    # a two-candidate chunking gives the same fits (groups are independent)
    bat2 = parts.fit_subtrees_batched(markers, r[:, 1:], b0, r[:, :1], smpl, cands, cfg, max_groups=2)
    for a, b in zip(bat, bat2):
        assert rel_err(b["trans"].cpu().numpy(), a["trans"].cpu().numpy()) < 1e-5
    out = parts.find_best_part_fits(markers, r[:, 1:], b0, r[:, :1], labels, smpl, model.parents, cfg, subtrees=cands)
    assert [s for s, _ in out["ranking"]] == [f["subtree"] for f in sorted(seq, key=lambda f: f["distance"])]
    assert out["chain"].tolist() == out["ranking"][0][0]
