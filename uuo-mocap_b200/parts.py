"""Part fits: ``find_best_part_fits`` of /root/reference/src/video_mocap/markers/markers_utils.py:274-638 on the
CUDA path.

For every candidate sub-tree of the skeleton the reference fits a yaw angle [1,1,1], a per-frame translation
[F,3] and the betas [1,10] so that the vertices owned by the sub-tree's joints explain a subset of the markers
(:454-562: ``10 * chamfer(markers_subset -> vertices_subset) + reg_betas * mse(betas, o_betas)``), scores the fit
with the bidirectional chamfer distance (:566-579) and labels the markers from the best one (:581-597).

Two drivers.  ``fit_subtrees`` follows the reference candidate by candidate: every closure evaluation is the dense
CUDA lbs() forward/backward (``SMPL.forward`` is an autograd function over uuo_lbs_forward/backward) plus the CUDA
nearest-neighbour kernel (``losses.chamfer_distance``) under torch's L-BFGS (the reference's optimiser; each candidate
keeps its own line search).  ``fit_subtrees_batched`` (optimizer.name = "adam") fits ALL
candidates at once: the candidates are independent fits, so they become hypothesis groups stacked along the frame
axis of ONE problem on the kernels of the fitting loop -- dense forward on the visiting-order twin, the vertices a
candidate does not own moved away (uuo_mask_vertices), the box-pruned correspondence kernel with last iteration's
matches as hint, the row-sparse backward, one fused Adam step over every candidate's parameters, the whole
iteration replayed from a CUDA graph.  Out of scope: the "reproject"/foot/ground terms (they need the HMR camera).
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib, ops
from ._lib import check, ptr
from .fit import FitProblem, ParamBlock, _model_of, rot_z, run_adam
from .losses import chamfer_distance
from .smpl import POSE_ROTMAT


def sub_hierarchies(parents, num_bones: int) -> List[List[int]]:
    """All connected sub-trees with exactly ``num_bones`` joints (each listed root first, the rest ascending),
    the candidates of utils/smpl_utils.py:106-160, in the reference's order (tests/golden/sub_hierarchies.json was
    produced by the reference's own function).  Grown joint by joint from every root by adding a child of a member;
    a visited set keeps every connected rooted set once."""
    parents = np.asarray(parents.detach().cpu() if isinstance(parents, torch.Tensor) else parents).astype(np.int64)
    J = parents.shape[0]
    num_bones = min(int(num_bones), J)
    children = [[] for _ in range(J)]
    for j in range(1, J):
        children[int(parents[j])].append(j)
    found = set()

    def grow(members: frozenset, root: int):
        if len(members) == num_bones:
            found.add((root, members))
            return
        for m in members:
            for c in children[m]:
                if c not in members:
                    nxt = members | {c}
                    key = (root, nxt)
                    if key not in seen:
                        seen.add(key)
                        grow(nxt, root)

    for root in range(J):
        seen = set()
        grow(frozenset([root]), root)
    # the reference lists them root by root from the last joint down, and under one root in the order of an
    # itertools.product over the children's own sub-tree tables (last child varying fastest, "no joint of this
    # child" first): that is the lexicographic order of this recursive key
    def key(members, node):
        if node not in members:
            return (0,)
        return (1,) + tuple(key(members, c) for c in children[node])

    out = [(root, m) for root, m in found]
    out.sort(key=lambda rm: (-rm[0], key(rm[1], rm[0])))
    return [[root] + sorted(m - {root}) for root, m in out]


def fit_subtrees(markers_subset: torch.Tensor, pose_body: torch.Tensor, o_betas: torch.Tensor, root_orient: torch.Tensor,
                 smpl_inference, subtrees: Sequence[Sequence[int]], config: Dict,
                 iter_fn: Optional[Callable] = None) -> List[Dict]:
    """The loop of markers_utils.py:413-600 over candidate sub-trees.  Returns one dict per sub-tree with the
    fitted ``betas [1,NB]``, ``trans [F,3]``, ``root_orient [F,1,3,3]`` (yaw applied), ``z_angle``, the final
    ``loss`` and the bidirectional chamfer ``distance`` used for ranking."""
    smpl = _model_of(smpl_inference)
    dev = smpl.device
    st = config["stages"]["part"]
    losses = st["losses"]
    opt_cfg = config.get("optimizer", {})
    T = lambda t: t.detach().to(dev, torch.float32).contiguous()
    mk, pose, root, ob = T(markers_subset), T(pose_body), T(root_orient), T(o_betas).reshape(1, -1)
    F = mk.shape[0]
    vertex_labels = torch.argmax(smpl.lbs_weights, dim=-1)                                  # :438-439
    results = []
    for subtree in subtrees:
        z_angle = torch.zeros((1, 1, 1), device=dev, requires_grad=True)                    # :416
        trans = torch.median(mk, dim=1)[0].clone().requires_grad_(True)                     # :417
        betas = ob.clone().requires_grad_(True)                                             # :418
        vidx = torch.cat([(vertex_labels == int(j)).nonzero(as_tuple=True)[0] for j in subtree], dim=0)   # :441-444
        it = [0]

        def loss_fn():
            z_root = rot_z(z_angle.expand(F, 1, 1)) @ root                                  # :458
            out = smpl(betas=betas.expand(F, -1), body_pose=pose, global_orient=z_root, transl=trans, pose2rot=False)
            vs = out.vertices[:, vidx]
            loss = chamfer_distance(mk, vs, single_directional=True)[0] * float(losses["chamfer"])   # :471-476
            if "reg_betas" in losses:                                                       # :516-517
                loss = loss + torch.nn.functional.mse_loss(betas, ob) * float(losses["reg_betas"])
            return loss

        params = [z_angle, trans, betas]
        name = str(opt_cfg.get("name", "lbfgs")).lower()
        last = [None]
        if name == "lbfgs":                                                                 # :428-435
            opt = torch.optim.LBFGS(params, max_iter=int(st["num_iters"]),
                                    tolerance_grad=float(opt_cfg.get("tolerance_grad", 1e-7)),
                                    tolerance_change=float(opt_cfg.get("tolerance_change", 1e-9)), lr=1.0,
                                    line_search_fn="strong_wolfe")

            def closure():
                opt.zero_grad()
                loss = loss_fn()
                loss.backward()
                last[0] = loss.detach()
                if iter_fn is not None:
                    iter_fn(stage="part", iteration=it[0], part_joints=np.array(subtree), loss=float(loss))
                it[0] += 1
                return loss
            opt.step(closure)
        else:
            opt = torch.optim.Adam(params, lr=float(st.get("lr_adam", opt_cfg.get("lr", 0.01))))
            for _ in range(int(st["num_iters"])):
                opt.zero_grad()
                loss = loss_fn()
                loss.backward()
                opt.step()
                last[0] = loss.detach()
        with torch.no_grad():                                                               # :564-580
            z_root = rot_z(z_angle.expand(F, 1, 1)) @ root
            out = smpl(betas=betas.expand(F, -1), body_pose=pose, global_orient=z_root, transl=trans, pose2rot=False)
            vs = out.vertices[:, vidx].contiguous()
            distance = float(chamfer_distance(mk, vs, single_directional=False)[0])
        results.append({"subtree": list(int(j) for j in subtree), "distance": distance,
                        "loss": None if last[0] is None else float(last[0]), "betas": betas.detach().clone(),
                        "trans": trans.detach().clone(), "root_orient": z_root.detach().clone(),
                        "z_angle": z_angle.detach().clone(), "vertex_indices": vidx, "vertices": out.vertices.detach()})
    return results


def fit_subtrees_batched(markers_subset: torch.Tensor, pose_body: torch.Tensor, o_betas: torch.Tensor,
                         root_orient: torch.Tensor, smpl_inference, subtrees: Sequence[Sequence[int]], config: Dict,
                         max_groups: int = 64) -> List[Dict]:
    """All candidate sub-trees of markers_utils.py:413-600 fitted TOGETHER with the fused Adam optimiser: candidate g is
    hypothesis group g of one stacked problem of G x F frames.  Parameters per candidate as in the reference (:416-418):
    z_angle [1] (one yaw for the whole sequence), trans [F,3] (starts at the per-frame marker median), betas [1,10];
    loss per candidate = chamfer * mean_f mean_m min_{v in subset} |m - v|^2 + reg_betas * mse(betas, o_betas)
    (:463-476, 516-517; unweighted: a marker at the origin counts like any other).  Returns the dicts of
    ``fit_subtrees`` (the ranking distance is evaluated per candidate after the fit)."""
    smpl = _model_of(smpl_inference)
    dev = smpl.device
    st = config["stages"]["part"]
    losses = st["losses"]
    opt_cfg = config.get("optimizer", {})
    T = lambda t: t.detach().to(dev, torch.float32).contiguous()
    mk, pose, root, ob = T(markers_subset), T(pose_body), T(root_orient), T(o_betas).reshape(1, -1)
    F, Ms = mk.shape[:2]
    NB = ob.shape[1]
    twin = smpl.visiting_twin()
    V = smpl.handle.V
    vertex_labels = torch.argmax(smpl.lbs_weights, dim=-1)                                  # :438-439
    results: List[Dict] = []
    for g0 in range(0, len(subtrees), max_groups):
        chunk = [list(int(j) for j in s) for s in subtrees[g0:g0 + max_groups]]
        G = len(chunk)
        keep_orig = torch.stack([torch.isin(vertex_labels, torch.tensor(s, device=dev)) for s in chunk])   # [G,V] :441-444
        keep_rows = keep_orig[:, twin.visit_ids.long()].to(torch.uint8).contiguous()        # row numbering of the twin
        rep = lambda t: t.repeat((G,) + (1,) * (t.dim() - 1)).contiguous()
        mk_s, pose_s, root_s = rep(mk), rep(pose), rep(root)
        prob = FitProblem(smpl, mk_s, pose_s, ob, None, groups=G)
        pb = ParamBlock.build({"trans": rep(torch.median(mk, dim=1)[0]), "z_angle": torch.zeros(G),
                               "betas": ob.repeat(G, 1)}, dev)
        g_yaw = torch.zeros(G * F, dtype=torch.float32, device=dev)
        g_pose = torch.zeros_like(pose_s)                                                    # pose is not a parameter here
        gb_rows = torch.zeros((G, NB), dtype=torch.float32, device=dev)                     # one d/d betas row per candidate
        w_ch, w_reg = float(losses["chamfer"]), float(losses.get("reg_betas", 0.0))
        ob_g = ob.repeat(G, 1).contiguous()
        L = _lib.lib()

        def closure():
            v, g = pb.views, pb.gviews
            pb.grad.zero_()
            prob.red.zero_()
            yaw_rows = torch.repeat_interleave(v["z_angle"], F).contiguous()
            brows = v["betas"]                                                               # betas_rows = G: row f // F
            with torch.cuda.device(dev):
                if prob.vertices is None:
                    prob.vertices = torch.empty((G * F, V, 3), dtype=torch.float32, device=dev)
                check(L.uuo_lbs_forward(twin.handle.handle, ptr(brows), G, ptr(root_s), ptr(pose_s), POSE_ROTMAT,
                                        ptr(yaw_rows), ptr(v["trans"]), G * F, ptr(prob.vertices), None, ptr(prob.ws),
                                        prob._st()))                                         # :458-466 (pose used as it is)
                check(L.uuo_mask_vertices(ptr(prob.vertices), ptr(keep_rows), G * F, V, F, 1.0e18, prob._st()))
                check(L.uuo_nn_argmin_presorted_fit(ptr(mk_s), ptr(prob.vertices), ptr(twin.visit_ids), None, ptr(prob.pos),
                                                    G * F, Ms, V, ptr(prob.d2), ptr(prob.idx), ptr(prob.pos), prob._st()))
                check(L.uuo_chamfer_rows(ptr(mk_s), ptr(prob.vertices), None, ptr(prob.pos), G * F, Ms, V,
                                         w_ch / (F * Ms), ptr(prob.loss_slot), ptr(prob.g_rows), prob._st()))   # :471-476
                fn = L.uuo_lbs_backward_rows
                if Ms <= 64:             # v_posed of the matched rows straight out of the dense forward's workspace
                    check(L.uuo_lbs_cache_vposed_rows(twin.handle.handle, G * F, ptr(prob.pos), Ms, ptr(prob.ws), prob._st()))
                    fn = L.uuo_lbs_backward_rows_cached
                check(fn(smpl.handle.handle, ptr(brows), G, ptr(root_s), ptr(pose_s), POSE_ROTMAT, ptr(yaw_rows), G * F,
                         ptr(prob.idx), 0, Ms, ptr(prob.g_rows), None, ptr(prob.ws), ptr(gb_rows), None, ptr(g_pose),
                         ptr(g_yaw), ptr(g["trans"]), prob._st()))
            g["z_angle"].copy_(g_yaw.view(G, F).sum(1))                                      # one yaw per candidate
            g["betas"].copy_(gb_rows)
            if w_reg != 0.0:                                                                 # :516-517
                prob.sqdiff(v["betas"], ob_g, w_reg / NB, g["betas"])
            return prob.loss_slot[0]

        iters = int(st["num_iters"])
        if iters > 0:
            run_adam(pb, closure, iters, float(st.get("lr_adam", opt_cfg.get("lr", 0.01))))
        with torch.no_grad():
            for gi, subtree in enumerate(chunk):
                z = pb.views["z_angle"][gi].reshape(1, 1, 1).clone()
                trans = pb.views["trans"][gi * F:(gi + 1) * F].clone()
                betas = pb.views["betas"][gi:gi + 1].clone()
                vidx = keep_orig[gi].nonzero(as_tuple=True)[0]
                z_root = rot_z(z.expand(F, 1, 1)) @ root
                out = smpl(betas=betas.expand(F, -1), body_pose=pose, global_orient=z_root, transl=trans, pose2rot=False)
                vs = out.vertices[:, vidx].contiguous()
                distance = float(chamfer_distance(mk, vs, single_directional=False)[0])     # :564-580
                results.append({"subtree": subtree, "distance": distance, "loss": None, "betas": betas, "trans": trans,
                                "root_orient": z_root.clone(), "z_angle": z, "vertex_indices": vidx,
                                "vertices": out.vertices.detach()})
    return results


def find_best_part_fits(markers: torch.Tensor, pose_body: torch.Tensor, betas: torch.Tensor, root_orient: torch.Tensor,
                        marker_labels: torch.Tensor, smpl_inference, hierarchy, config: Dict,
                        subtrees: Optional[Sequence[Sequence[int]]] = None, iter_fn: Optional[Callable] = None) -> Dict:
    """markers_utils.py:274-638 for the "cluster" mode and ``use_full_skeleton`` (the shipped video_mocap.yaml
    settings): the markers of the largest label group are fitted by every candidate sub-tree; returns the
    reference's output dict (betas, marker_labels, markers_subset, marker_weights, root_orient, trans,
    aabb_volume_ratio, chain).  The camera arguments of the reference feed only its "reproject" term and are
    not taken."""
    smpl = _model_of(smpl_inference)
    dev = smpl.device
    st = config["stages"]["part"]
    if "reproject" in st["losses"]:
        raise NotImplementedError("the reproject term needs the HMR camera inputs (outside the fitting hot path)")
    markers = markers.to(dev, torch.float32)
    marker_labels = marker_labels.to(dev)
    labels_mode = torch.mode(marker_labels, dim=0)[0]                                       # :294
    chain = torch.unique(labels_mode).tolist()                                              # cluster mode, :341-342
    indices = torch.cat([torch.where(labels_mode == j)[0] for j in chain], dim=0)           # :383-386
    markers_subset = markers[:, indices].contiguous()
    parents = np.asarray(hierarchy.detach().cpu() if isinstance(hierarchy, torch.Tensor) else hierarchy)
    if subtrees is None:
        if st.get("use_full_skeleton", False):
            subtrees = [list(range(parents.shape[0]))]                                      # :396-397
        else:
            subtrees = sub_hierarchies(parents, len(chain))                                 # :399
    batched = str(config.get("optimizer", {}).get("name", "lbfgs")).lower() == "adam" and iter_fn is None
    if batched:      # independent fits -> hypothesis groups of one stacked problem (fused Adam, graph replay)
        fits = fit_subtrees_batched(markers_subset, pose_body, betas, root_orient, smpl, subtrees, config)
    else:            # the reference's optimiser: one L-BFGS (own line search, own stopping test) per candidate
        fits = fit_subtrees(markers_subset, pose_body, betas, root_orient, smpl, subtrees, config, iter_fn)
    ranked = sorted(fits, key=lambda r: r["distance"])
    best = ranked[0]
    final_labels = torch.zeros_like(marker_labels)
    final_weights = torch.zeros_like(marker_labels, dtype=torch.float32)
    vertex_labels = torch.argmax(smpl.lbs_weights, dim=-1)
    # labels: vertex with the smallest frame-averaged distance to each marker (:585-589) -- the mean-distance kernel
    vi, _ = ops.mean_distance_argmin(markers_subset, best["vertices"].contiguous())
    final_labels[:, indices] = vertex_labels[vi.long()].to(final_labels.dtype).unsqueeze(0)
    if len(ranked) > 1:                                                                     # :603-609
        final_weights[:, indices] = ranked[1]["distance"] / ranked[0]["distance"]
        if indices.shape[0] == 1:
            final_weights *= 0
    mx = torch.max(final_weights)
    final_weights = final_weights / mx if float(mx) > 0 else final_weights                  # :617

    def aabb_volume(p):                       # utils/aabb.py:5-25: per-frame box volume [F]
        return torch.prod(p.max(1)[0] - p.min(1)[0], dim=-1)
    return {"betas": best["betas"], "marker_labels": final_labels, "markers_subset": markers_subset,
            "marker_weights": final_weights, "root_orient": best["root_orient"], "trans": best["trans"],
            "aabb_volume_ratio": aabb_volume(markers_subset) / aabb_volume(markers),
            "chain": np.array(best["subtree"], dtype=np.int32), "ranking": [(r["subtree"], r["distance"]) for r in ranked], "fits": fits}
