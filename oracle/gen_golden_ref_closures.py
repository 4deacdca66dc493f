"""Golden vectors of the stage closures produced by the REFERENCE'S OWN SOURCE (TEST INFRASTRUCTURE; build container only).

    python oracle/gen_golden_ref_closures.py      (needs /root/reference; writes tests/golden/closure_ref.npz)

``video_mocap`` cannot be imported here (pytorch3d, igl, trimesh, roma ... are not installable offline), so the
function / class definitions are taken out of the reference files with ``ast`` and compiled UNMODIFIED:
  src/video_mocap/optimization.py     optim_chamfer, optim_markers, compute_nearest_points, compute_root_orient_z,
                                      compute_root_orient_y, chamfer_distance_by_part, get_marker_mask, weighted_mse_loss
  src/video_mocap/losses/chamfer_distance.py   weighted_chamfer_distance
  src/video_mocap/losses/losses.py    MarkerLoss
  src/video_mocap/utils/points.py     closest_point, geometric_median
  src/video_mocap/markers/markers_utils.py   find_best_part_fits (with closure_fit_subtree inside it)
  src/video_mocap/utils/smpl_utils.py  get_joint_id, get_joint_name, get_sub_hierachies; utils/aabb.py  get_aabb[_volume]
  src/video_mocap/utils/smpl.py       SmplInference (its forward; the instance is built around the vendored smplx.SMPL
                                      on the synthetic data_struct instead of the licensed .pkl)
What is bound from outside: the pytorch3d names (-> oracle/p3d_oracle.py, the one library that is absent: "parity
unpinned" stays true for those ops only), ``trimesh.Trimesh`` (a stub: the reference builds one per frame and never
uses it on the use_mean branch), ``pdb.set_trace`` (a no-op: the reference stops in the debugger after the "temporal"
term, optimization.py:375), and ``torch.optim.LBFGS`` (a stub that evaluates the closure ONCE at chosen parameters and
records the loss and every parameter's .grad -- the closure is the parity contract).
The inputs are the ones of tests/golden/closure_top4.npz, so oracle/closures.py and the CUDA closures can be checked
against both files on identical data."""
import ast
import math
import os
import pdb
import sys
import types
from collections.abc import Callable
from typing import Dict, List

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import uuo_mocap_b200  # noqa: E402,F401
from uuo_mocap_b200 import synthetic  # noqa: E402
from oracle import p3d_oracle as p3d  # noqa: E402
from oracle.gen_golden import reference_smpl  # noqa: E402

REF = "/root/reference/src/video_mocap"
GOLD = os.path.join(ROOT, "tests", "golden")


def take(path, names):
    """The named top-level def / class nodes of a reference file, compiled as they are."""
    tree = ast.parse(open(path).read())
    nodes = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    missing = set(names) - {n.name for n in nodes}
    assert not missing, (path, missing)
    return compile(ast.Module(body=nodes, type_ignores=[]), path, "exec")


class OneEvaluation:
    """Stand-in for torch.optim.LBFGS: step(closure) evaluates the closure once, at `at` when given."""
    last = None
    at = None          # {param index: tensor} copied into the parameters before the evaluation
    history = []       # every instance, in creation order

    def __init__(self, params, **kw):
        self.params = list(params)
        self.kw = kw
        OneEvaluation.last = self
        OneEvaluation.history.append(self)

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    def step(self, closure):
        if OneEvaluation.at:
            with torch.no_grad():
                for i, v in OneEvaluation.at.items():
                    self.params[i].copy_(v.reshape(self.params[i].shape))
        self.loss = float(closure())
        self.grads = [None if p.grad is None else p.grad.detach().clone() for p in self.params]


def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    model = synthetic.make_model("top4")
    g = np.load(os.path.join(GOLD, "closure_top4.npz"))
    T = lambda k: torch.from_numpy(np.ascontiguousarray(g[k])).clone()

    # ---- namespace the reference functions run in
    trimesh = types.SimpleNamespace(Trimesh=lambda **kw: None)
    ns = {"torch": torch, "np": np, "F": F, "nn": nn, "math": math, "Dict": Dict, "List": List, "Callable": Callable,
          "trimesh": trimesh, "igl": None, "SDF": None, "MARKER_DISTANCE": 0.0095,           # utils/settings.py:1
          "chamfer_distance": p3d.chamfer_distance, "rotation_6d_to_matrix": p3d.rotation_6d_to_matrix,
          "matrix_to_rotation_6d": p3d.matrix_to_rotation_6d, "so3_relative_angle": p3d.so3_relative_angle,
          "axis_angle_to_matrix": p3d.axis_angle_to_matrix, "smplx": None}
    assert abs(float(open(REF + "/utils/settings.py").read().split("MARKER_DISTANCE")[1].split("=")[1].split()[0]) - 0.0095) < 1e-12
    exec(take(REF + "/utils/smpl.py", ["SmplInference"]), ns)
    exec(take(REF + "/losses/chamfer_distance.py", ["weighted_chamfer_distance"]), ns)
    exec(take(REF + "/losses/losses.py", ["MarkerLoss"]), ns)
    exec(take(REF + "/utils/points.py", ["closest_point", "geometric_median"]), ns)
    exec(take(REF + "/optimization.py", ["optim_chamfer", "optim_markers", "compute_nearest_points", "compute_root_orient_z",
                                         "compute_root_orient_y", "chamfer_distance_by_part", "get_marker_mask",
                                         "weighted_mse_loss"]), ns)
    # the reference's SmplInference.forward around the vendored smplx layer (no licensed .pkl: data_struct path)
    SmplInference = ns["SmplInference"]
    si = SmplInference.__new__(SmplInference)
    nn.Module.__init__(si)
    si.device = torch.device("cpu")
    si.smpl = reference_smpl(model)
    ns["SmplInference"] = SmplInference

    real_lbfgs, real_trace = torch.optim.LBFGS, pdb.set_trace
    torch.optim.LBFGS, pdb.set_trace = OneEvaluation, (lambda *a, **k: None)
    out = {"model_checksum": np.float64(model.checksum())}
    try:
        cfg = {"optimizer": {"tolerance_grad": 1e-7, "tolerance_change": 1e-9},
               "stages": {"chamfer": {"num_iters": 1, "yaw_lock": True, "single_directional": True,
                                      "losses": {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0, "trans_vel": 1.0}},
                          "marker": {"num_iters": 1, "use_sdf": False,
                                     "losses": {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0, "temporal": 1.0}},
                          "compute_locations": {"use_velocity": False, "use_barycentric": False, "use_mean": True},
                          "segment": {"granularity": "full"}}}
        markers = T("markers")
        Fn, M = markers.shape[:2]
        # -- optim_chamfer (optimization.py:147-285): params = [trans, z_angle, betas, pose_body]
        pose, betas, root, trans = (T(k) for k in ("ch_pose", "ch_betas", "ch_root", "ch_trans"))
        for t in (pose, betas, trans):
            t.requires_grad_(True)
        OneEvaluation.at = {1: T("ch_z")}
        ns["optim_chamfer"](markers, pose, T("ch_o_pose"), betas, T("ch_o_betas"), root, trans, torch.ones(Fn), None, si, cfg)
        ev = OneEvaluation.last
        assert ev.kw["lr"] == 0.1 and ev.kw["line_search_fn"] == "strong_wolfe"
        out.update(ch_loss=np.float32(ev.loss), ch_g_trans=ev.grads[0].numpy(), ch_g_z=ev.grads[1].numpy(),
                   ch_g_betas=ev.grads[2].numpy(), ch_g_pose=ev.grads[3].numpy(), ch_root_baked=root.detach().numpy())
        # the same closure without the optional trans_vel term (the shipped yaml)
        cfg2 = {**cfg, "stages": {**cfg["stages"], "chamfer": {**cfg["stages"]["chamfer"],
                                                                "losses": {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0}}}}
        pose, betas, root, trans = (T(k) for k in ("ch_pose", "ch_betas", "ch_root", "ch_trans"))
        for t in (pose, betas, trans):
            t.requires_grad_(True)
        ns["optim_chamfer"](markers, pose, T("ch_o_pose"), betas, T("ch_o_betas"), root, trans, torch.ones(Fn), None, si, cfg2)
        ev = OneEvaluation.last
        out.update(ch0_loss=np.float32(ev.loss), ch0_g_trans=ev.grads[0].numpy(), ch0_g_z=ev.grads[1].numpy(),
                   ch0_g_betas=ev.grads[2].numpy(), ch0_g_pose=ev.grads[3].numpy())

        # -- optim_chamfer with yaw_lock: False (:167-172, :194, :282-283): the free 3x3 "z_angle" IS the root inside the closure
        cfg3 = {**cfg, "stages": {**cfg["stages"], "chamfer": {**cfg2["stages"]["chamfer"], "yaw_lock": False}}}
        pose, betas, root, trans = (T(k) for k in ("ch_pose", "ch_betas", "ch_root", "ch_trans"))
        for t in (pose, betas, trans):
            t.requires_grad_(True)
        z33 = T("mk_root")                       # a slightly non-orthonormal 3x3 per frame to evaluate at
        OneEvaluation.at = {1: z33}
        ns["optim_chamfer"](markers, pose, T("ch_o_pose"), betas, T("ch_o_betas"), root, trans, torch.ones(Fn), None, si, cfg3)
        ev = OneEvaluation.last
        out.update(ch3_z=z33.numpy(), ch3_loss=np.float32(ev.loss), ch3_g_trans=ev.grads[0].numpy(), ch3_g_z=ev.grads[1].numpy(),
                   ch3_g_betas=ev.grads[2].numpy(), ch3_g_pose=ev.grads[3].numpy(), ch3_root_baked=root.detach().numpy())

        # -- compute_nearest_points (optimization.py:402-642), use_mean branch, one frame masked out
        OneEvaluation.at = None
        img_mask = torch.ones(Fn)
        img_mask[1] = 0
        si.smpl.faces = model.faces
        one_hot = ns["compute_nearest_points"](
            markers=markers, pose_body=T("ch_pose"), betas=T("ch_betas"), root_orient=T("mk_root"), trans=T("ch_trans"),
            smpl_inference=si, marker_labels=None, granularity="full", img_mask=img_mask, device=torch.device("cpu"),
            config=cfg, o_pose_body=None, window_size=1, use_velocity=False)
        out["np_img_mask"] = img_mask.numpy()
        out["np_vertex_idx"] = torch.argmax(one_hot, dim=1).numpy()
        assert float(one_hot.sum()) == M
        # the use_mean: False branch (per-frame closest_point, :503-510, 538-590) for the three granularities
        cfg_cp = {**cfg, "stages": {**cfg["stages"], "compute_locations": {"use_velocity": False, "use_barycentric": False,
                                                                          "use_mean": False}}}
        mask2 = torch.ones(Fn)
        mask2[Fn - 1] = 0                                     # the last frame carries no video prior
        lab = np.random.default_rng(9).integers(0, 24, size=(Fn, M))
        lab[Fn - 2, ::3] = 23                                 # label 23 is outside range(num_joints = 23): not updated there
        out["cp_img_mask"], out["cp_labels"] = mask2.numpy(), lab
        with torch.no_grad():
            vv = si(poses=p3d.normalize_rot(T("ch_pose")), betas=T("ch_betas").repeat(Fn, 1),
                    root_orient=p3d.normalize_rot(T("mk_root")), trans=T("ch_trans"))["vertices"].numpy()
        for gran in ("full", "marker", "part"):
            oh = ns["compute_nearest_points"](
                markers=markers, pose_body=T("ch_pose"), betas=T("ch_betas"), root_orient=T("mk_root"), trans=T("ch_trans"),
                smpl_inference=si, marker_labels=lab, granularity=gran, img_mask=mask2, device=torch.device("cpu"),
                config=cfg_cp, o_pose_body=None, window_size=1, use_velocity=False)
            # the reference scatters the weight 1 and then the two zero weights to the SAME index (:543-549), so this branch
            # returns an all-zero matrix; what it computed before that is closest_point of the last frame that updates a row
            assert float(oh.abs().sum()) == 0.0, gran
            want = np.zeros(M, dtype=np.int64)
            for c in range(M):
                fr = [f for f in range(Fn) if mask2[f] == 1 and (gran != "part" or lab[f, c] < 23)]
                want[c] = ns["closest_point"](markers[fr[-1]].numpy(), vv[fr[-1]])["vertex_indices"][c]
            out["cp_idx_" + gran] = want

        # -- optim_markers (optimization.py:288-399): params = [pose_body, betas, root_orient, trans]
        vidx = torch.from_numpy(g["mk_vidx"]).long()
        oh = torch.zeros(M, model.v_template.shape[0])
        oh[torch.arange(M), vidx] = 1.0
        pose, betas, root, trans = (T(k) for k in ("ch_pose", "ch_betas", "mk_root", "ch_trans"))
        for t in (pose, betas, root, trans):
            t.requires_grad_(True)
        ns["optim_markers"](markers, pose, T("ch_o_pose"), betas, T("ch_o_betas"), root, trans, oh, torch.ones(Fn), si, cfg)
        ev = OneEvaluation.last
        assert ev.kw["lr"] == 1.0
        out.update(mk_loss=np.float32(ev.loss), mk_g_pose=ev.grads[0].numpy(), mk_g_betas=ev.grads[1].numpy(),
                   mk_g_root=ev.grads[2].numpy(), mk_g_trans=ev.grads[3].numpy())

        # -- small helpers
        out["marker_mask"] = ns["get_marker_mask"](markers).numpy()
        ang = torch.tensor([[[0.0]], [[1e-7]], [[0.3]], [[-2.5]], [[3.14159]]])
        out["rot_z_angles"] = ang.numpy()
        out["rot_z"] = ns["compute_root_orient_z"](ang).numpy()
        out["rot_y"] = ns["compute_root_orient_y"](ang).numpy()
        with torch.no_grad():
            o = si(poses=p3d.normalize_rot(T("ch_pose")), betas=T("ch_betas").repeat(Fn, 1),
                   root_orient=p3d.normalize_rot(T("mk_root")), trans=T("ch_trans"))
        cp = ns["closest_point"](markers[0].numpy(), o["vertices"][0].numpy())            # utils/points.py:19-42
        out.update(cp_vertices=o["vertices"][0].numpy(), cp_idx=cp["vertex_indices"], cp_dist=cp["distances"].astype(np.float32))
        w = ns["get_marker_mask"](markers)
        out["wcd_value"] = np.float32(ns["weighted_chamfer_distance"](markers, o["vertices"], w, single_directional=True)[0].item())
        x = torch.rand(4, 5, 3)
        out["wmse"] = np.float32(ns["weighted_mse_loss"](x, x * 0.5, torch.arange(4.0).view(4, 1, 1)).item())
        out["wmse_x"] = x.numpy()
        labels = torch.from_numpy(np.random.default_rng(5).integers(0, 24, size=(Fn, M)))
        cbp = ns["chamfer_distance_by_part"](markers, o["vertices"], labels, si.smpl.lbs_weights, single_directional=True)
        out.update(cbp_labels=labels.numpy(), cbp_value=np.float32(float(cbp)))

        # -- find_best_part_fits (markers/markers_utils.py:274-638), "cluster" mode: every candidate sub-tree's
        # closure_fit_subtree evaluated once (run A: at z = 0.2 for the closure parity; run B: at the reference's own
        # start, z = 0, whose scores / ranking / labels / weights are the function's outputs with no optimiser step)
        import itertools
        from tqdm import tqdm
        sys.path.insert(0, "/root/reference/third_party/smplx")
        from smplx.joint_names import SMPL_JOINT_NAMES
        ns.update({"itertools": itertools, "tqdm": tqdm, "SMPL_JOINT_NAMES": SMPL_JOINT_NAMES, "os": os, "shutil": None})
        exec(take(REF + "/utils/aabb.py", ["get_aabb", "get_aabb_volume"]), ns)
        exec(take(REF + "/utils/smpl_utils.py", ["get_joint_id", "get_joint_name", "get_sub_hierachies"]), ns)
        exec(take(REF + "/markers/markers_utils.py", ["find_best_part_fits"]), ns)
        Fp = 8
        cap = synthetic.make_capture(model, Fp, num_markers=41, seed=23, noise_std=0.001)
        rot = synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)
        vlab = np.argmax(model.weights, axis=-1)
        rng = np.random.default_rng(5)
        vids = np.concatenate([rng.choice(np.where(vlab == j)[0], size=6, replace=False) for j in (16, 18, 20)])
        pts = synthetic.lbs_numpy(model, cap.gt_betas, rot.astype(np.float64), cap.gt_trans.astype(np.float64), vertex_ids=vids)
        pm = torch.from_numpy(pts.astype(np.float32))
        plabels = torch.from_numpy(np.repeat(vlab[vids][None], Fp, 0).astype(np.int64))
        pcfg = {"optimizer": {"tolerance_grad": 1e-7, "tolerance_change": 1e-9},
                "stages": {"part": {"mode": "cluster", "num_iters": 1, "losses": {"chamfer": 10.0, "reg_betas": 0.1}}}}
        r = torch.from_numpy(rot)
        pb0 = torch.from_numpy(cap.gt_betas * 0.5)
        out.update(pf_markers=pm.numpy(), pf_labels=plabels.numpy(), pf_rot=rot, pf_betas=pb0.numpy())
        for tag, at in (("pfA", {0: torch.tensor([0.2])}), ("pfB", None)):
            OneEvaluation.at, OneEvaluation.history = at, []
            import contextlib
            import io
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                res = ns["find_best_part_fits"](pm, r[:, 1:], pb0, r[:, :1], plabels, si, np.array(model.parents), None, None,
                                                None, None, None, pcfg)
            # the function prints its three best candidates as "<joint names> <bidirectional chamfer, 6 decimals>" (:611-613)
            top = [ln for ln in buf.getvalue().splitlines() if ln and ln[-1].isdigit() and "," in ln]
            out[tag + "_top3_dist"] = np.array([float(ln.rsplit(" ", 1)[1]) for ln in top], dtype=np.float64)
            out[tag + "_top3_joints"] = np.array([[SMPL_JOINT_NAMES.index(n.strip()) for n in ln.rsplit(" ", 1)[0].split(",")]
                                                  for ln in top], dtype=np.int32)
            evs = OneEvaluation.history
            out[tag + "_loss"] = np.array([e.loss for e in evs], dtype=np.float32)
            out[tag + "_g_z"] = np.stack([e.grads[0].numpy().reshape(-1) for e in evs])
            out[tag + "_g_trans"] = np.stack([e.grads[1].numpy() for e in evs])
            out[tag + "_g_betas"] = np.stack([e.grads[2].numpy() for e in evs])
            if tag == "pfB":
                out.update(pfB_chain=res["chain"], pfB_marker_labels=res["marker_labels"].numpy(),
                           pfB_marker_weights=res["marker_weights"].numpy(), pfB_aabb=res["aabb_volume_ratio"].detach().numpy(),
                           pfB_trans=res["trans"].detach().numpy(), pfB_root=res["root_orient"].detach().numpy())
        OneEvaluation.at = None
        out["pf_subtrees"] = np.array(ns["get_sub_hierachies"](np.array(model.parents), 3), dtype=np.int32)
    finally:
        torch.optim.LBFGS, pdb.set_trace = real_lbfgs, real_trace
    np.savez_compressed(os.path.join(GOLD, "closure_ref.npz"), **out)
    print({k: (float(v) if np.ndim(v) == 0 else np.shape(v)) for k, v in out.items()})


if __name__ == "__main__":
    main()
