"""Generate ``tests/golden/*.npz`` (TEST INFRASTRUCTURE; run in the build container only).

``smplx_*.npz`` are outputs of the REFERENCE ITSELF: the vendored, unmodified
``/root/reference/third_party/smplx`` is imported by path (never copied) and executed on the
synthetic model through ``smplx.SMPL(data_struct=...)`` (body_models.py:57,132-143).
``closure_*.npz`` come from the oracle restatement (``oracle/closures.py``), because the
reference's ``video_mocap`` package cannot be imported here (pytorch3d, igl, trimesh ...
missing); they pin the oracle against drift.  The pin against the REFERENCE is ``closure_ref.npz``:
``oracle/gen_golden_ref_closures.py`` compiles the reference's own stage functions from source and runs them on the
inputs stored here.

    python -m oracle.gen_golden            # from the repo root
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SMPLX = "/root/reference/third_party/smplx"
GOLD = os.path.join(ROOT, "tests", "golden")


def _load_pkg():
    sys.path.insert(0, ROOT)
    import uuo_mocap_b200  # noqa: F401  (import shim for the hyphenated package dir)
    from uuo_mocap_b200 import synthetic
    return synthetic


def reference_smpl(model):
    """The real smplx.SMPL built from the synthetic data_struct."""
    if REF_SMPLX not in sys.path:
        sys.path.insert(0, REF_SMPLX)
    import smplx
    from smplx.utils import Struct
    ds = Struct(v_template=model.v_template, shapedirs=model.shapedirs, posedirs=model.posedirs,
                J_regressor=model.J_regressor, kintree_table=model.kintree_table, weights=model.weights,
                f=model.faces)
    return smplx.SMPL(model_path="", data_struct=ds, batch_size=1)


def gen_smplx(synthetic, weights: str, F: int = 3, seed: int = 11):
    model = synthetic.make_model(weights)
    smpl = reference_smpl(model)
    mb = synthetic.make_lbs_microbench(F, seed=seed)
    rng = np.random.default_rng(seed + 1)
    out = {"model_checksum": np.float64(model.checksum()), "weights_kind": weights}
    gv = torch.from_numpy(mb["g_vertices"])
    gj = torch.from_numpy(mb["g_joints"])
    for tag, pose2rot in (("rot", False), ("aa", True)):
        if pose2rot:
            pose = torch.from_numpy(mb["pose_aa"]).clone()
            go, bp = pose[:, :1].reshape(F, 3), pose[:, 1:].reshape(F, 69)
        else:
            # NOT orthonormal on purpose for one frame: lbs(pose2rot=False) takes any 3x3
            rot = torch.from_numpy(mb["rot"]).clone()
            rot[0] += torch.from_numpy(rng.normal(0, 0.05, size=(24, 3, 3)).astype(np.float32))
            go, bp = rot[:, :1].clone(), rot[:, 1:].clone()
        betas = torch.from_numpy(mb["betas"]).clone()
        transl = torch.from_numpy(mb["transl"]).clone()
        for t in (go, bp, betas, transl):
            t.requires_grad_(True)
        o = smpl(betas=betas, body_pose=bp, global_orient=go, transl=transl, pose2rot=pose2rot)
        loss = (o.vertices * gv).sum() + (o.joints * gj).sum()
        loss.backward()
        out.update({
            f"{tag}_global_orient": go.detach().numpy(), f"{tag}_body_pose": bp.detach().numpy(),
            f"{tag}_betas": betas.detach().numpy(), f"{tag}_transl": transl.detach().numpy(),
            f"{tag}_vertices": o.vertices.detach().numpy(), f"{tag}_joints": o.joints.detach().numpy(),
            f"{tag}_g_global_orient": go.grad.numpy(), f"{tag}_g_body_pose": bp.grad.numpy(),
            f"{tag}_g_betas": betas.grad.numpy(), f"{tag}_g_transl": transl.grad.numpy(),
        })
    # shared betas row (the fitting loop's [1,10] broadcast, body_models.py:371-376)
    rot = torch.from_numpy(mb["rot"])
    o = smpl(betas=torch.from_numpy(mb["betas"][:1]), body_pose=rot[:, 1:], global_orient=rot[:, :1],
             transl=torch.from_numpy(mb["transl"]), pose2rot=False)
    out["shared_vertices"] = o.vertices.detach().numpy()
    out["shared_joints"] = o.joints.detach().numpy()
    out["g_vertices"] = mb["g_vertices"]
    out["g_joints"] = mb["g_joints"]
    np.savez_compressed(os.path.join(GOLD, f"smplx_{weights}.npz"), **out)
    print("wrote smplx_%s.npz" % weights)


def gen_closures(synthetic, F: int = 6, seed: int = 21):
    from oracle import closures, p3d_oracle as p3d
    from oracle.smpl_oracle import OracleSMPL
    model = synthetic.make_model("top4")
    smpl = OracleSMPL.from_model(model)
    cap = synthetic.make_capture(model, F, num_markers=41, seed=seed, n_ghost=2, dropout_p=0.1, dropout_block=2)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    markers = T(cap.markers)
    o_pose, o_betas = T(cap.prior_pose_body), T(cap.prior_betas)
    rng = np.random.default_rng(seed)
    out = {"model_checksum": np.float64(model.checksum()), "markers": cap.markers}

    # chamfer stage (video_mocap.yaml:40-47 weights) + optional trans_vel
    pose = (o_pose + T(rng.normal(0, 0.02, size=o_pose.shape).astype(np.float32))).requires_grad_(True)
    betas = (o_betas + 0.1).requires_grad_(True)
    root = T(cap.prior_root_orient).clone()
    trans = T(cap.prior_trans).clone().requires_grad_(True)
    z = T(rng.normal(0, 0.2, size=(F, 1, 1)).astype(np.float32)).requires_grad_(True)
    losses = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0, "trans_vel": 1.0}
    loss, idx = closures.chamfer_stage_loss(smpl, markers, pose, o_pose, betas, o_betas, root, trans, z, losses,
                                            return_idx=True)
    loss.backward()
    out.update(ch_pose=pose.detach().numpy(), ch_betas=betas.detach().numpy(), ch_root=root.numpy(),
               ch_trans=trans.detach().numpy(), ch_z=z.detach().numpy(), ch_o_pose=o_pose.numpy(),
               ch_o_betas=o_betas.numpy(), ch_loss=np.float32(loss.item()), ch_idx=idx.numpy(),
               ch_g_pose=pose.grad.numpy(), ch_g_betas=betas.grad.numpy(), ch_g_trans=trans.grad.numpy(),
               ch_g_z=z.grad.numpy())

    # marker stage (video_mocap.yaml:52-58 weights) + optional temporal
    pose = pose.detach().clone().requires_grad_(True)
    betas = betas.detach().clone().requires_grad_(True)
    trans = trans.detach().clone().requires_grad_(True)
    rootp = (root + T(rng.normal(0, 0.02, size=root.shape).astype(np.float32))).requires_grad_(True)
    vidx, _ = closures.mean_distance_argmin(smpl, markers, pose.detach(), betas.detach(), rootp.detach(),
                                            trans.detach(), T(cap.img_mask))
    losses = {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0, "temporal": 1.0}
    loss = closures.marker_stage_loss(smpl, markers, pose, o_pose, betas, o_betas, rootp, trans,
                                      torch.from_numpy(vidx), losses)
    loss.backward()
    out.update(mk_root=rootp.detach().numpy(), mk_vidx=vidx, mk_loss=np.float32(loss.item()),
               mk_g_pose=pose.grad.numpy(), mk_g_betas=betas.grad.numpy(), mk_g_trans=trans.grad.numpy(),
               mk_g_root=rootp.grad.numpy())

    # the reference's only deterministic seed: losses/chamfer_distance.py:24-36
    x = torch.arange(0, 16 * 7 * 3, dtype=torch.float32).reshape(16, 7, 3)
    y = torch.arange(0, 16 * 19 * 3, dtype=torch.float32).reshape(16, 19, 3)
    w = torch.ones(16, 7)
    w[:, ::2] = 0
    val, _ = p3d.weighted_chamfer_distance(x, y, w)
    out["smoke_weighted_chamfer"] = np.float32(val.item())
    np.savez_compressed(os.path.join(GOLD, "closure_top4.npz"), **out)
    print("wrote closure_top4.npz  (chamfer smoke value = %r)" % float(val))


def main():
    os.makedirs(GOLD, exist_ok=True)
    synthetic = _load_pkg()
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    gen_smplx(synthetic, "dense")
    gen_smplx(synthetic, "top4", F=2)
    gen_closures(synthetic)


if __name__ == "__main__":
    main()
