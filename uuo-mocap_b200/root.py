"""Root stage: ``optim_root`` of /root/reference/src/video_mocap/optimization.py:21-144 on the CUDA path
(SURVEY.md section 8 rows A22 -- the ``root_orient_vel`` term, :68-75,105-106 -- and N1).

The stage fits a yaw offset (one for the take, one per frame, or a free 3x3 per frame), the per-frame translation
and the betas with the body pose frozen.  The shipped ``video_mocap.yaml`` sets ``num_iters: 0`` for it and defines
no ``lr`` (a latent KeyError in the reference), so here ``lr`` defaults to 1.0.

Every closure evaluation is the dense CUDA lbs() forward/backward (``SMPL.forward``'s autograd pair over
uuo_lbs_forward / uuo_lbs_backward) and the CUDA correspondence kernel (``losses``); the [F]-sized velocity terms
are a handful of torch ops on the device.  Not taken: "reg_betas" (the reference reads an undefined ``o_betas``
there and raises NameError) and "ground" (stops in pdb).
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import math

import torch

from . import losses as L
from .fit import _model_of, get_marker_mask, normalize_rot, rot_z


def acos_linear_extrapolation(x: torch.Tensor, lo: float, hi: float) -> torch.Tensor:
    """pytorch3d.transforms.math.acos_linear_extrapolation: arccos on (lo, hi), first-order extension outside."""
    ext = lambda x0: math.acos(x0) + (x - x0) * (-1.0 / math.sqrt(1.0 - x0 * x0))
    inner = torch.acos(x.clamp(lo, hi))
    return torch.where(x <= lo, ext(lo), torch.where(x >= hi, ext(hi), inner))


def so3_relative_angle(R1: torch.Tensor, R2: torch.Tensor, cos_bound: float = 1e-4, eps: float = 1e-4) -> torch.Tensor:
    """pytorch3d.transforms.so3_relative_angle with its defaults (call site optimization.py:66-74): [N,3,3] x2 -> [N]."""
    R = torch.bmm(R1, R2.transpose(1, 2))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    if bool(((tr < -1.0 - eps) | (tr > 3.0 + eps)).any()):
        raise ValueError("A matrix has trace outside valid range [-1-eps,3+eps].")
    b = 1.0 - cos_bound
    return acos_linear_extrapolation((tr - 1.0) * 0.5, -b, b)


def root_closure_loss(smpl, markers, pose_body, betas, root_orient, trans, z_angle, marker_labels, stage: Dict):
    """The loss of closure_stage_translation (optimization.py:57-117), differentiable in trans / z_angle / betas."""
    F = pose_body.shape[0]
    if stage.get("constrained_rotation", False):                                          # :61-62
        z_root = rot_z(z_angle.expand(F, -1, -1)) @ root_orient
    elif stage.get("yaw_lock", False):                                                    # :63-64
        z_root = rot_z(z_angle) @ root_orient
    else:                                                                                 # :65-66
        z_root = normalize_rot(z_angle)
    terms = stage["losses"]
    for name in ("reg_betas", "ground"):
        if name in terms:
            raise NotImplementedError("optim_root term %r is broken in the reference (NameError / pdb) and not taken" % name)
    out = smpl(betas=betas.expand(F, -1), body_pose=pose_body, global_orient=z_root, transl=trans, pose2rot=False)
    loss = 0
    if "part_chamfer" in terms:                                                           # :88-96
        loss = loss + L.chamfer_distance_by_part(markers, out.vertices, marker_labels, smpl.lbs_weights,
                                                 single_directional=stage["single_directional"]) * float(terms["part_chamfer"])
    if "full_chamfer" in terms:                                                           # :97-104
        loss = loss + L.weighted_chamfer_distance(markers, out.vertices, get_marker_mask(markers).float(),
                                                  model=smpl.handle)[0] * float(terms["full_chamfer"])
    if "root_orient_vel" in terms:                                                        # :68-75,105-106
        vel0 = so3_relative_angle(normalize_rot(root_orient[1:, 0]), normalize_rot(root_orient[:-1, 0]))
        vel1 = so3_relative_angle(normalize_rot(z_root[1:, 0]), normalize_rot(z_root[:-1, 0]))
        loss = loss + torch.nn.functional.mse_loss(vel1, vel0) * float(terms["root_orient_vel"])
    if "trans_vel" in terms:                                                              # :107-111
        mm = markers.mean(dim=1)
        loss = loss + torch.nn.functional.mse_loss(trans[1:] - trans[:-1], mm[1:] - mm[:-1]) * float(terms["trans_vel"])
    return loss


def optim_root(markers, pose_body, betas, root_orient, trans, marker_labels, smpl_inference, config: Dict,
               verbose: bool = False, iter_fn: Optional[Callable] = None):
    """optimization.py:21-144.  ``trans`` [F,3] and ``betas`` [1,10] are stepped and ``root_orient`` [F,1,3,3] gets the
    fitted offset baked in, all IN PLACE like the reference.  Optimiser: L-BFGS / strong Wolfe (the reference's) or
    Adam with ``optimizer.name: adam``."""
    smpl = _model_of(smpl_inference)
    dev = smpl.device
    st = config["stages"]["root"]
    oc = config.get("optimizer", {})
    T = lambda t: t.detach().to(dev, torch.float32).contiguous()
    mk, pose, root = T(markers), T(pose_body), T(root_orient)
    labels = None if marker_labels is None else marker_labels.to(dev)
    F = root.shape[0]
    if st.get("constrained_rotation", False):                                             # :34-35
        z_angle = torch.zeros((1, root.shape[1], 1), device=dev)
    elif st.get("yaw_lock", False):                                                       # :36-37
        z_angle = torch.zeros((F, root.shape[1], 1), device=dev)
    else:                                                                                 # :38-43
        z_angle = torch.eye(3, device=dev).repeat(F, root.shape[1], 1, 1)
    z_angle.requires_grad_(True)
    p_trans = T(trans).clone().requires_grad_(True)
    p_betas = T(betas).reshape(1, -1).clone().requires_grad_(True)
    params = [p_trans, z_angle, p_betas]
    it = [0]

    def evaluate():
        loss = root_closure_loss(smpl, mk, pose, p_betas, root, p_trans, z_angle, labels, st)
        if not torch.is_tensor(loss):
            raise ValueError("optim_root: no loss term enabled in stages.root.losses")
        return loss

    def report(loss):
        if verbose:
            print("Root", it[0], float(loss))
        if iter_fn is not None:
            iter_fn(stage="root", iteration=it[0], pose_body=pose.cpu().numpy(), betas=p_betas.detach().cpu().numpy(),
                    trans=p_trans.detach().cpu().numpy(), root_orient=_baked(False).cpu().numpy())
        it[0] += 1

    def _baked(final: bool):
        with torch.no_grad():
            if st.get("constrained_rotation", False):
                return rot_z(z_angle.expand(F, -1, -1)) @ root
            if st.get("yaw_lock", False):
                return rot_z(z_angle) @ root
            # the closure poses the body with the offset ALONE (:66); only the write-back multiplies it in (:141)
            return normalize_rot(z_angle) @ root if final else normalize_rot(z_angle)

    n = int(st["num_iters"])
    if n > 0:
        if str(oc.get("name", "lbfgs")).lower() == "lbfgs":                               # :46-53
            opt = torch.optim.LBFGS(params, max_iter=n, tolerance_grad=float(oc.get("tolerance_grad", 1e-7)),
                                    tolerance_change=float(oc.get("tolerance_change", 1e-9)), lr=float(st.get("lr", 1.0)),
                                    line_search_fn="strong_wolfe")

            def closure():
                opt.zero_grad()
                loss = evaluate()
                loss.backward()
                report(loss)
                return loss
            opt.step(closure)
        else:
            opt = torch.optim.Adam(params, lr=float(st.get("lr_adam", st.get("lr", oc.get("lr", 0.01)))))
            for _ in range(n):
                opt.zero_grad()
                loss = evaluate()
                loss.backward()
                report(loss)
                opt.step()
    with torch.no_grad():                                                                 # :137-144
        root_orient.copy_(_baked(True).to(root_orient.device))
        trans.copy_(p_trans.detach().to(trans.device))
        betas.copy_(p_betas.detach().view_as(betas).to(betas.device))
    return {"z_angle": z_angle.detach(), "iterations": it[0]}
