"""Oracle restatement of the reference's stage closures (TEST INFRASTRUCTURE, CPU, fp32).

Follows ``/root/reference/src/video_mocap/optimization.py`` (line numbers below) on top of
``smpl_oracle`` (pinned to the vendored smplx) and ``p3d_oracle`` (pytorch3d restatement,
parity unpinned).  The optimiser in the reference is ``torch.optim.LBFGS`` (:176-183); the
north-star optimiser is Adam, so both drivers are provided and parity is defined on the
closure (loss + every gradient), which is optimiser independent.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

from . import p3d_oracle as p3d
from .smpl_oracle import OracleSMPL

MARKER_DISTANCE = 0.0095   # src/video_mocap/utils/settings.py:1


def get_marker_mask(markers: torch.Tensor) -> torch.Tensor:
    """optimization.py:703-715."""
    return torch.sum(torch.abs(markers), dim=-1) != 0.0


def marker_loss(markers, virtual_markers, marker_weights, marker_distance):
    """src/video_mocap/losses/losses.py:43-51."""
    ml = (torch.norm(markers - virtual_markers, dim=-1) - marker_distance) ** 2
    return ml * marker_weights


def _smpl_chunked(smpl: OracleSMPL, poses, betas_rows, root, trans, chunk):
    """SMPL forward in frame chunks so the [F,V,...] temporaries fit in RAM (values unchanged:
    every op on the path is frame-independent)."""
    verts, joints = [], []
    Fn = poses.shape[0]
    for s in range(0, Fn, chunk):
        e = min(Fn, s + chunk)
        o = smpl.smpl_inference(poses[s:e], betas_rows[s:e], root[s:e], trans[s:e])
        verts.append(o["vertices"])
        joints.append(o["joints"])
    return torch.cat(verts, 0), torch.cat(joints, 0)


def chamfer_stage_loss(smpl: OracleSMPL, markers, pose_body, o_pose_body, betas, o_betas, root_orient, trans,
                       z_angle, losses: Dict[str, float], yaw_lock: bool = True, chunk: int = 256,
                       return_idx: bool = False):
    """optimization.py:187-256 (closure_stage_chamfer) up to, not including, backward()."""
    Fn = pose_body.shape[0]
    if yaw_lock:
        z_root = p3d.compute_root_orient_z(z_angle) @ root_orient                    # :192
    else:
        z_root = p3d.normalize_rot(z_angle)                                          # :194
    verts, _ = _smpl_chunked(
        smpl, p3d.normalize_rot(pose_body), torch.repeat_interleave(betas, dim=0, repeats=Fn),
        p3d.normalize_rot(z_root), trans, chunk)                                     # :196-202
    loss = 0
    idx = None
    if "full_chamfer" in losses:                                                     # :232-239
        cham, idx = p3d.weighted_chamfer_distance(markers, verts, get_marker_mask(markers),
                                                  single_directional=True, return_idx=True)
        loss = loss + cham * losses["full_chamfer"]
    if "reg_pose_body" in losses:                                                    # :243-245
        loss = loss + F.mse_loss(pose_body, o_pose_body) * losses["reg_pose_body"]
    if "trans_vel" in losses:                                                        # :246-250
        tv = trans[1:] - trans[:-1]
        mm = torch.mean(markers, dim=1)
        mv = mm[1:] - mm[:-1]
        loss = loss + F.mse_loss(tv, mv) * losses["trans_vel"]
    if "reg_betas" in losses:                                                        # :255-256
        loss = loss + F.mse_loss(betas, o_betas) * losses["reg_betas"]
    if return_idx:
        return loss, idx
    return loss


def marker_stage_loss(smpl: OracleSMPL, markers, pose_body, o_pose_body, betas, o_betas, root_orient, trans,
                      vertex_idx: torch.Tensor, losses: Dict[str, float], chunk: int = 256):
    """optimization.py:329-375 (closure_stage_marker_pose); ``vertex_idx`` [M] is the argmax of
    the one-hot matrix, so the :345-351 contraction is a gather."""
    Fn = pose_body.shape[0]
    verts, _ = _smpl_chunked(
        smpl, p3d.normalize_rot(pose_body), torch.repeat_interleave(betas, dim=0, repeats=Fn),
        p3d.normalize_rot(root_orient), trans, chunk)                                # :335-340
    virtual = verts[:, vertex_idx]                                                   # :345-351
    loss = 0
    if "marker" in losses:                                                           # :356-362
        ml = marker_loss(markers, virtual, get_marker_mask(markers), MARKER_DISTANCE)
        loss = loss + torch.mean(ml) * losses["marker"]
    if "reg_pose_body" in losses:                                                    # :363-365
        loss = loss + F.mse_loss(pose_body, o_pose_body) * losses["reg_pose_body"]
    if "reg_betas" in losses:                                                        # :366-367
        loss = loss + F.mse_loss(betas, o_betas) * losses["reg_betas"]
    if "temporal" in losses:                                                         # :368-374 (sign as written)
        vel = pose_body[2:] - (2 * pose_body[1:-1]) - pose_body[0:-2]
        loss = loss + F.mse_loss(vel, torch.zeros_like(vel)) * losses["temporal"]
    return loss


def one_hot_gather_dense(vertices: torch.Tensor, one_hot: torch.Tensor) -> torch.Tensor:
    """optimization.py:345-351 exactly as written (dense [F,M,V,3] product) -- small inputs only."""
    M = one_hot.shape[0]
    v = torch.repeat_interleave(vertices.unsqueeze(1), repeats=M, dim=1)
    bc = one_hot.unsqueeze(0).unsqueeze(-1)
    bc = torch.repeat_interleave(bc, repeats=v.shape[0], dim=0)
    bc = torch.repeat_interleave(bc, repeats=3, dim=-1)
    return torch.sum(v * bc, dim=2)


def mean_distance_argmin(smpl: OracleSMPL, markers, pose_body, betas, root_orient, trans, img_mask,
                         chunk: int = 64):
    """compute_nearest_points with use_mean (optimization.py:422-430,464-486,595-603).

    numpy fp32: D[f,m,v] = ||v_{f,v} - m_{f,m}||_2 (np.linalg.norm), frames whose index is in
    valid_frames filled, others left zero (:465,:474-476), mean over frames with img_mask==1,
    argmin over V (first minimum).  Returns (idx[M] int64, mean_matrix[M,V] float32)."""
    Fn = pose_body.shape[0]
    mean_betas = torch.repeat_interleave(torch.mean(betas, dim=0, keepdim=True), dim=0, repeats=Fn)  # :424
    with torch.no_grad():
        verts, _ = _smpl_chunked(smpl, p3d.normalize_rot(pose_body), mean_betas,
                                 p3d.normalize_rot(root_orient), trans, 256)
    verts = verts.numpy()
    mk = markers.detach().numpy()
    mask = img_mask.detach().numpy()
    sel = np.where(mask == 1)[0]
    M, V = mk.shape[1], verts.shape[1]
    # np.mean(distance_matrix[sel], axis=0) on a float32 [n,M,V] array adds the frames one
    # after the other in fp32 (reduction over the OUTER axis) and divides by n: same here.
    acc = np.zeros((M, V), dtype=np.float32)
    for s in range(0, len(sel), chunk):
        fr = sel[s:s + chunk]
        d = np.linalg.norm(verts[fr][:, None, :, :] - mk[fr][:, :, None, :], axis=-1)  # [c,M,V] fp32
        for i in range(d.shape[0]):
            acc += d[i]
    mean = acc / np.float32(max(1, len(sel)))
    idx = np.argmin(mean, axis=-1)                                                   # :599
    return idx.astype(np.int64), mean


def part_fit_loss(smpl: OracleSMPL, markers_subset, pose_body, betas, o_betas, root_orient, trans, z_angle,
                  vertex_indices: torch.Tensor, losses: Dict[str, float], chunk: int = 256):
    """src/video_mocap/markers/markers_utils.py:454-522 (closure_fit_subtree), chamfer + reg_betas."""
    Fn = pose_body.shape[0]
    z_root = p3d.compute_root_orient_z(torch.repeat_interleave(z_angle, repeats=Fn, dim=0)) @ root_orient
    verts, _ = _smpl_chunked(smpl, pose_body, torch.repeat_interleave(betas, dim=0, repeats=Fn),
                             z_root, trans, chunk)
    vs = verts[:, vertex_indices]
    loss = p3d.chamfer_distance(markers_subset, vs, single_directional=True)[0] * losses["chamfer"]
    if "reg_betas" in losses:
        loss = loss + F.mse_loss(betas, o_betas) * losses["reg_betas"]
    return loss


def root_stage_loss(smpl: OracleSMPL, markers, pose_body, betas, root_orient, trans, z_angle, marker_labels,
                    stage: Dict, chunk: int = 256):
    """src/video_mocap/optimization.py:57-117 (closure_stage_translation of optim_root).  ``stage`` is
    config["stages"]["root"].  The "reg_betas" term there reads an undefined ``o_betas`` (NameError) and the
    "ground" term stops in pdb, so neither is restated."""
    Fn = pose_body.shape[0]
    if stage.get("constrained_rotation", False):                                    # :61-62
        z_root = p3d.compute_root_orient_z(torch.repeat_interleave(z_angle, repeats=Fn, dim=0)) @ root_orient
    elif stage.get("yaw_lock", False):                                              # :63-64
        z_root = p3d.compute_root_orient_z(z_angle) @ root_orient
    else:                                                                           # :65-66
        z_root = p3d.normalize_rot(z_angle)
    losses = stage["losses"]
    verts, _ = _smpl_chunked(smpl, pose_body, torch.repeat_interleave(betas, dim=0, repeats=Fn), z_root, trans, chunk)
    loss = 0
    if "part_chamfer" in losses:                                                    # :88-96
        loss = loss + chamfer_distance_by_part(markers, verts, marker_labels, smpl.lbs_weights,
                                               single_directional=stage["single_directional"]) * losses["part_chamfer"]
    if "full_chamfer" in losses:                                                    # :97-104
        loss = loss + p3d.weighted_chamfer_distance(markers, verts, get_marker_mask(markers).float(),
                                                    single_directional=stage["single_directional"])[0] * losses["full_chamfer"]
    if "root_orient_vel" in losses:                                                 # :68-75,105-106
        vel0 = p3d.so3_relative_angle(p3d.normalize_rot(root_orient[1:, 0]), p3d.normalize_rot(root_orient[:-1, 0]))
        vel1 = p3d.so3_relative_angle(p3d.normalize_rot(z_root[1:, 0]), p3d.normalize_rot(z_root[:-1, 0]))
        loss = loss + F.mse_loss(vel1, vel0) * losses["root_orient_vel"]
    if "trans_vel" in losses:                                                       # :107-111
        mm = torch.mean(markers, dim=1)
        loss = loss + F.mse_loss(trans[1:] - trans[:-1], mm[1:] - mm[:-1]) * losses["trans_vel"]
    return loss


def chamfer_distance_by_part(markers, vertices, marker_labels, vertex_weights, single_directional: bool = False):
    """src/video_mocap/optimization.py:682-700 ("part_chamfer")."""
    vertex_mask = torch.argmax(vertex_weights, dim=-1)
    marker_labels_mode = torch.mode(marker_labels, dim=0)[0]
    loss = 0
    for i in torch.unique(marker_labels_mode).tolist():
        vertices_part = vertices[:, vertex_mask == i]
        markers_part = markers[:, marker_labels_mode == i]
        loss_part = (p3d.chamfer_distance(vertices_part, markers_part, single_directional=single_directional)[0]
                     - MARKER_DISTANCE) ** 2
        loss = loss + loss_part
    return loss


# ------------------------------------------------------------------------------------------
# optimiser drivers
# ------------------------------------------------------------------------------------------
def run_adam(params, loss_fn: Callable[[], torch.Tensor], num_iters: int, lr: float,
             betas=(0.9, 0.999), eps: float = 1e-8, on_iter: Optional[Callable] = None):
    """torch.optim.Adam over the oracle closure: the CPU twin of the fused Adam step."""
    opt = torch.optim.Adam(params, lr=lr, betas=betas, eps=eps)
    last = None
    for it in range(num_iters):
        opt.zero_grad()
        loss = loss_fn()
        loss.backward()
        opt.step()
        last = float(loss)
        if on_iter is not None:
            on_iter(it, last)
    return last


def run_lbfgs(params, loss_fn: Callable[[], torch.Tensor], max_iter: int, lr: float,
              tolerance_grad: float = 1e-7, tolerance_change: float = 1e-9):
    """The reference's optimiser call (optimization.py:176-183, 279)."""
    opt = torch.optim.LBFGS(params, max_iter=max_iter, tolerance_grad=tolerance_grad,
                            tolerance_change=tolerance_change, lr=lr, line_search_fn="strong_wolfe")
    n_eval = [0]

    def closure():
        opt.zero_grad()
        loss = loss_fn()
        loss.backward()
        n_eval[0] += 1
        return loss

    opt.step(closure)
    return n_eval[0]
