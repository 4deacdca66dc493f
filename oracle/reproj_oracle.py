"""Oracle restatement of the reprojection stage (TEST INFRASTRUCTURE, CPU, fp32).

Follows ``/root/reference/src/video_mocap/utils/hmr_utils.py``: ``perspective_projection`` :14-52,
``get_3d_parameters`` :57-127, the axis swaps :130-145, ``optim_reprojection`` :170-425 (set-up :222-283, closure
:293-345), on top of ``smpl_oracle`` (pinned to the vendored smplx) and ``p3d_oracle`` (pytorch3d restatement).
PARITY UNPINNED for the stage as a whole: the reference holds no test or golden vector for it and its inputs come
from PHALP/HMR 2.0 pickles that are not in the tree; the restatement is checked line by line against the source.
"""
from __future__ import annotations

from typing import Dict

import torch

from . import p3d_oracle as p3d
from .smpl_oracle import OracleSMPL

FOCAL_LENGTH = 5000.0   # hmr_utils.py:67
IMG_SIZE = 256.0        # :72


def perspective_projection(points, translation, focal_length, camera_center, rotation):
    """hmr_utils.py:14-52, with the same einsum contractions."""
    B = points.shape[0]
    K = torch.zeros([B, 3, 3], dtype=points.dtype)
    K[:, 0, 0] = focal_length[:, 0]
    K[:, 1, 1] = focal_length[:, 1]
    K[:, 2, 2] = 1.0
    K[:, :-1, -1] = camera_center
    pts = torch.einsum("bij,bkj->bki", rotation, points) + translation.unsqueeze(1)
    proj = pts / pts[:, :, -1].unsqueeze(-1)
    return torch.einsum("bij,bkj->bki", K, proj)[:, :, :-1]


def hmr_to_mocap(pos):
    """convert_hmr_pos_to_mocap_pos, :130-136: (x, y, z) -> (x, z, -y)."""
    return torch.cat((pos[..., [0]], pos[..., [2]], pos[..., [1]] * -1), dim=-1)


def mocap_to_hmr(pos):
    """convert_mocap_pos_to_hmr_pos, :139-145: (x, y, z) -> (x, -z, y)."""
    return torch.cat((pos[..., [0]], pos[..., [2]] * -1, pos[..., [1]]), dim=-1)


def compute_root_orient_y(angle):
    """optimization.py:660-668."""
    v = torch.zeros(list(angle.shape[:-1]) + [3], dtype=angle.dtype)
    v[..., [1]] = angle
    return p3d.axis_angle_to_matrix(v)


def get_3d_parameters(smpl: OracleSMPL, betas, body_pose, global_orient, pred_cam, center, size, scale) -> Dict:
    """hmr_utils.py:57-127 (PHALP's camera recovery)."""
    N = pred_cam.shape[0]
    h, w = size[:, [0]], size[:, [1]]
    new = torch.max(size, dim=-1, keepdim=True)[0]
    top, left = (new - h) // 2, (new - w) // 2
    ratio = 1.0 / torch.round(new) * IMG_SIZE
    center = (center + torch.cat((left, top), dim=-1)) * ratio
    scale = scale * new * ratio
    focal = FOCAL_LENGTH * torch.ones(N, 2, dtype=pred_cam.dtype)
    joints = smpl.forward(betas, body_pose, global_orient, torch.zeros(N, 3), pose2rot=False).joints
    tmp = torch.stack([pred_cam[:, 1], pred_cam[:, 2], 2 * focal[:, 0] / (pred_cam[:, 0] * scale[:, 0] + 1e-9)], dim=1)
    cam_t = torch.cat((tmp[:, :2] + (center - IMG_SIZE / 2.0) * tmp[:, [2]] / focal, tmp[:, [2]]), dim=1)
    cc = torch.zeros(N, 2)
    eye = torch.eye(3).unsqueeze(0).expand(N, -1, -1)
    kp = perspective_projection(joints, cam_t, focal / IMG_SIZE, cc, eye)
    kp = (kp + 0.5) * IMG_SIZE
    return {"camera_center": cc, "focal_length": focal / IMG_SIZE, "pred_cam_t": cam_t, "pred_joints": joints,
            "pred_keypoints_2d_smpl": kp / IMG_SIZE}


def reprojection_setup(smpl: OracleSMPL, markers, pose_body, hmr_betas, root_orient, trans, pred_cam, cam_center,
                       cam_size, cam_scale) -> Dict:
    """hmr_utils.py:222-268: targets, mask and the initial values of the optimised translations."""
    j = get_3d_parameters(smpl, hmr_betas, pose_body, root_orient, pred_cam, cam_center, cam_size, cam_scale)
    target = torch.nan_to_num(j["pred_keypoints_2d_smpl"], 0)                        # [F,45,2]
    cam_t = j["pred_cam_t"]
    mask = torch.mean((cam_t == cam_t).float(), dim=-1).detach()                      # [F]
    smpl_t = torch.nan_to_num(cam_t, 0)                                               # the body takes the camera's place
    cam_t = trans.clone().detach()                                                    # and the camera the body's (:249-252)
    offset = mocap_to_hmr(torch.median(markers.reshape(-1, 3), dim=0, keepdim=True)[0]) - \
        torch.median(smpl_t, dim=0, keepdim=True)[0]
    return {"target": target, "mask": mask, "smpl_trans0": (smpl_t + offset).detach(),
            "cam_single0": torch.mean(cam_t - offset, dim=0, keepdim=True).detach(),
            "focal_length": torch.mean(j["focal_length"], dim=0, keepdim=True), "camera_center": j["camera_center"]}


def reprojection_loss(smpl: OracleSMPL, st: Dict, markers, pose_body, betas_rows, root_orient, y_angle, smpl_trans,
                      cam_single, weights: Dict[str, float]):
    """The closure of optim_reprojection (hmr_utils.py:293-345) for one hypothesis.  y_angle [1,1,1,1],
    smpl_trans [1,F,3], cam_single [1,3] are the optimised tensors.  -> (loss, reproject_loss, chamfer_loss)."""
    Fn = pose_body.shape[0]
    cam = torch.repeat_interleave(cam_single[:, None], dim=1, repeats=Fn)            # [1,F,3]
    ya = torch.repeat_interleave(y_angle, repeats=Fn, dim=1)                          # [1,F,1,1]
    y_root = compute_root_orient_y(ya) @ root_orient                                  # [1,F,1,3,3]
    Rm = compute_root_orient_y(-ya)[:, :, 0]                                          # [1,F,3,3]
    inv = (Rm @ (smpl_trans - cam)[..., None])[..., 0] + cam                          # :305-307 (the 4x4 padding is a no-op)
    j1 = smpl.forward(betas_rows, pose_body, root_orient, inv[0], pose2rot=False).joints
    eye = torch.eye(3).unsqueeze(0).expand(Fn, -1, -1)
    kp = perspective_projection(j1, cam[0], torch.repeat_interleave(st["focal_length"], dim=0, repeats=Fn),
                                st["camera_center"], eye).reshape(1, Fn, 45, 2) + 0.5
    l_rep = torch.mean((kp - st["target"][None]) ** 2 * st["mask"][None, :, None, None]) * weights["reprojection"]
    corr = torch.tensor([[[[1.0, 0, 0], [0, 0, 1.0], [0, -1.0, 0]]]]).repeat(Fn, 1, 1, 1)
    v2 = smpl.forward(betas_rows, pose_body, (corr @ y_root)[0], hmr_to_mocap(smpl_trans)[0], pose2rot=False).vertices
    l_ch = p3d.chamfer_distance(markers, v2, single_directional=True)[0] * weights["chamfer"]
    return l_rep + l_ch, l_rep, l_ch
