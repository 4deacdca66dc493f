"""Reprojection stage: ``optim_reprojection`` of /root/reference/src/video_mocap/utils/hmr_utils.py:170-425 on the
CUDA path (the stage right before the fitting path, ``num_iters: 0`` in the shipped
yaml).

For one yaw hypothesis the stage places the HMR 2.0 body in front of a single camera: it optimises the yaw about the
camera's vertical axis [1], the per-frame body translation [F,3] and one camera position [3] so that (a) the 45
projected SMPL joints fall on the 2-D joints HMR predicted (perspective camera recovered from the PHALP box
parameters, :57-127) and (b) the body, swapped into the mocap axes, lies on the markers (single-directional chamfer).

Every closure evaluation is two dense CUDA lbs() forward/backward pairs (``SMPL.forward``'s autograd function over
uuo_lbs_forward / uuo_lbs_backward) and the CUDA correspondence kernel (``losses.chamfer_distance``); the camera
algebra is a handful of [F,45]-sized torch ops on the device.  L-BFGS / strong Wolfe like the reference, or Adam.
One deviation: the reference reports joints / translations from the LAST closure evaluation of the line search
(``nonlocal`` variables, :296-307); here they are recomputed at the final parameters.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

from .fit import _model_of
from .losses import chamfer_distance

FOCAL_LENGTH = 5000.0    # HMR 2.0 default, hmr_utils.py:67
IMG_SIZE = 256.0         # :72


def perspective_projection(points: torch.Tensor, translation: torch.Tensor, focal_length: torch.Tensor,
                           camera_center: Optional[torch.Tensor] = None,
                           rotation: Optional[torch.Tensor] = None) -> torch.Tensor:
    """hmr_utils.py:14-52: points [B,N,3], translation [B,3], focal_length [B,2], camera_center [B,2],
    rotation [B,3,3] -> pixels [B,N,2].  K = [[fx,0,cx],[0,fy,cy],[0,0,1]] applied to (x/z, y/z, 1)."""
    if rotation is not None:
        points = points @ rotation.transpose(1, 2)
    p = points + translation.unsqueeze(1)
    xy = p[..., :2] / p[..., 2:3]
    out = xy * focal_length.unsqueeze(1)
    if camera_center is not None:
        out = out + camera_center.unsqueeze(1)
    return out


def convert_hmr_pos_to_mocap_pos(pos: torch.Tensor) -> torch.Tensor:
    """:130-136."""
    return torch.stack((pos[..., 0], pos[..., 2], -pos[..., 1]), dim=-1)


def convert_mocap_pos_to_hmr_pos(pos: torch.Tensor) -> torch.Tensor:
    """:139-145."""
    return torch.stack((pos[..., 0], -pos[..., 2], pos[..., 1]), dim=-1)


def rot_y(angle: torch.Tensor) -> torch.Tensor:
    """compute_root_orient_y (optimization.py:660-668): [...,1] -> [...,3,3]."""
    c, s = torch.cos(angle[..., 0]), torch.sin(angle[..., 0])
    z, o = torch.zeros_like(c), torch.ones_like(c)
    return torch.stack([c, z, s, z, o, z, -s, z, c], dim=-1).reshape(angle.shape[:-1] + (3, 3))


def get_3d_parameters(smpl_inference, pred_smpl_betas, pred_smpl_body_pose, pred_smpl_global_orient, pred_cam, center,
                      size, scale) -> Dict[str, torch.Tensor]:
    """hmr_utils.py:57-127: camera translation from the PHALP box parameters and the SMPL joints it projects."""
    smpl = _model_of(smpl_inference)
    N = pred_cam.shape[0]
    dev = pred_cam.device
    side = size.max(dim=-1, keepdim=True)[0]                                             # the padded square crop
    top, left = (side - size[:, [0]]) // 2, (side - size[:, [1]]) // 2
    ratio = 1.0 / torch.round(side) * IMG_SIZE
    center = (center + torch.cat((left, top), dim=-1)) * ratio
    scale = scale * side * ratio
    focal = torch.full((N, 2), FOCAL_LENGTH, device=dev, dtype=pred_cam.dtype)
    joints = smpl(betas=pred_smpl_betas, body_pose=pred_smpl_body_pose, global_orient=pred_smpl_global_orient,
                  transl=torch.zeros(N, 3, device=dev), pose2rot=False).joints
    tz = 2 * focal[:, 0] / (pred_cam[:, 0] * scale[:, 0] + 1e-9)
    cam_t = torch.cat((pred_cam[:, 1:3] + (center - IMG_SIZE / 2.0) * tz[:, None] / focal, tz[:, None]), dim=1)
    cc = torch.zeros(N, 2, device=dev)
    kp = perspective_projection(joints, cam_t, focal / IMG_SIZE, cc)
    kp = (kp + 0.5) * IMG_SIZE
    return {"camera_center": cc, "focal_length": focal / IMG_SIZE, "pred_cam_t": cam_t, "pred_joints": joints,
            "pred_keypoints_2d_smpl": kp / IMG_SIZE,
            "rotation": torch.eye(3, device=dev).unsqueeze(0).expand(N, -1, -1)}


def reprojection_setup(smpl, markers, pose_body, hmr_betas, root_orient, trans, pred_cam, cam_center, cam_size,
                       cam_scale) -> Dict[str, torch.Tensor]:
    """hmr_utils.py:222-268: 2-D targets, frame mask, initial body translations and camera position."""
    with torch.no_grad():
        j = get_3d_parameters(smpl, hmr_betas, pose_body, root_orient, pred_cam, cam_center, cam_size, cam_scale)
        cam_t = j["pred_cam_t"]
        mask = (cam_t == cam_t).float().mean(dim=-1)                                     # frames without a detection are NaN
        smpl_t = torch.nan_to_num(cam_t, 0)                                              # body <-> camera swap, :249-252
        offset = convert_mocap_pos_to_hmr_pos(torch.median(markers.reshape(-1, 3), dim=0, keepdim=True)[0]) - \
            torch.median(smpl_t, dim=0, keepdim=True)[0]
        return {"target": torch.nan_to_num(j["pred_keypoints_2d_smpl"], 0), "mask": mask, "smpl_trans0": smpl_t + offset,
                "cam_single0": (trans - offset).mean(dim=0, keepdim=True),
                "focal_length": j["focal_length"].mean(dim=0, keepdim=True), "camera_center": j["camera_center"]}


_CORRECTION = ((1.0, 0.0, 0.0), (0.0, 0.0, 1.0), (0.0, -1.0, 0.0))       # HMR axes -> mocap axes, :212-216


def _place(st, root_orient, y_angle, smpl_trans, cam_single):
    """Body placement of one closure evaluation (:297-307): yawed root, body translation rotated about the camera."""
    F = root_orient.shape[0]
    cam = cam_single.expand(F, 3)
    y_root = rot_y(y_angle.reshape(1, 1, 1).expand(F, 1, 1)) @ root_orient
    back = rot_y(-y_angle.reshape(1, 1))[0]                                              # [3,3]
    inv = (smpl_trans - cam) @ back.T + cam
    return cam, y_root, inv


def _project(smpl, st, pose_body, betas_rows, root_orient, inv, cam):
    F = pose_body.shape[0]
    joints = smpl(betas=betas_rows, body_pose=pose_body, global_orient=root_orient, transl=inv, pose2rot=False).joints
    return perspective_projection(joints, cam, st["focal_length"].expand(F, 2), st["camera_center"]) + 0.5


def reprojection_loss(smpl, st: Dict, markers, pose_body, betas_rows, root_orient, y_angle, smpl_trans, cam_single,
                      weights: Dict[str, float]):
    """The closure's loss (:293-329).  y_angle [1], smpl_trans [F,3], cam_single [1,3] carry gradients.
    -> (loss, reproject_loss, chamfer_loss, joints_2d [F,45,2])."""
    cam, y_root, inv = _place(st, root_orient, y_angle, smpl_trans, cam_single)
    kp = _project(smpl, st, pose_body, betas_rows, root_orient, inv, cam)                # note: the UN-yawed root, :309-314
    l_rep = torch.mean((kp - st["target"]) ** 2 * st["mask"][:, None, None]) * float(weights["reprojection"])
    corr = torch.tensor(_CORRECTION, device=root_orient.device)
    verts = smpl(betas=betas_rows, body_pose=pose_body, global_orient=corr @ y_root,
                 transl=convert_hmr_pos_to_mocap_pos(smpl_trans), pose2rot=False).vertices
    l_ch = chamfer_distance(markers, verts, single_directional=True, model=smpl.handle)[0] * float(weights["chamfer"])
    return l_rep + l_ch, l_rep, l_ch, kp


def optim_reprojection(markers, pose_body, betas, hmr_betas, root_orient, trans, pred_cam, cam_center, cam_size,
                       cam_scale, angle, img_mask, smpl_inference, num_iters: int, config: Dict, verbose: bool = False,
                       iter_fn: Optional[Callable] = None) -> Dict:
    """hmr_utils.py:170-425, same arguments and output dict (leading hypothesis axis of length 1 kept)."""
    smpl = _model_of(smpl_inference)
    dev = smpl.device
    T = lambda t: torch.as_tensor(t).detach().to(dev, torch.float32).contiguous()
    mk, pose, root, tr = T(markers), T(pose_body), T(root_orient), T(trans)
    F = pose.shape[0]
    betas_rows = T(betas).reshape(1, -1).expand(F, -1).contiguous()                      # detached in the reference, :202,286
    st = reprojection_setup(smpl, mk, pose, T(hmr_betas), root, tr, T(pred_cam), T(cam_center), T(cam_size), T(cam_scale))
    angle0 = float(angle)
    y_angle = torch.full((1,), angle0, device=dev, requires_grad=True)
    smpl_trans = st["smpl_trans0"].clone().requires_grad_(True)
    cam_single = st["cam_single0"].clone().requires_grad_(True)
    weights = config["stages"]["reprojection_part"]["losses"]
    oc = config.get("optimizer", {})
    corr = torch.tensor(_CORRECTION, device=dev)
    it = [0]

    def evaluate():
        loss, l_rep, l_ch, kp = reprojection_loss(smpl, st, mk, pose, betas_rows, root, y_angle, smpl_trans, cam_single, weights)
        if verbose:
            print("Reprojection", it[0], float(loss), float(l_rep), angle0 * 57.29577951308232, float(y_angle) * 57.29577951308232)
        if iter_fn is not None:
            with torch.no_grad():
                _, y_root, _ = _place(st, root, y_angle, smpl_trans, cam_single)
                iter_fn(stage="reprojection", iteration=it[0], pose_body=pose.cpu().numpy(), betas=betas_rows.cpu().numpy(),
                        trans=convert_hmr_pos_to_mocap_pos(smpl_trans).cpu().numpy(), root_orient=(corr @ y_root).cpu().numpy(),
                        pred_angle=float(y_angle), initial_angle=angle0, pred_2d_joints=kp.detach().cpu().numpy(),
                        gt_2d_joints=st["target"].cpu().numpy())
        it[0] += 1
        return loss

    params = [y_angle, smpl_trans, cam_single]
    if int(num_iters) > 0:
        if str(oc.get("name", "lbfgs")).lower() == "lbfgs":                              # :281-288
            opt = torch.optim.LBFGS(params, max_iter=int(num_iters), tolerance_grad=float(oc.get("tolerance_grad", 1e-7)),
                                    tolerance_change=float(oc.get("tolerance_change", 1e-9)), lr=1.0,
                                    line_search_fn="strong_wolfe")

            def closure():
                opt.zero_grad()
                loss = evaluate()
                loss.backward()
                return loss
            opt.step(closure)
        else:
            opt = torch.optim.Adam(params, lr=float(oc.get("lr", 0.01)))
            for _ in range(int(num_iters)):
                opt.zero_grad()
                loss = evaluate()
                loss.backward()
                opt.step()
    with torch.no_grad():                                                                # :349-425
        cam, y_root, inv = _place(st, root, y_angle, smpl_trans, cam_single)
        kp = _project(smpl, st, pose, betas_rows, root, inv, cam)
        world_root = corr @ y_root
        verts_world = smpl(betas=betas_rows, body_pose=pose, global_orient=world_root, transl=inv, pose2rot=False).vertices
        reproject_error = float(torch.mean((kp - st["target"]) ** 2 * st["mask"][:, None, None]))
        chamfer_error = float(chamfer_distance(mk, verts_world, single_directional=True, model=smpl.handle)[0])
    return {
        "pose_body": pose[None].clone(), "betas": betas_rows[None].clone(), "root_orient": world_root[None].clone(),
        "trans": convert_hmr_pos_to_mocap_pos(smpl_trans.detach())[None].clone(), "joints_2d": kp[None].clone(),
        "joints_2d_gt": st["target"][None], "cam_trans": convert_hmr_pos_to_mocap_pos(cam)[None].clone(),
        "camera_center": st["camera_center"].clone(), "focal_length": st["focal_length"].clone(),
        "reproject_mask": st["mask"].clone(), "input_angle": angle0, "output_angle": float(y_angle),
        "metrics": {"chamfer": chamfer_error, "reproject": reproject_error},
    }
