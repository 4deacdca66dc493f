"""The stage order of ``multimodal_video_mocap`` (/root/reference/src/video_mocap/multimodal.py:184-708) around the
fitting path, on tensors: [root stage] -> yaw hypotheses (chamfer stage -> correspondences -> marker stage, best by
weighted chamfer distance) -> final correspondence + marker rounds -> the reference's output dict.

What stays with the caller (outside the path, SURVEY.md section 8 scope table): reading the capture / the video prior
(``Markers``, ``ImgSmpl``), temporal alignment (``pad``), marker segmentation (``segment_rigid``), rendering.  The
part-fit and reprojection stages (``parts.find_best_part_fits``, ``reprojection.optim_reprojection``) need the
segmentation labels / the HMR camera and are called by the caller before this function when it has them.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch

from . import fit
from .root import optim_root
from .stageii import resample_prior


def fit_capture(markers: torch.Tensor, pose_body: torch.Tensor, betas: torch.Tensor, root_orient: torch.Tensor,
                trans: torch.Tensor, smpl_inference, config: Dict, img_mask: Optional[torch.Tensor] = None,
                marker_labels: Optional[torch.Tensor] = None, mocap_frame_rate: float = 120.0,
                save_stages: bool = True, verbose: bool = False, img_frame_rate: Optional[float] = None) -> Dict:
    """markers [F,M,3]; the video prior: pose_body [F,23,3,3], betas [1,10], root_orient [F,1,3,3], trans [F,3] (the
    ``o_*`` tensors of multimodal.py:172-186).  Returns the dict of multimodal.py:680-706 (``betas`` repeated per frame,
    rotations normalised, ``stages`` with numpy snapshots) plus ``scores`` / ``best_angle`` of the hypothesis loop.
    ``img_frame_rate``: rate of the video prior when it differs from ``mocap_frame_rate``; the prior is then resampled
    to the marker rate first (multimodal.py:144-182)."""
    smpl = fit._model_of(smpl_inference)
    dev = smpl.device
    if img_frame_rate is not None and img_frame_rate != mocap_frame_rate:
        n_src = torch.as_tensor(trans).shape[0]
        trans, root_orient, pose_body, _ = resample_prior(torch.as_tensor(trans), torch.as_tensor(root_orient),
                                                          torch.as_tensor(pose_body), torch.as_tensor(trans),
                                                          img_frame_rate, mocap_frame_rate)
        if img_mask is not None:
            # the mask follows the prior to the marker rate: output frame i reads source frame int(i * img / mocap), the
            # frame resample_prior starts its blend from (the reference leaves the mask at the video rate, which only
            # lines up when the rates agree)
            im = torch.as_tensor(img_mask).reshape(-1)
            if im.shape[0] != n_src:
                raise ValueError("img_mask must have one entry per frame of the video prior")
            src = torch.tensor([min(int(i * (img_frame_rate / mocap_frame_rate)), n_src - 1)
                                for i in range(trans.shape[0])], dtype=torch.long, device=im.device)
            img_mask = im[src]
    T = lambda t: torch.as_tensor(t).detach().to(dev, torch.float32).contiguous()
    mk = torch.nan_to_num(T(markers), nan=0.0)                                          # :188
    F = min(mk.shape[0], trans.shape[0])                                                # :191-199
    mk, o_pose, o_betas = mk[:F].contiguous(), T(pose_body)[:F].contiguous(), T(betas).reshape(1, -1)
    root, tr = T(root_orient)[:F].clone(), T(trans)[:F].clone()
    bet = o_betas.clone()
    mask = torch.ones(F, device=dev) if img_mask is None else T(img_mask).reshape(-1)[:F].contiguous()
    if mask.shape[0] != F:
        raise ValueError("img_mask holds %d frames, the capture %d" % (mask.shape[0], F))
    snap = lambda p: {k: v.detach().cpu().numpy() for k, v in p.items()}
    stages = {}

    st_root = config["stages"].get("root", {"num_iters": 0})
    if st_root.get("num_iters", 0) > 0:                                                 # :437-455
        labels = marker_labels if marker_labels is not None else torch.zeros(mk.shape[:2], dtype=torch.long)
        optim_root(mk, o_pose, bet, root, tr, labels, smpl, config, verbose=verbose)
        stages["root"] = snap({"trans": tr, "root_orient": fit.normalize_rot(root), "betas": bet[0],
                               "pose_body": fit.normalize_rot(o_pose)})

    n_ang = int(config.get("num_root_orient_angles", 1))                                # :462
    angles = [2.0 * math.pi * i / n_ang for i in range(n_ang)]
    best, scores, fitted = fit.fit_hypotheses(mk, o_pose.clone(), o_pose, bet, o_betas, root, tr, mask, smpl, config, angles)
    p = fitted[best]                                                                    # :601-608
    pose, root, tr, bet = p["pose_body"], p["root_orient"], p["trans"], p["betas"]
    if config["stages"]["chamfer"]["num_iters"] > 0:                                    # :497-503 (test/test.py:131-143)
        c = p["chamfer"]
        stages["chamfer"] = snap({"trans": c["trans"], "root_orient": c["root_orient"], "betas": c["betas"][0],
                                  "pose_body": c["pose_body"]})
    if config["stages"]["marker"]["num_iters"] > 0:
        stages["marker"] = snap({"trans": tr, "root_orient": root, "betas": bet[0], "pose_body": pose})

    labels_out = marker_labels
    for _ in range(int(config.get("stage_repeats", 0))):                                # :610-678
        pose_stage = pose.detach().clone()
        if config["stages"]["marker"]["num_iters"] > 0:
            vidx = fit.compute_nearest_points(mk, pose, bet, root, tr, smpl, img_mask=mask, config=config,
                                              as_one_hot=False)
            if config.get("recompute_marker_labels", False):
                labels_out = fit.compute_marker_labels_from_coords(smpl, vidx, F)
            fit.optim_markers(mk, pose, pose_stage, bet, o_betas, root, tr, vidx, mask, smpl, config, verbose=verbose)
        root, pose = fit.normalize_rot(root), fit.normalize_rot(pose)
        stages["marker_final"] = snap({"trans": tr, "root_orient": root, "betas": bet[0], "pose_body": pose})

    out = {"trans": tr.detach().cpu(), "root_orient": fit.normalize_rot(root).detach().cpu(),
           "pose_body": fit.normalize_rot(pose).detach().cpu(),
           "betas": bet.mean(dim=0, keepdim=True).repeat_interleave(F, dim=0).detach().cpu(),
           "mocap_frame_rate": mocap_frame_rate, "mocap_markers": mk.detach().cpu().numpy(), "markers_labels": labels_out,
           "scores": scores, "best_angle": angles[best]}
    if save_stages:
        out["stages"] = stages
    return out
