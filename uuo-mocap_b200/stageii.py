"""Output side of the fitting path: the ``*_stageii.npz`` files the reference writes after a fit
(/root/reference/src/video_mocap/test/test.py:115-143; SURVEY.md section 8(f) row N2, writer half).

Keys (AMASS / MoSh++ stage-II layout): ``betas [10]``, ``trans [F,3]``, ``poses [F,72]`` axis-angle (root first),
``mocap_frame_rate``, ``mocap_markers [F,M,3]``, ``gender``; one extra file per intermediate stage named
``<name>_stageii.<stage>.npz`` with that stage's ``trans`` / ``betas`` / ``poses``.

Host-side format code: tensors may live on the GPU, they are converted on their device and copied out once.
Input side (N2, reader half): ``resample_prior`` brings the video prior to the marker frame rate
(multimodal.py:144-182).  Not built: the PHALP/HMR ``.pkl`` and ``.c3d`` file parsers (``ezc3d`` is not installable here).
"""
from __future__ import annotations

from typing import Dict, Mapping, Optional

import numpy as np
import torch


def matrix_to_quaternion(m: torch.Tensor) -> torch.Tensor:
    """[...,3,3] -> unit quaternions [...,4] (w, x, y, z) with w >= 0.  Same construction as
    pytorch3d.transforms.matrix_to_quaternion (the best-conditioned of the four candidates), followed by
    standardize_quaternion."""
    m = m.to(torch.float32)
    m00, m01, m02 = m[..., 0, 0], m[..., 0, 1], m[..., 0, 2]
    m10, m11, m12 = m[..., 1, 0], m[..., 1, 1], m[..., 1, 2]
    m20, m21, m22 = m[..., 2, 0], m[..., 2, 1], m[..., 2, 2]
    q_abs = torch.sqrt(torch.clamp(torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                                                1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], dim=-1), min=0.0))
    cand = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1),
    ], dim=-2)
    cand = cand / (2.0 * q_abs[..., None].clamp(min=0.1))
    best = torch.argmax(q_abs, dim=-1)
    q = torch.gather(cand, -2, best[..., None, None].expand(best.shape + (1, 4)))[..., 0, :]
    return torch.where(q[..., :1] < 0, -q, q)


def matrix_to_axis_angle(m: torch.Tensor) -> torch.Tensor:
    """[...,3,3] -> axis-angle [...,3], angle in [0, pi] (pytorch3d.transforms.matrix_to_axis_angle =
    quaternion_to_axis_angle(matrix_to_quaternion(.)), call site test/test.py:119)."""
    q = matrix_to_quaternion(m)
    norms = torch.norm(q[..., 1:], p=2, dim=-1, keepdim=True)
    angles = 2.0 * torch.atan2(norms, q[..., :1])
    small = angles.abs() < 1e-6
    s = torch.where(small, 0.5 - angles * angles / 48.0, torch.sin(0.5 * angles) / torch.where(small, torch.ones_like(angles), angles))
    return q[..., 1:] / s


def quaternion_to_matrix(q: torch.Tensor) -> torch.Tensor:
    """(w, x, y, z) [...,4] -> [...,3,3] (pytorch3d.transforms.quaternion_to_matrix: scaled by 2 / |q|^2)."""
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def unitquat_slerp(q0: torch.Tensor, q1: torch.Tensor, alpha: torch.Tensor) -> torch.Tensor:
    """Shortest-arc slerp of unit quaternions [...,4], ``alpha`` broadcastable to [...,1].

    The reference calls ``roma.utils.unitquat_slerp`` (q0 * exp(alpha * log(q0^-1 q1)), multimodal.py:168-169) on
    pytorch3d's WXYZ quaternions although roma reads XYZW.  The curve is the constant-speed great arc of S^3 from q0
    to q1 (flipped when <q0, q1> < 0), which depends on the 4-vectors only and not on the component order, so the
    closed form below gives the same rotations (``roma`` is not installable here: checked against scipy's Slerp)."""
    d = (q0 * q1).sum(-1, keepdim=True)
    q1 = torch.where(d < 0, -q1, q1)
    omega = torch.acos(d.abs().clamp(max=1.0))
    so = torch.sin(omega)
    small = so < 1e-6
    den = torch.where(small, torch.ones_like(so), so)
    w0 = torch.where(small, 1.0 - alpha, torch.sin((1.0 - alpha) * omega) / den)
    w1 = torch.where(small, alpha + torch.zeros_like(so), torch.sin(alpha * omega) / den)
    q = w0 * q0 + w1 * q1
    return q / q.norm(dim=-1, keepdim=True)


def resample_prior(trans: torch.Tensor, root_orient: torch.Tensor, pose_body: torch.Tensor,
                   foot_contacts: torch.Tensor, img_freq: float, mocap_freq: float):
    """Video prior at ``img_freq`` -> the marker rate ``mocap_freq`` (multimodal.py:144-182): output frame i reads
    source frame int(i * img_freq / mocap_freq) and the next one, translations and foot contacts blended linearly,
    rotations ([F,1,3,3], [F,J-1,3,3]) by quaternion slerp; past the last source frame the frame is repeated.
    Returns (trans, root_orient, pose_body, foot_contacts) with round(F * mocap_freq / img_freq) frames; the
    inputs are returned untouched when the rates agree (:145)."""
    if mocap_freq == img_freq:
        return trans, root_orient, pose_body, foot_contacts
    n_src = trans.shape[0]
    n_new = round(n_src * (mocap_freq / img_freq))
    pos = [i * (img_freq / mocap_freq) for i in range(n_new)]                   # python doubles, like :152-153
    frame = [int(p) for p in pos]
    dev = trans.device
    f0 = torch.tensor(frame, dtype=torch.long, device=dev)
    inside = f0 + 1 < n_src
    f1 = torch.where(inside, f0 + 1, f0)
    alpha = torch.tensor([p - f for p, f in zip(pos, frame)], dtype=torch.float32, device=dev)
    alpha = torch.where(inside, alpha, torch.zeros_like(alpha))

    def lerp(x):
        a = alpha.reshape((-1,) + (1,) * (x.dim() - 1)).to(x.dtype)
        return torch.where(inside.reshape(a.shape), x[f1] * a + x[f0] * (1.0 - a), x[f0])

    def slerp(r):
        q = unitquat_slerp(matrix_to_quaternion(r[f0]), matrix_to_quaternion(r[f1]), alpha.reshape(-1, 1, 1))
        return torch.where(inside.reshape(-1, 1, 1, 1), quaternion_to_matrix(q).to(r.dtype), r[f0])

    return lerp(trans), slerp(root_orient), slerp(pose_body), lerp(foot_contacts)


def _np(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


def poses_axis_angle(root_orient: torch.Tensor, pose_body: torch.Tensor) -> np.ndarray:
    """root_orient [F,1,3,3] + pose_body [F,J-1,3,3] -> poses [F, 3J] (test/test.py:118-119)."""
    rot = torch.cat((torch.as_tensor(root_orient), torch.as_tensor(pose_body)), dim=1)
    return _np(torch.flatten(matrix_to_axis_angle(rot), start_dim=1))


def stage_filename(output_filename: str, stage: str) -> str:
    """test/test.py:139-140 (the replacement is applied to the file name only, not to its directories)."""
    import os
    head, tail = os.path.split(output_filename)
    return os.path.join(head, tail.replace("_stageii", "_stageii." + stage))


def save_stageii(output_filename: str, betas, trans, root_orient, pose_body, mocap_frame_rate, mocap_markers,
                 gender: str = "neutral", stages: Optional[Mapping[str, Mapping]] = None) -> Dict[str, np.ndarray]:
    """Write the fit like test/test.py:115-143.  ``betas`` [1,10] or [10]; ``stages``: {name: {"trans", "betas",
    "root_orient", "pose_body"}} for the per-stage files.  Returns the main file's dict."""
    b = _np(betas)
    out = {
        "betas": b[0] if b.ndim == 2 else b,                                    # :116 multimodal_smpl["betas"][0]
        "trans": _np(trans),
        "poses": poses_axis_angle(root_orient, pose_body),
        "mocap_frame_rate": np.asarray(mocap_frame_rate),
        "mocap_markers": _np(mocap_markers),
        "gender": gender,
    }
    np.savez(output_filename, **out)
    for name, st in (stages or {}).items():                                     # :129-141
        rec = dict(out)
        rec["trans"] = _np(st["trans"])
        rec["betas"] = _np(st["betas"])
        rec["poses"] = poses_axis_angle(torch.as_tensor(st["root_orient"]), torch.as_tensor(st["pose_body"]))
        np.savez(stage_filename(output_filename, name), **rec)
    return out


def load_stageii(filename: str) -> Dict[str, np.ndarray]:
    with np.load(filename, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}
