"""Evaluation metrics on the CUDA path -- same names and argument meaning as
/root/reference/src/video_mocap/evaluation/metrics.py (the step right after the
fitting path).  Every function is one or two C-ABI calls (``csrc/metrics.cu``) on the current stream and returns a
0-dim CUDA tensor like the reference's ``torch.mean(...)``.  CUDA tensors only.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


def _req(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.UuoError("%s must be a CUDA tensor (uuo-mocap_b200 has no CPU path)" % name)
    return t.detach().float().contiguous()


def _point_error(a, b, ids: Optional[List[int]], velocity: bool, freq: float) -> torch.Tensor:
    a, b = _req(a, "pred"), _req(b, "gt")
    if a.shape != b.shape or a.dim() != 3 or a.shape[-1] != 3:
        raise ValueError("pred and gt must both be [F, P, 3]")
    F, P, _ = a.shape
    dev = a.device
    sel = torch.as_tensor(list(ids), dtype=torch.int32, device=dev) if ids is not None else None
    if sel is not None and (int(sel.min()) < 0 or int(sel.max()) >= P):
        raise IndexError("joint id out of range")
    scratch = torch.empty(1024, dtype=torch.float32, device=dev)
    out = torch.empty(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(_lib.lib().uuo_point_error_mean(ptr(a), ptr(b), F, P, ptr(sel), 0 if sel is None else sel.numel(),
                                              1 if velocity else 0, float(freq), ptr(scratch), ptr(out), stream_ptr(dev)))
    return out[0]


def compute_MPJPE(pred_joints, gt_joints):
    """metrics.py:49-54."""
    return _point_error(pred_joints, gt_joints, None, False, 1.0)


def compute_MPJPE_joints(pred_joints, gt_joints, joints_ids: List[int]):
    """metrics.py:57-63."""
    return _point_error(pred_joints, gt_joints, joints_ids, False, 1.0)


def compute_MPJVE(pred_joints, gt_joints, freq: float):
    """metrics.py:65-73."""
    return _point_error(pred_joints, gt_joints, None, True, freq)


def compute_MPJVE_joints(pred_joints, gt_joints, freq: float, joints_ids: List[int]):
    """metrics.py:76-85."""
    return _point_error(pred_joints, gt_joints, joints_ids, True, freq)


def compute_V2V(pred_vertices, gt_vertices):
    """metrics.py:126-131."""
    return _point_error(pred_vertices, gt_vertices, None, False, 1.0)


def compute_similarity_transform(S1: torch.Tensor, S2: torch.Tensor) -> torch.Tensor:
    """metrics.py:138-190: S1 [B,N,3] after the similarity transform (scale, rotation, translation) closest to S2."""
    S1, S2 = _req(S1, "S1"), _req(S2, "S2")
    if S1.shape != S2.shape or S1.dim() != 3 or S1.shape[-1] != 3:
        raise ValueError("S1 and S2 must both be [B, N, 3]")
    out = torch.empty_like(S1)
    with torch.cuda.device(S1.device):
        check(_lib.lib().uuo_similarity_transform(ptr(S1), ptr(S2), S1.shape[0], S1.shape[1], ptr(out),
                                                  stream_ptr(S1.device)))
    return out


def compute_PA_MPJPE(pred_joints, gt_joints):
    """metrics.py:88-94."""
    return _point_error(compute_similarity_transform(pred_joints, gt_joints), gt_joints, None, False, 1.0)


def compute_PA_MPJPE_joints(pred_joints, gt_joints, joints_ids: List[int]):
    """metrics.py:97-104."""
    return _point_error(compute_similarity_transform(pred_joints, gt_joints), gt_joints, joints_ids, False, 1.0)


def compute_PA_MPJVE(pred_joints, gt_joints, freq: float):
    """metrics.py:107-116."""
    return _point_error(compute_similarity_transform(pred_joints, gt_joints), gt_joints, None, True, freq)


def compute_PA_MPJVE_joints(pred_joints, gt_joints, freq: float, joints_ids: List[int]):
    """metrics.py:119-129."""
    return _point_error(compute_similarity_transform(pred_joints, gt_joints), gt_joints, joints_ids, True, freq)


def point_mesh_distance(points: torch.Tensor, vertices: torch.Tensor, faces: torch.Tensor) -> torch.Tensor:
    """dist [F,M]: unsigned distance from every point to the closest triangle of the frame's mesh.
    points [F,M,3], vertices [F,V,3], faces [NF,3] (shared) or [F,NF,3] (the reference repeats them per frame)."""
    points, vertices = _req(points, "points"), _req(vertices, "vertices")
    if faces.dim() == 3:
        faces = faces[0]
    if not faces.is_cuda:
        raise _lib.UuoError("faces must be a CUDA tensor")
    faces = faces.to(torch.int32).contiguous()
    F, M, _ = points.shape
    V, NF = vertices.shape[1], faces.shape[0]
    if int(faces.min()) < 0 or int(faces.max()) >= V:
        raise IndexError("face index out of range")
    dist = torch.empty((F, M), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        check(_lib.lib().uuo_point_mesh_distance(ptr(points), ptr(vertices), ptr(faces), F, M, V, NF, ptr(dist),
                                                 stream_ptr(points.device)))
    return dist


def compute_marker_to_surface_distance(vertices, faces, markers):
    """metrics.py:26-46: mean over frames and markers of |signed distance to the mesh| (igl.signed_distance there)."""
    return point_mesh_distance(markers, vertices, faces).mean()
