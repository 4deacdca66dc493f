"""Oracle restatement of the pytorch3d ops the reference calls (TEST INFRASTRUCTURE, CPU).

PARITY UNPINNED: pytorch3d is an un-vendored, un-pinned dependency of the reference
(``/root/reference/install.sh:7`` -> 0.7.4/0.7.5 for torch 2.0.1+cu117) and cannot be
installed in this offline image, and the reference holds no golden vector for these ops.
What is below restates the published pytorch3d 0.7.x algorithms:

* ``pytorch3d.transforms.rotation_conversions``: ``matrix_to_rotation_6d``,
  ``rotation_6d_to_matrix``, ``axis_angle_to_quaternion``, ``quaternion_to_matrix``,
  ``axis_angle_to_matrix``;
* ``pytorch3d.transforms.so3``: ``so3_relative_angle`` / ``so3_rotation_angle`` (defaults ``cos_angle=False``,
  ``cos_bound=1e-4``, ``eps=1e-4``) on ``pytorch3d.transforms.math.acos_linear_extrapolation``;
* ``pytorch3d.ops.knn_points`` (K=1, squared L2, candidates scanned in index order with a
  strict ``<`` so the FIRST minimum wins, distance accumulated as
  ``((dx*dx) + dy*dy) + dz*dz`` in fp32 with separately rounded products, which is what the
  C++ CPU implementation ``knn_cpu.cpp`` computes);
* ``pytorch3d.loss.chamfer_distance`` with ``weights``, ``single_directional=True``,
  ``point_reduction="mean"``, ``batch_reduction="mean"``.

Reference call sites: ``src/video_mocap/losses/chamfer_distance.py:5-21``,
``src/video_mocap/optimization.py:66-74,192-202,232-239,336-338,662-679``,
``src/video_mocap/markers/markers_utils.py:471-475,575-579``.
Cross-checks that stand in for the missing pin: scipy ``Rotation`` for the conversions,
and the reference's own numpy ``closest_point`` (``src/video_mocap/utils/points.py:19-42``)
for the nearest-vertex indices (``tests/test_oracle_p3d.py``).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------
# rotation conversions
# ------------------------------------------------------------------------------------------
def acos_linear_extrapolation(x: torch.Tensor, bounds: Tuple[float, float]) -> torch.Tensor:
    """pytorch3d.transforms.math.acos_linear_extrapolation: arccos inside [lo, hi], its first-order Taylor
    expansion at the bound outside (keeps the gradient finite at +-1)."""
    import math
    lo, hi = bounds
    if lo > hi or lo <= -1.0 or hi >= 1.0:
        raise ValueError("bad extrapolation bounds")
    out = torch.empty_like(x)
    below, above = x <= lo, x >= hi
    inside = (~below) & (~above)
    out[inside] = torch.acos(x[inside])
    taylor = lambda v, x0: math.acos(x0) + (v - x0) * (-1.0 / math.sqrt(1.0 - x0 * x0))
    out[below] = taylor(x[below], lo)
    out[above] = taylor(x[above], hi)
    return out


def so3_rotation_angle(R: torch.Tensor, eps: float = 1e-4, cos_angle: bool = False, cos_bound: float = 1e-4):
    """pytorch3d.transforms.so3.so3_rotation_angle: angle of a batch of rotations from the trace."""
    if R.dim() != 3 or R.shape[1:] != (3, 3):
        raise ValueError("Input has to be a batch of 3x3 Tensors.")
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    if ((tr < -1.0 - eps) + (tr > 3.0 + eps)).any():
        raise ValueError("A matrix has trace outside valid range [-1-eps,3+eps].")
    phi_cos = (tr - 1.0) * 0.5
    if cos_angle:
        return phi_cos
    if cos_bound > 0.0:
        b = 1.0 - cos_bound
        return acos_linear_extrapolation(phi_cos, (-b, b))
    return torch.acos(phi_cos)


def so3_relative_angle(R1: torch.Tensor, R2: torch.Tensor, cos_angle: bool = False, cos_bound: float = 1e-4,
                       eps: float = 1e-4) -> torch.Tensor:
    """pytorch3d.transforms.so3.so3_relative_angle: angle of R1 R2^T (call site optimization.py:66-74)."""
    return so3_rotation_angle(torch.bmm(R1, R2.permute(0, 2, 1)), eps=eps, cos_angle=cos_angle, cos_bound=cos_bound)


def matrix_to_rotation_6d(matrix: torch.Tensor) -> torch.Tensor:
    """First two ROWS of the matrix, flattened."""
    return matrix[..., :2, :].clone().reshape(matrix.shape[:-2] + (6,))


def rotation_6d_to_matrix(d6: torch.Tensor) -> torch.Tensor:
    """Gram-Schmidt; F.normalize divides by max(norm, 1e-12); rows stacked."""
    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = F.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def normalize_rot(m: torch.Tensor) -> torch.Tensor:
    """src/video_mocap/multimodal.py:34-35 -- the 6D round trip used everywhere."""
    return rotation_6d_to_matrix(matrix_to_rotation_6d(m))


def axis_angle_to_quaternion(axis_angle: torch.Tensor) -> torch.Tensor:
    angles = torch.norm(axis_angle, p=2, dim=-1, keepdim=True)
    half = angles * 0.5
    eps = 1e-6
    small = angles.abs() < eps
    k = torch.empty_like(angles)
    k[~small] = torch.sin(half[~small]) / angles[~small]
    k[small] = 0.5 - (angles[small] * angles[small]) / 48
    return torch.cat([torch.cos(half), axis_angle * k], dim=-1)


def quaternion_to_matrix(q: torch.Tensor) -> torch.Tensor:
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack(
        (
            1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
            two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
            two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j),
        ),
        -1,
    )
    return o.reshape(q.shape[:-1] + (3, 3))


def axis_angle_to_matrix(axis_angle: torch.Tensor) -> torch.Tensor:
    return quaternion_to_matrix(axis_angle_to_quaternion(axis_angle))


def compute_root_orient_z(angle: torch.Tensor) -> torch.Tensor:
    """src/video_mocap/optimization.py:672-679 -- [...,1] yaw angle -> [...,3,3]."""
    z = torch.zeros(list(angle.shape[:-1]) + [3], dtype=angle.dtype, device=angle.device)
    z[..., [2]] = angle
    return axis_angle_to_matrix(z)


# ------------------------------------------------------------------------------------------
# nearest neighbour, chamfer
# ------------------------------------------------------------------------------------------
def sqdist_matrix(p1: torch.Tensor, p2: torch.Tensor) -> torch.Tensor:
    """[N,P1,3],[N,P2,3] -> [N,P1,P2]; ((dx*dx)+(dy*dy))+(dz*dz), each op rounded to fp32."""
    d = p1[:, :, None, :] - p2[:, None, :, :]
    sq = d * d
    return (sq[..., 0] + sq[..., 1]) + sq[..., 2]


class _KnnK1(torch.autograd.Function):
    """knn_points(K=1): forward dists/idx, backward of pytorch3d's knn_points_backward
    (grad_p1 = 2*g*(p1-p2[idx]); grad_p2[idx] -= that)."""

    @staticmethod
    def forward(ctx, p1, p2, chunk):
        N, P1, _ = p1.shape
        dists = torch.empty(N, P1, dtype=p1.dtype, device=p1.device)
        idx = torch.empty(N, P1, dtype=torch.long, device=p1.device)
        for s in range(0, N, chunk):
            e = min(N, s + chunk)
            d = sqdist_matrix(p1[s:e], p2[s:e])
            mn, am = torch.min(d, dim=-1)     # CPU torch.min returns the first minimal index
            # make the tie-break explicit instead of trusting the backend
            first = (d == mn[..., None]).to(torch.uint8).argmax(dim=-1)
            dists[s:e] = mn
            idx[s:e] = first
        ctx.save_for_backward(p1, p2, idx)
        ctx.mark_non_differentiable(idx)
        return dists, idx

    @staticmethod
    def backward(ctx, g_dists, _g_idx):
        p1, p2, idx = ctx.saved_tensors
        nn = torch.gather(p2, 1, idx[..., None].expand(-1, -1, 3))
        diff = 2.0 * g_dists[..., None] * (p1 - nn)
        g1 = diff
        g2 = torch.zeros_like(p2)
        g2.scatter_add_(1, idx[..., None].expand(-1, -1, 3), -diff)
        return g1, g2, None


def knn_k1(p1: torch.Tensor, p2: torch.Tensor, chunk: int = 64) -> Tuple[torch.Tensor, torch.Tensor]:
    return _KnnK1.apply(p1, p2, chunk)


def chamfer_distance(x: torch.Tensor, y: torch.Tensor, weights: Optional[torch.Tensor] = None,
                     single_directional: bool = False, chunk: int = 64):
    """pytorch3d.loss.chamfer_distance, full clouds, point_reduction=batch_reduction="mean"."""
    N, P1, _ = x.shape
    P2 = y.shape[1]
    if weights is not None:
        if weights.shape[0] != N:
            raise ValueError("weights must be of shape (N,).")
        if not (weights >= 0).all():
            raise ValueError("weights cannot be negative.")
        if weights.sum() == 0.0:
            w = weights.view(N, 1)
            return (x.sum((1, 2)) * w[:, 0]).sum() * 0.0, None
    cham_x, idx_x = knn_k1(x, y, chunk)
    if weights is not None:
        cham_x = cham_x * weights.view(N, 1)
    cham_x = cham_x.sum(1) / float(max(P1, 1))
    div = weights.sum() if weights is not None else max(N, 1)
    cham_x = cham_x.sum() / div
    if single_directional:
        return cham_x, None
    cham_y, _ = knn_k1(y, x, chunk)
    if weights is not None:
        cham_y = cham_y * weights.view(N, 1)
    cham_y = cham_y.sum(1) / float(max(P2, 1))
    cham_y = cham_y.sum() / div
    return cham_x + cham_y, None


def weighted_chamfer_distance(x: torch.Tensor, y: torch.Tensor, x_weights: torch.Tensor,
                              single_directional: bool = False, return_idx: bool = False):
    """src/video_mocap/losses/chamfer_distance.py:5-21 without materialising [F*M,V,3]:
    every (frame, marker) is its own cloud of ONE query point against that frame's vertices,
    weight = the marker mask, always single-directional."""
    Fn, M, _ = x.shape
    w = torch.flatten(x_weights)
    if w.sum() == 0.0:
        return x.sum() * 0.0, None
    d, idx = knn_k1(x, y)                       # [F,M]
    cham = (d.reshape(-1) * w).sum() / w.sum()  # P1 == 1 -> the point mean is a no-op
    if return_idx:
        return cham, idx
    return cham, None
