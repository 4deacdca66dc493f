"""Differentiable point-cloud losses on the CUDA correspondence kernel.

``chamfer_distance`` replaces ``pytorch3d.loss.chamfer_distance`` as the reference calls it (full clouds,
no normals, ``point_reduction = batch_reduction = "mean"``): call sites
/root/reference/src/video_mocap/markers/markers_utils.py:471-475 (part fits, single-directional),
:575-579 (part-fit score, bidirectional) and optimization.py:697 (``chamfer_distance_by_part``).
The nearest neighbours come from ``uuo_nn_argmin`` (bit-exact first-minimum semantics); the backward is
pytorch3d's ``knn_points`` backward: d/dx = 2 g (x - y[idx]), d/dy[idx] -= the same.  Nothing of size
[N,P1,P2] is materialised.  CUDA tensors only.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops

MARKER_DISTANCE = ops.MARKER_DISTANCE


class _NearestSq(torch.autograd.Function):
    """dists[N,P1] = min_j |x_ni - y_nj|^2 with gradients to both clouds."""

    @staticmethod
    def forward(ctx, x, y, model):
        d, idx = ops.nn_argmin(x, y, model)
        ctx.save_for_backward(x, y, idx)
        return d

    @staticmethod
    def backward(ctx, g):
        x, y, idx = ctx.saved_tensors
        ix = idx.long().unsqueeze(-1).expand(-1, -1, 3)
        diff = 2.0 * g.unsqueeze(-1) * (x - torch.gather(y, 1, ix))
        gy = None
        if ctx.needs_input_grad[1]:
            gy = torch.zeros_like(y)
            gy.scatter_add_(1, ix, -diff)
        return (diff if ctx.needs_input_grad[0] else None), gy, None


def nearest_sq(x: torch.Tensor, y: torch.Tensor, model=None) -> torch.Tensor:
    return _NearestSq.apply(x.float().contiguous(), y.float().contiguous(), model)


def chamfer_distance(x: torch.Tensor, y: torch.Tensor, weights: Optional[torch.Tensor] = None,
                     single_directional: bool = False, model=None):
    """x [N,P1,3], y [N,P2,3], weights [N] or None -> (loss, None), pytorch3d semantics:
    cham_x = sum_n w_n * mean_i min_j |x_ni - y_nj|^2 / (sum_n w_n  or  N); + the y->x term unless
    ``single_directional``.  ``model``: a ModelHandle whose V == P2 switches the x->y search to the pruned kernel."""
    if not (x.is_cuda and y.is_cuda):
        raise ops._lib.UuoError("chamfer_distance needs CUDA tensors (uuo-mocap_b200 has no CPU path)")
    N, P1, _ = x.shape
    P2 = y.shape[1]
    if y.shape[0] != N:
        raise ValueError("x and y must have the same batch size")
    if weights is not None:
        if weights.shape[0] != N:
            raise ValueError("weights must be of shape (N,).")
        if not bool((weights >= 0).all()):
            raise ValueError("weights cannot be negative.")
        if float(weights.sum()) == 0.0:
            return (x.sum((1, 2)) * weights.reshape(N)).sum() * 0.0, None
    div = weights.sum() if weights is not None else float(max(N, 1))
    cham_x = nearest_sq(x, y, model)
    if weights is not None:
        cham_x = cham_x * weights.view(N, 1)
    total = (cham_x.sum(1) / float(max(P1, 1))).sum() / div
    if not single_directional:
        cham_y = nearest_sq(y, x)
        if weights is not None:
            cham_y = cham_y * weights.view(N, 1)
        total = total + (cham_y.sum(1) / float(max(P2, 1))).sum() / div
    return total, None


def weighted_chamfer_distance(x: torch.Tensor, y: torch.Tensor, x_weights: torch.Tensor, model=None):
    """src/video_mocap/losses/chamfer_distance.py:5-21, differentiable in both clouds: every (frame, marker) is a cloud
    of one query point against its frame's vertices, weighted by ``x_weights`` [F,M] (the marker mask) and divided
    by their sum; always single-directional there.  -> (loss, None)."""
    if not (x.is_cuda and y.is_cuda):
        raise ops._lib.UuoError("weighted_chamfer_distance needs CUDA tensors (uuo-mocap_b200 has no CPU path)")
    w = x_weights.to(torch.float32)
    s = w.sum()
    if float(s) == 0.0:
        return x.sum() * 0.0, None
    return (nearest_sq(x, y, model) * w).sum() / s, None


def chamfer_distance_by_part(markers: torch.Tensor, vertices: torch.Tensor, marker_labels: torch.Tensor,
                             vertex_weights: torch.Tensor, single_directional: bool = False) -> torch.Tensor:
    """optimization.py:682-700 ("part_chamfer"): for every body part that owns markers (mode of the per-frame
    labels), (chamfer(vertices of the part -> markers of the part) - MARKER_DISTANCE)^2, summed over parts.
    markers [F,M,3], vertices [F,V,3] (differentiable), marker_labels [F,M] int, vertex_weights [V,P]."""
    vertex_mask = torch.argmax(vertex_weights, dim=-1)
    labels_mode = torch.mode(marker_labels, dim=0)[0]
    loss = 0
    for i in torch.unique(labels_mode).tolist():
        vertices_part = vertices[:, vertex_mask == i]
        markers_part = markers[:, labels_mode == i]
        if vertices_part.shape[1] == 0:
            continue
        part = chamfer_distance(vertices_part.contiguous(), markers_part.contiguous(),
                                single_directional=single_directional)[0]
        loss = loss + (part - MARKER_DISTANCE) ** 2
    return loss
