"""Tensor-level wrappers of the correspondence / loss / optimiser entry points of libuuomocap.

Each function mirrors a reference call site (cited) and is one C-ABI call on the current CUDA
stream.  Inputs must be CUDA tensors; there is no CPU fallback.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

MARKER_DISTANCE = 0.0095   # /root/reference/src/video_mocap/utils/settings.py:1


def _req(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.UuoError("%s must be a CUDA tensor (uuo-mocap_b200 has no CPU path)" % name)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def nn_argmin(markers: torch.Tensor, vertices: torch.Tensor, model=None,
              order: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """pytorch3d knn_points(K=1) as used by chamfer_distance
    (src/video_mocap/losses/chamfer_distance.py:5-21): markers [F,M,3], vertices [F,V,3]
    -> (dist2 [F,M] fp32, idx [F,M] int32; first minimum wins).  With ``model`` (a ModelHandle whose V
    matches) or an explicit ``order`` (int32 CUDA permutation of the vertex ids) the box-pruned kernel runs --
    identical outputs, a fraction of the arithmetic when consecutive entries of the order are spatially close."""
    markers, vertices = _req(markers, "markers"), _req(vertices, "vertices")
    F, M, _ = markers.shape
    V = vertices.shape[1]
    if vertices.shape[0] != F:
        raise ValueError("markers and vertices must have the same number of frames")
    d = torch.empty((F, M), dtype=torch.float32, device=markers.device)
    i = torch.empty((F, M), dtype=torch.int32, device=markers.device)
    with torch.cuda.device(markers.device):
        if order is not None:
            if not (order.is_cuda and order.dtype == torch.int32 and order.numel() == V and order.is_contiguous()):
                raise ValueError("order must be a contiguous int32 CUDA tensor with V entries")
            check(_lib.lib().uuo_nn_argmin_ordered(ptr(markers), ptr(vertices), ptr(order), F, M, V, ptr(d), ptr(i),
                                                   stream_ptr(markers.device)))
        elif model is not None and model.V == V:
            check(_lib.lib().uuo_nn_argmin_ordered(ptr(markers), ptr(vertices), model.vertex_order, F, M, V, ptr(d),
                                                   ptr(i), stream_ptr(markers.device)))
        else:
            check(_lib.lib().uuo_nn_argmin(ptr(markers), ptr(vertices), F, M, V, ptr(d), ptr(i),
                                           stream_ptr(markers.device)))
    return d, i


def nn_argmin_presorted(markers: torch.Tensor, vertices: torch.Tensor, ids: Optional[torch.Tensor] = None,
                        want_pos: bool = False, weights: Optional[torch.Tensor] = None,
                        hint: Optional[torch.Tensor] = None):
    """uuo_nn_argmin_presorted: ``vertices`` [F,V,3] already lie in visiting order (a twin model's output),
    ``ids`` [V] int32 gives the original id of every row.  Returns (dist2, idx) or (dist2, idx, pos).
    With ``weights`` [F,M] and/or ``hint`` [F,M] int32 (uuo_nn_argmin_presorted_fit): markers of weight 0 are not
    searched and report (0, ids[0], 0) -- they multiply by 0 in weighted_chamfer_distance and in its gradient; the
    hint (a guess of ``pos``, e.g. the previous iteration's) only tightens the pruning, never changes the outputs."""
    markers, vertices = _req(markers, "markers"), _req(vertices, "vertices")
    F, M, _ = markers.shape
    V = vertices.shape[1]
    if vertices.shape[0] != F:
        raise ValueError("markers and vertices must have the same number of frames")
    if ids is not None and not (ids.is_cuda and ids.dtype == torch.int32 and ids.numel() == V and ids.is_contiguous()):
        raise ValueError("ids must be a contiguous int32 CUDA tensor with V entries")
    dev = markers.device
    d = torch.empty((F, M), dtype=torch.float32, device=dev)
    i = torch.empty((F, M), dtype=torch.int32, device=dev)
    p = torch.empty((F, M), dtype=torch.int32, device=dev) if want_pos else None
    with torch.cuda.device(dev):
        if weights is not None or hint is not None:
            if weights is not None:
                weights = _req(weights, "weights")
                if weights.shape != (F, M):
                    raise ValueError("weights must be [F, M]")
            if hint is not None and not (hint.is_cuda and hint.dtype == torch.int32 and hint.shape == (F, M)
                                         and hint.is_contiguous()):
                raise ValueError("hint must be a contiguous int32 CUDA tensor [F, M]")
            check(_lib.lib().uuo_nn_argmin_presorted_fit(ptr(markers), ptr(vertices), ptr(ids), ptr(weights), ptr(hint),
                                                         F, M, V, ptr(d), ptr(i), ptr(p), stream_ptr(dev)))
        else:
            check(_lib.lib().uuo_nn_argmin_presorted(ptr(markers), ptr(vertices), ptr(ids), F, M, V, ptr(d), ptr(i), ptr(p),
                                                     stream_ptr(dev)))
    return (d, i, p) if want_pos else (d, i)


def get_marker_mask(markers: torch.Tensor) -> torch.Tensor:
    """src/video_mocap/optimization.py:703-715 (float 0/1 instead of bool)."""
    markers = _req(markers, "markers")
    mask = torch.empty(markers.shape[:-1], dtype=torch.float32, device=markers.device)
    with torch.cuda.device(markers.device):
        check(_lib.lib().uuo_marker_mask(ptr(markers), mask.numel(), ptr(mask), stream_ptr(markers.device)))
    return mask


def mean_distance_argmin(markers: torch.Tensor, vertices: torch.Tensor, frame_mask: Optional[torch.Tensor] = None,
                         acc: Optional[torch.Tensor] = None, finalize: bool = True, count: Optional[float] = None):
    """compute_nearest_points(use_mean) (optimization.py:464-486,595-603): returns (idx [M] int32, acc [M,V]).
    With ``finalize=False`` only accumulates (frame-range shards all-reduce ``acc`` and the count, then
    call again with F=0 frames or use ``mean_argmin``)."""
    markers, vertices = _req(markers, "markers"), _req(vertices, "vertices")
    F, M, _ = markers.shape
    V = vertices.shape[1]
    dev = markers.device
    if acc is None:
        acc = torch.zeros((M, V), dtype=torch.float32, device=dev)
    fm = _req(frame_mask, "frame_mask") if frame_mask is not None else None
    if fm is not None and fm.numel() != F:
        raise ValueError("frame_mask must hold one entry per frame (%d), got %d" % (F, fm.numel()))
    if vertices.shape[0] != F:
        raise ValueError("markers and vertices must have the same number of frames")
    with torch.cuda.device(dev):
        check(_lib.lib().uuo_mean_dist_accumulate(ptr(markers), ptr(vertices), ptr(fm), F, M, V, ptr(acc),
                                                  stream_ptr(dev)))
    if not finalize:
        return None, acc
    if count is None:
        count = float(F if fm is None else int((fm == 1).sum().item()))
    return mean_argmin(acc, max(count, 1.0)), acc


def mean_distance_candidate_sums(markers: torch.Tensor, vertices: torch.Tensor, cand: torch.Tensor,
                                 frame_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sums[m] (float64) = sum over the frames with frame_mask == 1 of ||vertices[f, cand[m]] - markers[f, m]||: the
    material for the bound of ``mean_distance_accumulate_bounded`` (uuo_mean_dist_candidate_sums)."""
    markers, vertices = _req(markers, "markers"), _req(vertices, "vertices")
    F, M, _ = markers.shape
    V = vertices.shape[1]
    if vertices.shape[0] != F:
        raise ValueError("markers and vertices must have the same number of frames")
    cand = cand.to(markers.device, torch.int32).contiguous()
    if cand.numel() != M:
        raise ValueError("one candidate vertex per marker (%d), got %d" % (M, cand.numel()))
    fm = _req(frame_mask, "frame_mask") if frame_mask is not None else None
    if fm is not None and fm.numel() != F:
        raise ValueError("frame_mask must hold one entry per frame (%d), got %d" % (F, fm.numel()))
    sums = torch.zeros((M,), dtype=torch.float64, device=markers.device)
    with torch.cuda.device(markers.device):
        check(_lib.lib().uuo_mean_dist_candidate_sums(ptr(markers), ptr(vertices), ptr(fm), F, M, V, ptr(cand),
                                                      ptr(sums), stream_ptr(markers.device)))
    return sums


def mean_bound_from_sums(sums: torch.Tensor, frames: int) -> torch.Tensor:
    """fp32 bound >= the in-order fp32 sum of the same terms: recursive summation of n non-negative terms is off by at
    most (n-1) u / (1 - (n-1) u) relative (u = 2^-24), the shards' partial sums add ``world`` more roundings; twice
    that plus 2^-10 is generous, and the result is rounded UP to fp32."""
    rel = 2.0 ** -10 + max(int(frames), 1) * 2.0 ** -23
    b = (sums.double() * (1.0 + rel)).to(torch.float32)
    return torch.nextafter(b, torch.full_like(b, float("inf")))


def mean_distance_accumulate_bounded(markers: torch.Tensor, vertices: torch.Tensor, bound: torch.Tensor,
                                     frame_mask: Optional[torch.Tensor] = None,
                                     acc: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``mean_distance_argmin(finalize=False)`` for callers that only need the per-marker argmin: entries whose running
    sum passes ``bound[m]`` stop being evaluated (uuo_mean_dist_accumulate_bounded; see include/uuo_mocap.h)."""
    markers, vertices = _req(markers, "markers"), _req(vertices, "vertices")
    F, M, _ = markers.shape
    V = vertices.shape[1]
    dev = markers.device
    if vertices.shape[0] != F:
        raise ValueError("markers and vertices must have the same number of frames")
    bound = _req(bound, "bound")
    if bound.numel() != M:
        raise ValueError("one bound per marker (%d), got %d" % (M, bound.numel()))
    if acc is None:
        acc = torch.zeros((M, V), dtype=torch.float32, device=dev)
    fm = _req(frame_mask, "frame_mask") if frame_mask is not None else None
    if fm is not None and fm.numel() != F:
        raise ValueError("frame_mask must hold one entry per frame (%d), got %d" % (F, fm.numel()))
    with torch.cuda.device(dev):
        check(_lib.lib().uuo_mean_dist_accumulate_bounded(ptr(markers), ptr(vertices), ptr(fm), F, M, V, ptr(bound),
                                                          ptr(acc), stream_ptr(dev)))
    return acc


def mean_argmin(acc: torch.Tensor, count: float) -> torch.Tensor:
    acc = _req(acc, "acc")
    M, V = acc.shape
    idx = torch.empty((M,), dtype=torch.int32, device=acc.device)
    with torch.cuda.device(acc.device):
        check(_lib.lib().uuo_mean_argmin(ptr(acc), M, V, float(count), ptr(idx), stream_ptr(acc.device)))
    return idx


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, step, betas=(0.9, 0.999), eps=1e-8):
    """torch.optim.Adam single-tensor update, in place, one fused kernel over the flat block."""
    for t in (param, grad, exp_avg, exp_avg_sq):
        if not (t.is_cuda and t.is_contiguous() and t.dtype == torch.float32):
            raise _lib.UuoError("adam_step needs contiguous fp32 CUDA tensors")
    with torch.cuda.device(param.device):
        check(_lib.lib().uuo_adam_step(ptr(param), ptr(grad), ptr(exp_avg), ptr(exp_avg_sq), param.numel(),
                                       float(lr), float(betas[0]), float(betas[1]), float(eps), int(step),
                                       stream_ptr(param.device)))


def adam_step_dev(param, grad, exp_avg, exp_avg_sq, lr, step_counter, betas=(0.9, 0.999), eps=1e-8):
    """adam_step with the step number in device memory (int32[1], incremented on the stream): graph-capturable."""
    for t in (param, grad, exp_avg, exp_avg_sq):
        if not (t.is_cuda and t.is_contiguous() and t.dtype == torch.float32):
            raise _lib.UuoError("adam_step needs contiguous fp32 CUDA tensors")
    if not (step_counter.is_cuda and step_counter.dtype == torch.int32 and step_counter.numel() == 1):
        raise _lib.UuoError("step_counter must be an int32 CUDA tensor with one element")
    with torch.cuda.device(param.device):
        check(_lib.lib().uuo_adam_step_dev(ptr(param), ptr(grad), ptr(exp_avg), ptr(exp_avg_sq), param.numel(),
                                           float(lr), float(betas[0]), float(betas[1]), float(eps), ptr(step_counter),
                                           stream_ptr(param.device)))


PROFILING = [False]


def profile_enable(on: bool):
    PROFILING[0] = bool(on)
    check(_lib.lib().uuo_profile_enable(1 if on else 0))


def profile_read():
    """{kernel name: (total_ms, launches)} of the scopes recorded since profile_enable(True)."""
    import ctypes as C
    l = _lib.lib()
    out = {}
    n = l.uuo_profile_read(-1, None, 0, None, None)
    for s in range(n):
        name = C.create_string_buffer(64)
        ms, cnt = C.c_float(0), C.c_int(0)
        l.uuo_profile_read(s, name, 64, C.byref(ms), C.byref(cnt))
        out[name.value.decode()] = (ms.value, cnt.value)
    return out
