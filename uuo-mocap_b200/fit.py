"""The video_mocap.yaml fitting loop on the CUDA path: stage closures, Adam / L-BFGS drivers and the
reference's stage functions (same names and argument meaning as
/root/reference/src/video_mocap/optimization.py).

A closure evaluation is a fixed sequence of C-ABI calls, no autograd graph and no [F,M,V] tensor:
  chamfer stage (optimization.py:187-275), on the visiting-order twin model (``SMPL.visiting_twin``):
      uuo_lbs_forward(RAW6D, yaw) -> uuo_nn_argmin_presorted_fit (marker mask as weights, the previous
      iteration's matches as hint) -> uuo_chamfer_rows -> uuo_lbs_cache_vposed_rows -> uuo_lbs_backward_rows_cached
      (v_posed rows taken from the dense forward's workspace) -> uuo_sqdiff_accumulate (video-pose prior, betas prior)
      [-> uuo_transvel_accumulate]
  marker stage (optimization.py:329-394):
      uuo_lbs_forward_rows -> uuo_marker_loss_rows -> uuo_lbs_backward_rows_cached -> priors
      [-> uuo_temporal_accumulate]
followed by uuo_adam_step(_dev) over the flat parameter block (north-star optimiser; with one process and no
per-iteration callback the whole iteration is replayed from a CUDA graph) or torch.optim.LBFGS (the
reference's optimiser) driving the same closure.

Frame-range sharding (``Shard``): every rank owns a contiguous frame range; per iteration the ranks
all-reduce ONE 16-float buffer (d/d betas + loss) and, only when a temporal term is enabled, exchange
a 1- or 2-frame halo with their neighbours (torch.distributed P2P: NCCL over NVLink on GPUs, gloo in
the CPU tests).  All ``mean`` normalisers use GLOBAL counts.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Optional

import numpy as np
import torch

from . import _lib, ops
from ._lib import check, ptr, stream_ptr
from .dist import Shard, exchange_halo_into, halo_buffers
from .smpl import POSE_RAW6D, SMPL, SmplInference

MARKER_DISTANCE = ops.MARKER_DISTANCE


def _model_of(smpl_inference) -> SMPL:
    return smpl_inference.smpl if isinstance(smpl_inference, SmplInference) else smpl_inference


class FitProblem:
    """Device buffers of one shard of one sequence (or of H hypotheses stacked along the frame axis)."""

    def __init__(self, smpl_inference, markers: torch.Tensor, o_pose_body: torch.Tensor, o_betas: torch.Tensor,
                 shard: Optional[Shard] = None, groups: int = 1, F_global: Optional[int] = None,
                 wsum_global: Optional[float] = None):
        self.smpl = _model_of(smpl_inference)
        self.mh = self.smpl.handle
        dev = self.mh.device
        self.dev = dev
        self.shard = shard or Shard()
        self.groups = groups                              # hypotheses stacked along frames (MHT)
        self.markers = markers.to(dev, torch.float32).contiguous()
        self.F, self.M = self.markers.shape[:2]
        if self.F % groups:
            raise ValueError("frames must divide evenly into hypothesis groups")
        self.o_pose_body = o_pose_body.to(dev, torch.float32).contiguous()
        self.o_betas = o_betas.to(dev, torch.float32).reshape(-1, self.mh.NB).contiguous()
        self.mask = ops.get_marker_mask(self.markers)
        # global normalisers (src/video_mocap/losses/chamfer_distance.py:20 divides by sum(weights))
        # (a caller that drives several shards from one process passes them in: tests/test_shard_gpu.py)
        if F_global is None or wsum_global is None:
            fg = torch.tensor([float(self.F // groups), float(self.mask.sum().item()) / groups], dtype=torch.float64)
            fg = self.shard.all_reduce_sum(fg)
            F_global, wsum_global = int(fg[0].item()), float(fg[1].item())
        self.F_global, self.wsum_global = int(F_global), float(wsum_global)
        self.ws = self.mh.workspace(self.F)
        self.vertices = None
        # dense forward of the chamfer stage runs on the visiting-order twin (rows of `vertices` are renumbered,
        # see SMPL.visiting_twin); idx holds ORIGINAL vertex ids, pos the matching row of `vertices`
        self.twin = self.smpl.visiting_twin()
        self.idx = torch.empty((self.F, self.M), dtype=torch.int32, device=dev)
        self.pos = torch.full((self.F, self.M), -1, dtype=torch.int32, device=dev)   # also the search hint of the next call
        self.d2 = torch.empty((self.F, self.M), dtype=torch.float32, device=dev)
        self.g_rows = torch.empty((self.F, self.M, 3), dtype=torch.float32, device=dev)
        self.virt = torch.empty((self.F, self.M, 3), dtype=torch.float32, device=dev)
        # one row per hypothesis group: [d/d betas (<=10) | pad | .]; the loss of the whole problem sits at [0, 15].
        # This block is what the shards sum every iteration (<= 1024 floats).
        if groups * 16 > 1024:
            raise ValueError("at most 64 stacked hypotheses per problem")
        self.red = torch.zeros((groups, 16), dtype=torch.float32, device=dev)
        self.loss_slot = self.red.view(-1)[15:16]
        self.gbetas_rows = None
        self.mm = None                                     # per-frame marker mean (trans_vel)
        self.tv_halo = None                                # (prev, next) [1,6] = [trans | marker mean] of the neighbour frame
        self.tp_halo = None                                # (prev, next) [2,D] pose rows of the neighbours (temporal)
        self.n_closure = 0

    # -- thin C-ABI call helpers -------------------------------------------------------------
    def _st(self):
        return stream_ptr(self.dev)

    def lbs_forward(self, betas_rows, root, pose, yaw, trans):
        if self.vertices is None:
            self.vertices = torch.empty((self.F, self.mh.V, 3), dtype=torch.float32, device=self.dev)
        check(_lib.lib().uuo_lbs_forward(self.twin.handle.handle, ptr(betas_rows), betas_rows.shape[0], ptr(root),
                                         ptr(pose), POSE_RAW6D, ptr(yaw), ptr(trans), self.F, ptr(self.vertices), None,
                                         ptr(self.ws), self._st()))
        return self.vertices

    def lbs_forward_rows(self, betas_rows, root, pose, yaw, trans, rows, shared):
        R = rows.shape[-1]
        check(_lib.lib().uuo_lbs_forward_rows(self.mh.handle, ptr(betas_rows), betas_rows.shape[0], ptr(root), ptr(pose),
                                              POSE_RAW6D, ptr(yaw), ptr(trans), self.F, ptr(rows), int(shared), R,
                                              ptr(self.virt), None, ptr(self.ws), self._st()))
        return self.virt

    def lbs_backward_rows(self, betas_rows, root, pose, yaw, rows, shared, g_rows, g_betas, g_root, g_pose, g_yaw, g_trans,
                          cached_pos=None, cached=False):
        """``cached``: the workspace holds v_posed of these rows (uuo_lbs_forward_rows just ran with them);
        ``cached_pos`` [F,R]: pick them out of the dense forward's v_posed first (chamfer stage)."""
        R = rows.shape[-1]
        fn = _lib.lib().uuo_lbs_backward_rows
        if R <= 64 and (cached or cached_pos is not None):
            if cached_pos is not None:
                check(_lib.lib().uuo_lbs_cache_vposed_rows(self.twin.handle.handle, self.F, ptr(cached_pos), R, ptr(self.ws),
                                                           self._st()))
            fn = _lib.lib().uuo_lbs_backward_rows_cached
        check(fn(self.mh.handle, ptr(betas_rows), betas_rows.shape[0], ptr(root), ptr(pose),
                                               POSE_RAW6D, ptr(yaw), self.F, ptr(rows), int(shared), R,
                                               ptr(g_rows), None, ptr(self.ws), ptr(g_betas), ptr(g_root), ptr(g_pose),
                                               ptr(g_yaw), ptr(g_trans), self._st()))

    def sqdiff(self, x, y, scale, g):
        check(_lib.lib().uuo_sqdiff_accumulate(ptr(x), ptr(y), x.numel(), float(scale), ptr(self.loss_slot), ptr(g),
                                               self._st()))

    def betas_rows(self, betas):
        """[G,NB] parameter -> rows handed to the kernels: one row per hypothesis group (betas_rows = G; G = 1: the row
        shared by all frames).  The kernels map frame f to row f // (F / G)."""
        return betas.reshape(self.groups, -1)

    def reduce_gbetas(self, g_betas_param):
        """d/d betas and the loss: the kernels leave one d/d betas row per hypothesis group; ONE sum over the shards of the
        [G,16] block (peer-memory kernel when the shard has a PeerComm, else torch.distributed)."""
        nb = self.mh.NB
        if self.groups > 1:
            self.red[:, :nb] = self.gbetas_rows
        self.shard.iter_all_reduce_sum_(self.red)
        g_betas_param.copy_(self.red[:, :nb].reshape(g_betas_param.shape))

    def gbetas_target(self):
        if self.groups == 1:
            return self.red
        if self.gbetas_rows is None:
            self.gbetas_rows = torch.empty((self.groups, self.mh.NB), dtype=torch.float32, device=self.dev)
        return self.gbetas_rows


@dataclass
class ParamBlock:
    """Flat fp32 parameter / gradient block with named views (one fused Adam launch over all of it)."""
    flat: torch.Tensor
    grad: torch.Tensor
    views: Dict[str, torch.Tensor]
    gviews: Dict[str, torch.Tensor]

    @classmethod
    def build(cls, tensors: Dict[str, torch.Tensor], device):
        n = sum(t.numel() for t in tensors.values())
        flat = torch.empty(n, dtype=torch.float32, device=device)
        grad = torch.zeros(n, dtype=torch.float32, device=device)
        views, gviews, off = {}, {}, 0
        for k, t in tensors.items():
            m = t.numel()
            # keep every view 16-byte aligned?  not required: the kernels only need 4-byte alignment here
            views[k] = flat[off:off + m].view(t.shape)
            gviews[k] = grad[off:off + m].view(t.shape)
            views[k].copy_(t.detach().to(device, torch.float32))
            off += m
        return cls(flat, grad, views, gviews)


def chamfer_closure(prob: FitProblem, pb: ParamBlock, root_orient: torch.Tensor, losses: Dict[str, float]) -> torch.Tensor:
    """closure_stage_chamfer (optimization.py:187-258): fills pb.grad, returns the loss (device scalar).
    Parameters: trans [F,3], z_angle, betas [G,10], pose_body [F,23,3,3]; root_orient is constant.
    ``z_angle`` [F] is the yaw of ``yaw_lock: True`` (:192, root = Rz(z) @ root_orient); a [F,1,3,3] ``z_angle`` is the
    free 3x3 of ``yaw_lock: False`` (:194), which the reference uses AS the root orientation inside the closure."""
    v, g = pb.views, pb.gviews
    F, M, Fg = prob.F, prob.M, prob.F_global
    pb.grad.zero_()
    prob.red.zero_()
    brows = prob.betas_rows(v["betas"])
    free_root = v["z_angle"].dim() == 4
    root_in, yaw_in = (v["z_angle"], None) if free_root else (root_orient, v["z_angle"])
    with torch.cuda.device(prob.dev):
        verts = prob.lbs_forward(brows, root_in, v["pose_body"], yaw_in, v["trans"])
        w_ch = float(losses.get("full_chamfer", 0.0))
        if w_ch != 0.0 and prob.wsum_global > 0:
            # dropped markers (mask 0) weigh nothing in the loss or the gradient: the kernel does not search them;
            # prob.pos still holds the previous iteration's matches: the kernel starts its search there (in == out)
            check(_lib.lib().uuo_nn_argmin_presorted_fit(ptr(prob.markers), ptr(verts), ptr(prob.twin.visit_ids),
                                                         ptr(prob.mask), ptr(prob.pos), F, M, prob.mh.V, ptr(prob.d2),
                                                         ptr(prob.idx), ptr(prob.pos), prob._st()))
            check(_lib.lib().uuo_chamfer_rows(ptr(prob.markers), ptr(verts), ptr(prob.mask), ptr(prob.pos), F, M,
                                              prob.mh.V, w_ch / prob.wsum_global, ptr(prob.loss_slot), ptr(prob.g_rows),
                                              prob._st()))
        else:
            prob.g_rows.zero_()
            prob.idx.zero_()
            prob.pos.zero_()
        prob.lbs_backward_rows(brows, root_in, v["pose_body"], yaw_in, prob.idx, False, prob.g_rows,
                               prob.gbetas_target(), g["z_angle"] if free_root else None, g["pose_body"],
                               None if free_root else g["z_angle"], g["trans"], cached_pos=prob.pos)
        if "reg_pose_body" in losses:                                                   # :243-245
            prob.sqdiff(v["pose_body"], prob.o_pose_body, losses["reg_pose_body"] / (Fg * v["pose_body"][0].numel()),
                        g["pose_body"])
        if "trans_vel" in losses:                                                       # :246-250
            if prob.groups > 1:
                raise ValueError("trans_vel is not defined across stacked hypotheses")
            if prob.mm is None:
                prob.mm = torch.empty((F, 3), dtype=torch.float32, device=prob.dev)
                check(_lib.lib().uuo_marker_mean(ptr(prob.markers), F, M, ptr(prob.mm), prob._st()))
                prob.tv_halo = halo_buffers(1, 6, prob.shard, prob.dev)
                exchange_halo_into(prob.mm, 1, prob.shard, *prob.tv_halo, col=3)        # constant: sent once
            prev, nxt = prob.tv_halo
            exchange_halo_into(v["trans"], 1, prob.shard, prev, nxt, col=0)
            check(_lib.lib().uuo_transvel_accumulate(ptr(v["trans"]), ptr(prob.mm), ptr(prev), ptr(nxt), F,
                                                     losses["trans_vel"] / ((Fg - 1) * 3), ptr(prob.loss_slot),
                                                     ptr(g["trans"]), prob._st()))
        prob.reduce_gbetas(g["betas"])
        if "reg_betas" in losses:                                                       # :255-256 (same on every rank)
            ob = prob.o_betas.expand_as(v["betas"]).contiguous()
            prob.sqdiff(v["betas"], ob, losses["reg_betas"] / v["betas"][0].numel(), g["betas"])
    prob.n_closure += 1
    return prob.loss_slot[0]


def marker_closure(prob: FitProblem, pb: ParamBlock, vertex_idx: torch.Tensor, losses: Dict[str, float]) -> torch.Tensor:
    """closure_stage_marker_pose (optimization.py:329-377).  Parameters: pose_body, betas, root_orient, trans."""
    v, g = pb.views, pb.gviews
    F, M, Fg = prob.F, prob.M, prob.F_global
    pb.grad.zero_()
    prob.red.zero_()
    brows = prob.betas_rows(v["betas"])
    with torch.cuda.device(prob.dev):
        # rows mode of the C-ABI: 1 = [M] for every frame; G = [G,M], one set per stacked hypothesis; 0 = [F,M] per frame
        if vertex_idx.dim() == 1:
            shared = 1
        elif prob.groups > 1 and vertex_idx.shape[0] == prob.groups and prob.groups != F:
            shared = prob.groups
        else:
            shared = 0
        virt = prob.lbs_forward_rows(brows, v["root_orient"], v["pose_body"], None, v["trans"], vertex_idx, shared)
        if "marker" in losses:                                                          # :356-362 (mean over ALL F*M)
            check(_lib.lib().uuo_marker_loss_rows(ptr(prob.markers), ptr(virt), ptr(prob.mask), F * M, MARKER_DISTANCE,
                                                  losses["marker"] / (Fg * M), ptr(prob.loss_slot), ptr(prob.g_rows),
                                                  prob._st()))
        else:
            prob.g_rows.zero_()
        prob.lbs_backward_rows(brows, v["root_orient"], v["pose_body"], None, vertex_idx, shared, prob.g_rows,
                               prob.gbetas_target(), g["root_orient"], g["pose_body"], None, g["trans"], cached=True)
        if "reg_pose_body" in losses:                                                   # :363-365
            prob.sqdiff(v["pose_body"], prob.o_pose_body, losses["reg_pose_body"] / (Fg * v["pose_body"][0].numel()),
                        g["pose_body"])
        if "temporal" in losses:                                                        # :368-374
            if prob.groups > 1:
                raise ValueError("the temporal term is not defined across stacked hypotheses")
            D = v["pose_body"][0].numel()
            x = v["pose_body"].view(F, D)
            if prob.tp_halo is None:
                prob.tp_halo = halo_buffers(2, D, prob.shard, prob.dev)
            prev, nxt = prob.tp_halo
            exchange_halo_into(x, 2, prob.shard, prev, nxt)
            check(_lib.lib().uuo_temporal_accumulate(ptr(x), ptr(prev), ptr(nxt), F, D,
                                                     losses["temporal"] / ((Fg - 2) * D), ptr(prob.loss_slot),
                                                     ptr(g["pose_body"]), prob._st()))
        prob.reduce_gbetas(g["betas"])
        if "reg_betas" in losses:                                                       # :366-367
            ob = prob.o_betas.expand_as(v["betas"]).contiguous()
            prob.sqdiff(v["betas"], ob, losses["reg_betas"] / v["betas"][0].numel(), g["betas"])
    prob.n_closure += 1
    return prob.loss_slot[0]


# ------------------------------------------------------------------------------------------------
# optimiser drivers
# ------------------------------------------------------------------------------------------------
_CAPTURE_STREAMS: Dict[torch.device, "torch.cuda.Stream"] = {}      # device -> side stream of the graph captures


def run_adam(pb: ParamBlock, closure: Callable[[], torch.Tensor], num_iters: int, lr: float,
             betas=(0.9, 0.999), eps: float = 1e-8, on_iter: Optional[Callable] = None,
             graph: Optional[bool] = None, warm: int = 3, shard: Optional[Shard] = None) -> Optional[torch.Tensor]:
    """closure -> fused Adam over the flat block.  No host synchronisation inside the loop unless ``on_iter`` asks
    for the loss.  After ``warm`` eager iterations ONE iteration (every kernel of the closure + the update, ~20-35
    launches) is captured in a CUDA graph and replayed for the rest: the step number lives in device memory
    (uuo_adam_step_dev), so nothing on the host changes between iterations.  ``graph=None``: on unless a callback
    wants the loss, the per-kernel profiler is running, or the closure's exchange with the other shards is issued
    from the host (``shard.graph_safe`` is False: torch.distributed transport instead of the peer-memory kernels)."""
    m = torch.zeros_like(pb.flat)
    v = torch.zeros_like(pb.flat)
    counter = torch.zeros(1, dtype=torch.int32, device=pb.flat.device)
    multi = torch.distributed.is_available() and torch.distributed.is_initialized()
    if graph is None:
        safe = shard.graph_safe if shard is not None else not multi
        graph = on_iter is None and not ops.PROFILING[0] and safe
    loss = None

    def one():
        out = closure()
        ops.adam_step_dev(pb.flat, pb.grad, m, v, lr, counter, betas, eps)
        return out

    it = 0
    eager = num_iters if not graph or num_iters <= warm + 1 else warm
    for _ in range(eager):
        loss = one()
        it += 1
        if on_iter is not None:
            on_iter(it - 1, loss)
    if it < num_iters:
        dev = pb.flat.device
        g = torch.cuda.CUDAGraph()
        with torch.cuda.device(dev):
            # capture_begin / capture_end directly: the torch.cuda.graph context manager first empties the caching
            # allocator and the pinned-host cache (cudaFree of every idle block: GBs at these sizes, blocking, and the
            # blocks have to be cudaMalloc'ed again), on every stage of every call
            side = _CAPTURE_STREAMS.get(dev)
            if side is None:
                side = _CAPTURE_STREAMS[dev] = torch.cuda.Stream(dev)
            cur = torch.cuda.current_stream(dev)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                # thread_local: NCCL's watchdog thread may touch the CUDA API while this thread captures
                g.capture_begin(capture_error_mode="thread_local" if multi else "global")
                try:
                    loss = one()
                finally:
                    g.capture_end()
            cur.wait_stream(side)
            while it < num_iters:        # the capture itself executed nothing
                g.replay()
                it += 1
            torch.cuda.current_stream(dev).synchronize()     # the graph (and its memory pool) dies with this frame
    return loss


def run_lbfgs(pb: ParamBlock, closure: Callable[[], torch.Tensor], max_iter: int, lr: float,
              tolerance_grad: float = 1e-7, tolerance_change: float = 1e-9) -> int:
    """The reference's optimiser call (optimization.py:176-183,279) around the CUDA closure."""
    p = pb.flat.requires_grad_(True)
    opt = torch.optim.LBFGS([p], max_iter=max_iter, tolerance_grad=tolerance_grad, tolerance_change=tolerance_change,
                            lr=lr, line_search_fn="strong_wolfe")
    n = [0]

    def cl():
        with torch.no_grad():
            loss = closure()
        p.grad = pb.grad
        n[0] += 1
        return loss.clone()

    opt.step(cl)
    pb.flat.requires_grad_(False)
    return n[0]


def _optimise(pb, closure, stage_cfg, opt_cfg, lbfgs_lr, on_iter=None, shard: Optional[Shard] = None):
    """``optimizer.name``: "lbfgs" (the reference's optimiser and the default when the yaml names none, like
    config/video_mocap.yaml) or "adam" (the fused north-star optimiser: ``default_config`` selects it)."""
    name = str(opt_cfg.get("name", "lbfgs")).lower()
    iters = int(stage_cfg["num_iters"])
    if iters <= 0:
        return
    if name == "adam":
        run_adam(pb, closure, iters, float(stage_cfg.get("lr_adam", opt_cfg.get("lr", 0.01))), on_iter=on_iter,
                 shard=shard)
    elif name == "lbfgs":
        if shard is not None and shard.world > 1:
            # torch's L-BFGS takes its direction, line-search and stopping decisions from the rank-local flat
            # gradient: shards would step the shared betas differently and call the closure's collectives a
            # different number of times.  Only the fused Adam path is defined across shards.
            raise NotImplementedError("optimizer.name = 'lbfgs' is not defined across frame-range shards; use 'adam'")
        run_lbfgs(pb, closure, iters, lbfgs_lr, float(opt_cfg.get("tolerance_grad", 1e-7)),
                  float(opt_cfg.get("tolerance_change", 1e-9)))
    else:
        raise ValueError("optimizer.name must be 'adam' or 'lbfgs'")


# ------------------------------------------------------------------------------------------------
# the reference's stage functions
# ------------------------------------------------------------------------------------------------
def get_marker_mask(markers: torch.Tensor) -> torch.Tensor:
    """optimization.py:703-715."""
    return ops.get_marker_mask(markers) != 0


def compute_marker_labels_from_coords(smpl_inference, barycentric_coords_one_hot: torch.Tensor, num_frames: int) -> torch.Tensor:
    """optimization.py:645-659: body-part label [F,M] of every marker = the joint with the largest skinning weight at
    the marker's vertex.  ``barycentric_coords_one_hot``: the [M,V] one-hot matrix or the vertex-index vector [M]."""
    smpl = _model_of(smpl_inference)
    vertex_ids = torch.argmax(smpl.lbs_weights, dim=-1)
    c = barycentric_coords_one_hot.to(smpl.device)
    coords_ids = torch.argmax(c, dim=-1) if c.dim() == 2 else c.long()
    return vertex_ids[coords_ids].unsqueeze(0).repeat_interleave(int(num_frames), dim=0)


def weighted_mse_loss(input: torch.Tensor, target: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
    """optimization.py:718-724."""
    return torch.mean(torch.nn.functional.mse_loss(input, target, reduction="none") * weights)


def weighted_chamfer_distance(x: torch.Tensor, y: torch.Tensor, x_weights: torch.Tensor,
                              single_directional: bool = False, return_idx: bool = False, model=None):
    """src/video_mocap/losses/chamfer_distance.py:5-21 (always single-directional there): value only, no
    [F*M,V,3] intermediate.  x markers [F,M,3], y vertices [F,V,3], x_weights [F,M]."""
    d, idx = ops.nn_argmin(x, y, model)
    w = x_weights.to(d.dtype).reshape(-1)
    s = w.sum()
    val = (d.reshape(-1) * w).sum() / s if float(s) != 0.0 else d.sum() * 0.0
    return (val, idx) if return_idx else (val, None)


def rot_z(angle: torch.Tensor) -> torch.Tensor:
    """compute_root_orient_z (optimization.py:672-679): [...,1] -> [...,3,3]."""
    c, s = torch.cos(angle[..., 0]), torch.sin(angle[..., 0])
    z, o = torch.zeros_like(c), torch.ones_like(c)
    return torch.stack([c, -s, z, s, c, z, z, z, o], dim=-1).reshape(angle.shape[:-1] + (3, 3))


def rot_z_apply(angle: torch.Tensor, m: torch.Tensor) -> torch.Tensor:
    """rot_z(angle) @ m written out row by row (Rz only mixes rows 0 and 1): elementwise kernels instead of a batched
    library GEMM of 3x3 matrices.  angle [...,1], m [...,3,3] (broadcast against each other)."""
    c, s = torch.cos(angle[..., 0])[..., None], torch.sin(angle[..., 0])[..., None]
    r0, r1, r2 = m[..., 0, :], m[..., 1, :], m[..., 2, :]
    a, b = c * r0 - s * r1, s * r0 + c * r1
    return torch.stack((a, b, r2.expand_as(a)), dim=-2)


def normalize_rot(m: torch.Tensor) -> torch.Tensor:
    """multimodal.py:34-35 (6D round trip) for the final packing step -- small torch ops on the device."""
    a1, a2 = m[..., 0, :], m[..., 1, :]
    b1 = torch.nn.functional.normalize(a1, dim=-1)
    b2 = torch.nn.functional.normalize(a2 - (b1 * a2).sum(-1, keepdim=True) * b1, dim=-1)
    return torch.stack((b1, b2, torch.cross(b1, b2, dim=-1)), dim=-2)


def optim_chamfer(markers, pose_body, o_pose_body, betas, o_betas, root_orient, trans, img_mask, marker_labels,
                  smpl_inference, config: Dict, initial_angle: float = 0, repeat: int = 0, verbose: bool = False,
                  iter_fn: Callable = None, shard: Optional[Shard] = None, groups: int = 1, problem=None):
    """optimization.py:147-285.  Updates trans, betas, pose_body and root_orient IN PLACE like the reference
    (its parameters are leaf tensors stepped by the optimiser; root_orient gets the fitted yaw baked in)."""
    st = config["stages"]["chamfer"]
    yaw_lock = bool(st.get("yaw_lock", True))
    prob = problem or FitProblem(smpl_inference, markers, o_pose_body, o_betas, shard, groups)
    F = prob.F
    dev = prob.dev
    root = root_orient.detach().to(dev, torch.float32).contiguous()
    # :165-172: a yaw angle per frame, or (yaw_lock: False) a free 3x3 per frame that starts at the identity
    z0 = torch.zeros(F, device=dev) if yaw_lock else torch.eye(3, device=dev).expand(F, 1, 3, 3).contiguous()
    pb = ParamBlock.build({"trans": trans, "z_angle": z0, "betas": betas.reshape(-1, prob.mh.NB),
                           "pose_body": pose_body}, dev)
    losses = dict(st["losses"])
    closure = lambda: chamfer_closure(prob, pb, root, losses)
    cb = None
    if verbose or iter_fn is not None:
        def cb(it, loss):
            if verbose:
                print("Chamfer", it, float(loss))
    _optimise(pb, closure, st, config.get("optimizer", {}), 0.1, cb, prob.shard)
    with torch.no_grad():
        trans.copy_(pb.views["trans"])
        betas.copy_(pb.views["betas"].view_as(betas))
        pose_body.copy_(pb.views["pose_body"])
        if yaw_lock:
            root_orient.copy_(rot_z_apply(pb.views["z_angle"].view(F, 1, 1), root))  # :280-281
        else:
            root_orient.copy_(normalize_rot(pb.views["z_angle"]) @ root)            # :282-283
    return prob


def optim_markers(markers, pose_body, o_pose_body, betas, o_betas, root_orient, trans, barycentric_coords_one_hot,
                  img_mask, smpl_inference, config: Dict, initial_angle: float = 0, repeat: int = 0,
                  verbose: bool = False, iter_fn: Callable = None, shard: Optional[Shard] = None, groups: int = 1,
                  problem=None):
    """optimization.py:288-399.  ``barycentric_coords_one_hot`` is the reference's [M,V] one-hot matrix or the
    equivalent vertex-index vector [M]."""
    st = config["stages"]["marker"]
    if st.get("use_sdf", False):
        raise NotImplementedError("use_sdf needs the reference's SDF checkpoints (not shipped)")
    prob = problem or FitProblem(smpl_inference, markers, o_pose_body, o_betas, shard, groups)
    dev = prob.dev
    vidx = barycentric_coords_one_hot
    if vidx.dim() == 2 and vidx.is_floating_point():      # the reference's [M,V] one-hot matrix
        vidx = torch.argmax(vidx, dim=1)
    elif vidx.dim() == 2 and tuple(vidx.shape) not in ((prob.F, prob.M), (prob.groups, prob.M)):
        raise ValueError("vertex indices must be [M], [groups, M] or [F, M]")
    vidx = vidx.to(dev, torch.int32).contiguous()
    pb = ParamBlock.build({"pose_body": pose_body, "betas": betas.reshape(-1, prob.mh.NB), "root_orient": root_orient,
                           "trans": trans}, dev)
    losses = dict(st["losses"])
    closure = lambda: marker_closure(prob, pb, vidx, losses)
    cb = None
    if verbose:
        cb = lambda it, loss: print("Marker", it, float(loss))
    _optimise(pb, closure, st, config.get("optimizer", {}), 1.0, cb, prob.shard)
    with torch.no_grad():
        pose_body.copy_(pb.views["pose_body"])
        betas.copy_(pb.views["betas"].view_as(betas))
        root_orient.copy_(pb.views["root_orient"])
        trans.copy_(pb.views["trans"])
    return prob


def compute_nearest_points(markers, pose_body, betas, root_orient, trans, smpl_inference, marker_labels=None,
                           granularity: str = "full", img_mask=None, device=None, config: Optional[Dict] = None,
                           o_pose_body=None, window_size: int = 1, use_velocity: bool = True,
                           shard: Optional[Shard] = None, as_one_hot: bool = True, problem: Optional[FitProblem] = None):
    """optimization.py:402-642.  The shipped ``use_mean: True`` branch (:422-430,464-486,595-603): SMPL forward with the
    mean betas, mean over the frames with img_mask == 1 of the marker-vertex distances, argmin.  ``use_mean: False``:
    per-frame ``closest_point`` (see ``_closest_point_last_frames``).  ``use_barycentric`` (igl / trimesh) is not taken.
    Returns the [M,V] one-hot matrix like the reference (or the index vector).  ``problem``: reuse that FitProblem's
    vertex buffer and workspace (the stages before and after hold them anyway) instead of allocating [F,V,3] again."""
    use_mean = True
    if config is not None:
        cl = config["stages"].get("compute_locations", {})
        if cl.get("use_barycentric", False):
            raise NotImplementedError("compute_locations.use_barycentric needs igl / trimesh (not on the CUDA path)")
        use_mean = bool(cl.get("use_mean", True))
    smpl = _model_of(smpl_inference)
    if use_mean:
        # the mean-distance kernel cuts (marker, vertex tile) pairs that can no longer win: that needs tiles of
        # neighbouring vertices, i.e. the visiting-order twin (same body, vertices renumbered; fit.FitProblem uses it too)
        smpl = smpl.visiting_twin()
    mh = smpl.handle
    dev = mh.device
    shard = shard or Shard()
    mk = markers.to(dev, torch.float32).contiguous()
    F, M = mk.shape[:2]
    mean_betas = torch.mean(betas.to(dev, torch.float32).reshape(-1, mh.NB), dim=0, keepdim=True).contiguous()   # :424
    if problem is not None and problem.F == F and problem.dev == dev and problem.mh.V == mh.V:
        if problem.vertices is None:
            problem.vertices = torch.empty((F, mh.V, 3), dtype=torch.float32, device=dev)
        vertices, ws = problem.vertices, problem.ws
    else:
        vertices = torch.empty((F, mh.V, 3), dtype=torch.float32, device=dev)
        ws = mh.workspace(F)
    pose = pose_body.detach().to(dev, torch.float32).contiguous()
    root = root_orient.detach().to(dev, torch.float32).contiguous()
    tr = trans.detach().to(dev, torch.float32).contiguous()
    with torch.cuda.device(dev):
        check(_lib.lib().uuo_lbs_forward(mh.handle, ptr(mean_betas), 1, ptr(root), ptr(pose), POSE_RAW6D, None, ptr(tr), F,
                                         ptr(vertices), None, ptr(ws), stream_ptr(dev)))
    fm = img_mask.to(dev, torch.float32).reshape(-1).contiguous() if img_mask is not None else None
    if fm is not None and fm.numel() != F:
        raise ValueError("img_mask must hold one entry per frame (%d), got %d" % (F, fm.numel()))
    if use_mean:
        acc, cnt = _mean_distance_sums(mk, vertices, fm, shard)
        ids = smpl.visit_ids.long()
        acc_model = torch.empty_like(acc)                                            # back to the model's numbering:
        acc_model[:, ids] = acc                                                      # the FIRST minimum wins there
        idx = ops.mean_argmin(acc_model, max(cnt, 1.0))
    else:
        idx = _closest_point_last_frames(mk, vertices, fm, marker_labels, granularity, pose.shape[1], shard)
    if not as_one_hot:
        return idx
    one_hot = torch.zeros((M, mh.V), dtype=torch.float32, device=dev)                # :601-603
    one_hot[torch.arange(M, device=dev), idx.long()] = 1.0
    return one_hot


BOUNDED_MIN_RATIO = 2.0      # probe sum / candidate sum from which the bounded mean-distance kernel is used


def _mean_distance_sums(mk, vertices, fm, shard):
    """[M,V] sums of the marker-vertex distances over the valid frames of ALL shards (optimization.py:479-486,596-599),
    exact wherever the per-marker minimum can be; returns (acc, number of valid frames).  One vertex per marker -- the
    nearest one in the first valid frame of the lowest rank that has one -- gives an upper bound of the minimum
    (uuo_mean_dist_candidate_sums, summed over the shards), which lets uuo_mean_dist_accumulate_bounded stop early on
    everything that is far away."""
    F, M = mk.shape[:2]
    dev = mk.device
    valid = torch.ones(F, dtype=torch.bool, device=dev) if fm is None else fm == 1
    n_valid = int(valid.sum().item())
    counts = torch.zeros(max(shard.world, 1), dtype=torch.float64)
    counts[shard.rank] = n_valid
    counts = shard.all_reduce_sum(counts)
    total = int(counts.sum().item())
    acc = torch.zeros((M, vertices.shape[1]), dtype=torch.float32, device=dev)
    if total == 0:
        return acc, 0.0
    owner = int((counts > 0).nonzero()[0].item())
    cand = torch.zeros(M, dtype=torch.float32, device=dev)
    if shard.rank == owner:
        f0 = int(valid.nonzero()[0].item())
        vi, _ = ops.mean_distance_argmin(mk[f0:f0 + 1].contiguous(), vertices[f0:f0 + 1].contiguous())
        cand = vi.to(torch.float32)
    shard.all_reduce_sum_(cand)                                                      # vertex ids < 2^24: exact
    cand = cand.to(torch.int32)
    # the candidates' sums, and as a probe the sums of vertices elsewhere on the body (the candidate of another marker)
    both = torch.cat([ops.mean_distance_candidate_sums(mk, vertices, cand, fm),
                      ops.mean_distance_candidate_sums(mk, vertices, torch.roll(cand, M // 2), fm)])
    shard.all_reduce_sum_(both)
    sums, probe = both[:M], both[M:]
    # The bound only cuts work when a marker column follows ONE point of the body, so that far vertices collect much
    # larger sums than the candidate.  Columns that hold a different marker in every frame (shuffled clouds) give all
    # vertices about the same sum: then nothing can be cut and the plain kernel (wider tiles, no checks) is faster.
    if M > 1 and float(torch.median(probe / sums.clamp_min(1e-30)).item()) >= BOUNDED_MIN_RATIO:
        bound = ops.mean_bound_from_sums(sums, total)
        ops.mean_distance_accumulate_bounded(mk, vertices, bound, fm, acc)
    else:
        ops.mean_distance_argmin(mk, vertices, fm, acc=acc, finalize=False)
    shard.all_reduce_sum_(acc)                                                       # [M,V] fp32, once per stage
    return acc, float(total)


def _closest_point_last_frames(mk, vertices, fm, marker_labels, granularity, num_joints, shard):
    """The ``use_mean: False`` branch of compute_nearest_points (optimization.py:503-510, 538-590) as the reference runs
    it (window_size = 1): every valid frame is its own window whose running minimum starts at infinity, so each frame's
    per-frame ``closest_point`` result (utils/points.py:19-42: argmin over vertices of the UN-squared distance)
    overwrites the previous one -- the output is the correspondence of the LAST frame with img_mask == 1 ("full" and
    "marker" granularity), or per marker column the last valid frame in which its label is one of the body joints
    ("part").  The frame's [M,V] distance matrix comes from the mean-distance kernel run on that one frame (same op
    order as np.linalg.norm, first minimum wins)."""
    F, M = mk.shape[:2]
    dev = mk.device
    valid = torch.ones(F, dtype=torch.bool, device=dev) if fm is None else fm == 1
    if granularity == "part":
        if marker_labels is None:
            raise ValueError("granularity 'part' needs marker_labels [F,M]")
        lab = torch.as_tensor(marker_labels).to(dev)
        ok = valid.view(F, 1) & (lab >= 0) & (lab < num_joints)
    else:
        ok = valid.view(F, 1).expand(F, M)
    fr = torch.arange(F, device=dev).view(F, 1).expand(F, M)
    last = torch.where(ok, fr, torch.full_like(fr, -1)).max(dim=0)[0]                 # [M] local frame, -1 = none
    # frame-range shards: the globally last frame wins; its owner contributes the indices
    offset = 0
    if shard.world > 1:
        sizes = torch.zeros(shard.world, dtype=torch.float64)
        sizes[shard.rank] = F
        sizes = shard.all_reduce_sum(sizes)
        offset = int(sizes[:shard.rank].sum().item())
    glast = torch.where(last >= 0, last + offset, last).to(torch.float32)
    gmax = glast.clone()
    if shard.world > 1:
        torch.distributed.all_reduce(gmax, op=torch.distributed.ReduceOp.MAX, group=shard.group)
    idx = torch.zeros(M, dtype=torch.float32, device=dev)
    mine = (glast == gmax) & (last >= 0)
    for f in torch.unique(last[mine]).tolist():
        cols = mine & (last == f)
        vi, _ = ops.mean_distance_argmin(mk[f:f + 1].contiguous(), vertices[f:f + 1].contiguous())
        idx = torch.where(cols, vi.to(torch.float32), idx)
    shard.all_reduce_sum_(idx)
    return idx.to(torch.int32)


def fit_hypotheses(markers, pose_body, o_pose_body, betas, o_betas, root_orient, trans, img_mask, smpl_inference,
                   config: Dict, angles, shard: Optional[Shard] = None):
    """The MHT loop of multimodal.py:462-606 batched: every yaw hypothesis becomes a block of frames of ONE
    stacked problem (chamfer stage -> correspondences -> marker stage), scored by the weighted chamfer
    distance (:576-599).  Returns (best index, per-hypothesis scores, list of fitted parameter dicts)."""
    smpl = _model_of(smpl_inference)
    dev = smpl.handle.device
    H = len(angles)
    F = markers.shape[0]
    T = lambda t: t.detach().to(dev, torch.float32)
    rep = lambda t: T(t).repeat((H,) + (1,) * (t.dim() - 1)).contiguous()
    mk, opose = rep(markers), rep(o_pose_body)
    pose, tr = rep(pose_body), rep(trans)
    ang = torch.as_tensor(list(angles), dtype=torch.float32, device=dev)
    root = rot_z_apply(ang.view(H, 1, 1, 1), T(root_orient).unsqueeze(0)).reshape(H * F, 1, 3, 3).contiguous()   # :465
    bet = T(betas).reshape(1, -1).repeat(H, 1).contiguous()
    obet = T(o_betas).reshape(1, -1)
    lbfgs = str(config.get("optimizer", {}).get("name", "lbfgs")).lower() == "lbfgs"
    if lbfgs and H > 1:
        # the reference runs an independent L-BFGS per hypothesis (its own line search and stopping test); one stacked
        # problem would couple them through the flat vector, so fit them one after another and merge
        parts = [fit_hypotheses(markers, pose_body, o_pose_body, betas, o_betas, root_orient, trans, img_mask,
                                smpl_inference, config, [a], shard) for a in angles]
        scores = [p[1][0] for p in parts]
        fitted = [p[2][0] for p in parts]
        return int(np.argmin(scores)), scores, fitted
    prob = FitProblem(smpl, mk, opose, obet, shard, groups=H)
    optim_chamfer(mk, pose, opose, bet, obet, root, tr, None, None, smpl, config, problem=prob, groups=H)
    chamfer_snap = {"pose_body": normalize_rot(pose).clone(), "root_orient": normalize_rot(root).clone(),
                    "trans": tr.clone(), "betas": bet.clone()}                                                  # :497-503
    # correspondences per hypothesis (they do not mix frames of different hypotheses)
    vidx = []
    for h in range(H):
        sl = slice(h * F, (h + 1) * F)
        vidx.append(compute_nearest_points(mk[sl], pose[sl], bet[h:h + 1], root[sl], tr[sl], smpl, img_mask=img_mask,
                                           config=config, shard=shard, as_one_hot=False))
    # marker stage of all hypotheses as ONE stacked problem: one set of vertex indices per hypothesis [H,M]
    vall = torch.stack([v.to(dev, torch.int32).reshape(-1) for v in vidx], dim=0).contiguous()
    if H == 1:
        vall = vall[0]
    optim_markers(mk, pose, opose, bet, obet, root, tr, vall, None, smpl, config, shard=shard, groups=H, problem=prob)
    scores, fitted = [], []
    for h in range(H):
        sl = slice(h * F, (h + 1) * F)
        # the reference carries the 6D-normalised rotations on from here (multimodal.py:566-574, 601-606)
        p = {"pose_body": normalize_rot(pose[sl]).clone(), "betas": bet[h:h + 1].clone(),
             "root_orient": normalize_rot(root[sl]).clone(), "trans": tr[sl].clone()}
        out = smpl(betas=p["betas"], body_pose=p["pose_body"], global_orient=p["root_orient"], transl=p["trans"],
                   pose2rot=False)
        wmask = ops.get_marker_mask(mk[sl])
        val, _ = weighted_chamfer_distance(mk[sl], out.vertices, wmask)
        tot = torch.stack([val * wmask.sum(), wmask.sum()]).double().cpu()
        tot = (shard or Shard()).all_reduce_sum(tot)
        scores.append(float(tot[0] / tot[1]) if float(tot[1]) > 0 else 0.0)
        p["vertex_idx"] = vidx[h]
        p["chamfer"] = {k: (v[sl] if k != "betas" else v[h:h + 1]).clone() for k, v in chamfer_snap.items()}
        fitted.append(p)
    best = int(np.argmin(scores))
    return best, scores, fitted


def load_config(filename: str) -> Dict:
    """src/video_mocap/utils/config.py:6-18: YAML whose ``parent:`` names another config file (a path, as in
    config/mht_rotation.yaml:2), loaded recursively and deep-merged with mergedeep.merge({}, parent, child)
    semantics (mergedeep is not installed here, hence the 6-line merge)."""
    import os
    import yaml

    def merge(a, b):
        out = dict(a)
        for k, v in b.items():
            out[k] = merge(out[k], v) if isinstance(v, dict) and isinstance(out.get(k), dict) else v
        return out

    with open(filename, "r") as stream:
        output = yaml.safe_load(stream)
    parent = output.get("parent")
    if parent is not None:
        if not os.path.exists(parent):       # the reference resolves it against the working directory
            alt = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(filename))), parent)
            parent = alt if os.path.exists(alt) else parent
        output = merge(load_config(parent), output)
    return output


def default_config(num_iters: int = 200, optimizer: str = "adam", lr: float = 0.01) -> Dict:
    """The schema and loss weights of config/video_mocap.yaml:1-58 as a dict (stages that are disabled
    there, or are outside the hot path, are left out), with the optimiser switchable between the
    north-star fused Adam and the reference's L-BFGS."""
    return {
        "name": "video_mocap", "parent": None, "stage_repeats": 1, "recompute_marker_labels": False,
        "find_best_part_fits": True, "num_root_orient_angles": 4,
        "optimizer": {"tolerance_grad": 1.0e-7, "tolerance_change": 1.0e-9, "name": optimizer, "lr": lr},
        "stages": {
            "chamfer": {"num_iters": num_iters, "losses": {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0},
                        "single_directional": True, "yaw_lock": True},
            "compute_locations": {"use_velocity": False, "use_barycentric": False, "use_mean": True},
            "marker": {"num_iters": num_iters, "losses": {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0},
                       "use_sdf": False},
        },
    }
