"""Host-side mirror of the reference's SMPL layer, backed by libuuomocap (sm_100a CUDA).

Drop-in surface (same names, argument meaning and error behaviour as the reference):
  * ``SMPL.forward(betas, body_pose, global_orient, transl, return_verts, return_full_pose, pose2rot)``
    -> object with ``.vertices [B,6890,3]``, ``.joints [B,45,3]``  -- smplx.SMPL.forward,
    /root/reference/third_party/smplx/smplx/body_models.py:315-399
  * ``SmplInference(device).forward(poses, betas, root_orient, trans) -> {"joints", "vertices"}``,
    ``.get_lbs_weights()``, ``.smpl.{faces,parents,lbs_weights}``
    -- /root/reference/src/video_mocap/utils/smpl.py:9-52
Outputs take part in autograd: the layer is a ``torch.autograd.Function`` whose forward and backward
are single C-ABI calls.  PyTorch is only the owner of device memory and streams here.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr, stream_ptr

POSE_ROTMAT, POSE_AXISANGLE, POSE_RAW6D = 0, 1, 2


@dataclass
class SMPLOutput:
    """Same fields as smplx.utils.SMPLOutput (utils.py:57-60)."""
    vertices: Optional[torch.Tensor] = None
    joints: Optional[torch.Tensor] = None
    full_pose: Optional[torch.Tensor] = None
    global_orient: Optional[torch.Tensor] = None
    transl: Optional[torch.Tensor] = None
    betas: Optional[torch.Tensor] = None
    body_pose: Optional[torch.Tensor] = None
    v_shaped: Optional[torch.Tensor] = None


class ModelHandle:
    """Owns one ``uuo_model`` (device constants).  Immutable after creation."""

    def __init__(self, v_template, shapedirs, posedirs_2d, J_regressor, weights, parents, extra_joint_ids,
                 device: torch.device):
        device = torch.device(device)
        if device.type != "cuda":
            raise _lib.UuoError("uuo-mocap_b200 runs on CUDA devices only (no CPU fallback); got %s" % device)
        f32 = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.float32))
        i32 = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.int32))
        vt, sd, pd, jr, w = f32(v_template), f32(shapedirs), f32(posedirs_2d), f32(J_regressor), f32(weights)
        par = i32(parents).copy()
        par[0] = -1
        ex = i32(extra_joint_ids)
        self.V, self.J, self.NB = vt.shape[0], jr.shape[0], sd.shape[-1]
        self.P, self.NE = pd.shape[0], ex.shape[0]
        if pd.shape[1] != 3 * self.V:
            raise ValueError("posedirs must be [P, 3V] (the layout smplx registers)")
        self.device = device
        self.index = device.index if device.index is not None else torch.cuda.current_device()
        h = C.c_void_p()
        as_p = lambda a: a.ctypes.data_as(C.c_void_p)
        check(_lib.lib().uuo_model_create(as_p(vt), as_p(sd), as_p(pd), as_p(jr), as_p(w), as_p(par), as_p(ex),
                                          self.V, self.J, self.NB, self.P, self.NE, self.index, C.byref(h)))
        self.handle = h
        self.max_nnz = _lib.lib().uuo_model_info(h, 6)
        # device pointer (int32[V]) of the spatially coherent vertex order used by the pruned correspondence kernel
        self.vertex_order = C.c_void_p(_lib.lib().uuo_model_vertex_order(h))

    def __del__(self):
        h = getattr(self, "handle", None)
        try:
            if h is not None and _lib is not None and _lib._lib is not None:
                _lib._lib.uuo_model_destroy(h)
                self.handle = None
        except Exception:   # interpreter shutdown
            pass

    def workspace(self, F: int) -> torch.Tensor:
        n = _lib.lib().uuo_lbs_workspace_bytes(self.handle, F)
        return torch.empty((n + 3) // 4, dtype=torch.float32, device=self.device)


def _c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class _LBSFunction(torch.autograd.Function):
    """vertices, joints = lbs(betas, global_orient, body_pose, transl[, yaw]); one C call each way."""

    @staticmethod
    def forward(ctx, mh: ModelHandle, pose_mode: int, betas, global_orient, body_pose, transl, yaw):
        betas, global_orient, body_pose = _c(betas.detach()), _c(global_orient.detach()), _c(body_pose.detach())
        transl = _c(transl.detach()) if transl is not None else None
        yaw = _c(yaw.detach()) if yaw is not None else None
        F = global_orient.shape[0]
        dev = mh.device
        vertices = torch.empty((F, mh.V, 3), dtype=torch.float32, device=dev)
        joints = torch.empty((F, mh.J + mh.NE, 3), dtype=torch.float32, device=dev)
        ws = mh.workspace(F)
        with torch.cuda.device(dev):
            check(_lib.lib().uuo_lbs_forward(mh.handle, ptr(betas), betas.shape[0], ptr(global_orient), ptr(body_pose),
                                             pose_mode, ptr(yaw), ptr(transl), F, ptr(vertices), ptr(joints),
                                             ptr(ws), stream_ptr(dev)))
        ctx.mh, ctx.pose_mode, ctx.F = mh, pose_mode, F
        ctx.has_transl, ctx.has_yaw = transl is not None, yaw is not None
        ctx.save_for_backward(betas, global_orient, body_pose, yaw, ws)
        ctx.set_materialize_grads(False)
        return vertices, joints

    @staticmethod
    def backward(ctx, g_vertices, g_joints):
        betas, global_orient, body_pose, yaw, ws = ctx.saved_tensors
        mh, F, dev = ctx.mh, ctx.F, ctx.mh.device
        need = ctx.needs_input_grad   # (mh, pose_mode, betas, go, bp, transl, yaw)
        g_vertices, g_joints = _c(g_vertices), _c(g_joints)
        mk = lambda like, on: torch.empty_like(like) if on else None
        g_betas = mk(betas, need[2])
        g_go = mk(global_orient, need[3])
        g_bp = mk(body_pose, need[4])
        g_tr = torch.empty((F, 3), dtype=torch.float32, device=dev) if (need[5] and ctx.has_transl) else None
        g_yaw = torch.empty((F,), dtype=torch.float32, device=dev) if (need[6] and ctx.has_yaw) else None
        with torch.cuda.device(dev):
            check(_lib.lib().uuo_lbs_backward(mh.handle, ptr(betas), betas.shape[0], ptr(global_orient), ptr(body_pose),
                                              ctx.pose_mode, ptr(yaw), F, ptr(g_vertices), ptr(g_joints), ptr(ws),
                                              ptr(g_betas), ptr(g_go), ptr(g_bp), ptr(g_yaw), ptr(g_tr),
                                              stream_ptr(dev)))
        return None, None, g_betas, g_go, g_bp, g_tr, g_yaw


def lbs_apply(mh: ModelHandle, betas, global_orient, body_pose, transl=None, yaw=None, pose_mode=POSE_ROTMAT):
    """Functional form used by SMPL.forward and the tests.  Shapes: betas [1|G|F,NB] (G | F: one row per group of F/G
    consecutive frames, the stacked hypotheses of the batched MHT loop);
    rot modes: global_orient [F,1,3,3], body_pose [F,J-1,3,3]; axis-angle: [F,1,3] / [F,J-1,3]."""
    F = global_orient.shape[0]
    J = mh.J
    if pose_mode == POSE_AXISANGLE:
        go, bp = global_orient.reshape(F, 1, 3), body_pose.reshape(F, J - 1, 3)
    else:
        go, bp = global_orient.reshape(F, 1, 3, 3), body_pose.reshape(F, J - 1, 3, 3)
    if betas.shape[0] < 1 or betas.shape[0] > F or F % betas.shape[0]:
        raise ValueError("betas batch (%d) must be 1, the pose batch (%d) or a divisor of it" % (betas.shape[0], F))
    if betas.shape[1] != mh.NB:
        raise ValueError("betas must have %d columns" % mh.NB)
    if yaw is not None:
        yaw = yaw.reshape(F)
    return _LBSFunction.apply(mh, pose_mode, betas, go, bp, transl, yaw)


class SMPL(torch.nn.Module):
    """smplx.SMPL look-alike built from a ``data_struct`` (body_models.py:57,132-143) or from a
    ``synthetic.SyntheticModel``.  Only the fields the fitting path reads are kept."""

    NUM_JOINTS = 23
    NUM_BODY_JOINTS = 23
    SHAPE_SPACE_DIM = 300

    def __init__(self, data_struct, device="cuda", num_betas: int = 10, extra_joint_ids=None, batch_size: int = 1):
        super().__init__()
        from .synthetic import EXTRA_JOINT_VERTEX_IDS
        device = torch.device(device)
        g = lambda k: np.asarray(getattr(data_struct, k))
        v_template = g("v_template").astype(np.float32)
        shapedirs = g("shapedirs")[:, :, :num_betas].astype(np.float32)          # body_models.py:158-172
        posedirs = g("posedirs").astype(np.float32)
        if posedirs.ndim == 3:                                                     # pkl layout [V,3,P] -> [P,3V] (:255-259)
            posedirs = np.ascontiguousarray(posedirs.reshape(-1, posedirs.shape[-1]).T)
        J_regressor = g("J_regressor").astype(np.float32)
        weights = g("weights").astype(np.float32)
        if hasattr(data_struct, "kintree_table"):
            parents = np.asarray(data_struct.kintree_table)[0].astype(np.int64).copy()
        else:
            parents = g("parents").astype(np.int64).copy()
        parents[0] = -1                                                            # :264-265
        faces = g("f") if hasattr(data_struct, "f") else g("faces")
        if extra_joint_ids is None:
            extra_joint_ids = getattr(data_struct, "extra_joint_ids", EXTRA_JOINT_VERTEX_IDS)
        self.num_betas = num_betas
        self.batch_size = batch_size
        self.faces = np.asarray(faces)
        self.extra_joint_ids = np.asarray(extra_joint_ids, dtype=np.int64)
        self._twin = None
        self.visit_ids = None          # set on a twin: int32 [V], original vertex id of every row of its output
        self.handle = ModelHandle(v_template, shapedirs, posedirs, J_regressor, weights, parents, extra_joint_ids, device)
        T = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=device)
        self.register_buffer("faces_tensor", T(self.faces.astype(np.int64), torch.long))
        self.register_buffer("v_template", T(v_template))
        self.register_buffer("shapedirs", T(shapedirs))
        self.register_buffer("posedirs", T(posedirs))
        self.register_buffer("J_regressor", T(J_regressor))
        self.register_buffer("lbs_weights", T(weights))
        self.register_buffer("parents", T(parents, torch.long))

    @property
    def device(self):
        return self.handle.device

    def vertex_order(self) -> np.ndarray:
        """The model's spatially coherent vertex permutation (uuo_model_vertex_order), as a host array."""
        out = np.empty(self.handle.V, dtype=np.int32)
        check(_lib.lib().uuo_model_vertex_order_host(self.handle.handle, out.ctypes.data_as(C.c_void_p)))
        return out

    def visiting_twin(self) -> "SMPL":
        """The same body with its vertices RENUMBERED in visiting order (row p of the twin's vertices is vertex
        ``visit_ids[p]`` of this model).  The fitting loop runs the dense forward on the twin so that the
        correspondence kernel streams tiles of 32 consecutive, spatially close vertices (uuo_nn_argmin_presorted);
        joints and every per-frame quantity are unchanged.  Internal: never handed to the caller."""
        if self._twin is None:
            perm = self.vertex_order().astype(np.int64)
            inv = np.empty_like(perm)
            inv[perm] = np.arange(perm.shape[0])
            V = perm.shape[0]
            cpu = lambda t: t.detach().cpu().numpy()
            pd = cpu(self.posedirs)                                                    # [P, 3V]
            st = _Struct(v_template=cpu(self.v_template)[perm], shapedirs=cpu(self.shapedirs)[perm],
                         posedirs=np.ascontiguousarray(pd.reshape(pd.shape[0], V, 3)[:, perm].reshape(pd.shape[0], 3 * V)),
                         J_regressor=np.ascontiguousarray(cpu(self.J_regressor)[:, perm]),
                         weights=cpu(self.lbs_weights)[perm], parents=cpu(self.parents), f=self.faces)
            tw = SMPL(st, device=self.device, num_betas=self.num_betas, extra_joint_ids=inv[self.extra_joint_ids],
                      batch_size=self.batch_size)
            tw.visit_ids = torch.as_tensor(perm.astype(np.int32), device=self.device)
            self._twin = tw
        return self._twin

    def get_num_verts(self):
        return self.handle.V

    def forward(self, betas: Optional[torch.Tensor] = None, body_pose: Optional[torch.Tensor] = None,
                global_orient: Optional[torch.Tensor] = None, transl: Optional[torch.Tensor] = None,
                return_verts: bool = True, return_full_pose: bool = False, pose2rot: bool = True, **kwargs):
        dev = self.device
        J = self.handle.J
        n = max([t.shape[0] for t in (betas, body_pose, global_orient) if t is not None] + [self.batch_size])
        if global_orient is None:
            global_orient = (torch.zeros(n, 3, device=dev) if pose2rot else
                             torch.eye(3, device=dev).expand(n, 1, 3, 3))
        if body_pose is None:
            body_pose = (torch.zeros(n, (J - 1) * 3, device=dev) if pose2rot else
                         torch.eye(3, device=dev).expand(n, J - 1, 3, 3))
        if betas is None:
            betas = torch.zeros(1, self.num_betas, device=dev)
        batch = max(betas.shape[0], global_orient.shape[0], body_pose.shape[0])      # :371-376
        if global_orient.shape[0] != batch or body_pose.shape[0] != batch:
            raise ValueError("global_orient and body_pose must share the batch size")
        if betas.shape[0] != batch and betas.shape[0] != 1:
            betas = betas.expand(int(batch / betas.shape[0]), -1)
        mode = POSE_AXISANGLE if pose2rot else POSE_ROTMAT
        vertices, joints = lbs_apply(self.handle, betas, global_orient, body_pose, transl, None, mode)
        full_pose = None
        if return_full_pose:
            full_pose = torch.cat([global_orient.reshape(batch, 1, -1), body_pose.reshape(batch, J - 1, -1)], dim=1)
        return SMPLOutput(vertices=vertices if return_verts else None, joints=joints, betas=betas,
                          global_orient=global_orient, body_pose=body_pose, full_pose=full_pose, transl=transl)


class SmplInference:
    """src/video_mocap/utils/smpl.py:9-52 with the smplx layer replaced by the CUDA one."""

    def __init__(self, device, model=None, weights: str = "top4"):
        from . import synthetic
        self.device = torch.device(device)
        if model is None:   # the licensed SMPL_NEUTRAL.pkl is not shipped: synthetic SMPL-shaped model
            model = synthetic.make_model(weights)
        self.smpl = model if isinstance(model, SMPL) else SMPL(_as_struct(model), device=self.device)

    def __call__(self, *a, **k):
        return self.forward(*a, **k)

    def forward(self, poses, betas, root_orient, trans):
        if betas.shape[1] != 10:
            raise ValueError("Betas array must have 10 beta values")      # utils/smpl.py:35-36
        out = self.smpl(body_pose=poses, betas=betas, global_orient=root_orient, transl=trans, pose2rot=False)
        return {"joints": out.joints, "vertices": out.vertices}

    def get_lbs_weights(self):
        return self.smpl.lbs_weights


class _Struct:
    def __init__(self, **kw):
        self.__dict__.update(kw)


def _as_struct(model):
    """SyntheticModel -> the data_struct smplx.SMPL reads (body_models.py:144-265)."""
    if hasattr(model, "kintree_table") and hasattr(model, "weights") and hasattr(model, "f"):
        return model
    return _Struct(v_template=model.v_template, shapedirs=model.shapedirs, posedirs=model.posedirs,
                   J_regressor=model.J_regressor, kintree_table=model.kintree_table, weights=model.weights,
                   f=model.faces, extra_joint_ids=model.extra_joint_ids)
