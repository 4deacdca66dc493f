"""Synthetic SMPL-shaped body model and marker-capture workloads.

The licensed ``SMPL_NEUTRAL.pkl`` is not available offline, so every config in
BASELINE.json runs on a synthetic model with the real SMPL dimensions
(6890 vertices, 24 joints, 10 betas, 207 pose-corrective basis rows); see
SURVEY.md section 8(d).  The arrays returned by :func:`make_model` have exactly
the fields ``smplx.SMPL(data_struct=...)`` reads
(/root/reference/third_party/smplx/smplx/body_models.py:144-265): ``v_template``,
``shapedirs``, ``posedirs`` ([V,3,P], reshaped by smplx to [P,3V]),
``J_regressor``, ``kintree_table``, ``weights``, ``f``.

Marker semantics follow the reference's synthetic-marker code: a marker sits
9.5 mm off the surface (src/video_mocap/utils/settings.py:1), a dropped marker is
the exact origin (src/video_mocap/markers/markers_noise.py:39-66, detected by
src/video_mocap/optimization.py:703-715) and markers are unlabeled, i.e. permuted
per frame (src/video_mocap/markers/markers_synthetic.py:65-70).

Everything here is numpy on the host; nothing in this file touches the GPU.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np

NUM_VERTS = 6890
NUM_JOINTS = 24
NUM_BETAS = 10
NUM_POSE_BASIS = 207
MARKER_DISTANCE = 0.0095

# standard SMPL kinematic tree (kintree_table[0] of the pkl; joint names at
# src/video_mocap/utils/smpl_utils.py:12-35)
SMPL_PARENTS = np.array(
    [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21],
    dtype=np.int64,
)

# extra joints = fixed vertex picks, order of
# third_party/smplx/smplx/vertex_joint_selector.py:36-68 with the ids of
# third_party/smplx/smplx/vertex_ids.py:24-46 ('smplh')
EXTRA_JOINT_VERTEX_IDS = np.array(
    [332, 6260, 2800, 4071, 583,            # nose, reye, leye, rear, lear
     3216, 3226, 3387, 6617, 6624, 6787,    # LBigToe LSmallToe LHeel RBigToe RSmallToe RHeel
     2746, 2319, 2445, 2556, 2673,          # l thumb index middle ring pinky
     6191, 5782, 5905, 6016, 6133],         # r thumb index middle ring pinky
    dtype=np.int64,
)

# a 1.7 m T-pose skeleton (x = left, y = up, z = forward), metres
_REST_JOINTS = np.array([
    [0.00, 0.93, 0.00],    # 0 pelvis
    [0.09, 0.84, 0.00],    # 1 l hip
    [-0.09, 0.84, 0.00],   # 2 r hip
    [0.00, 1.05, -0.02],   # 3 spine1
    [0.10, 0.47, 0.00],    # 4 l knee
    [-0.10, 0.47, 0.00],   # 5 r knee
    [0.00, 1.18, -0.01],   # 6 spine2
    [0.10, 0.07, -0.03],   # 7 l ankle
    [-0.10, 0.07, -0.03],  # 8 r ankle
    [0.00, 1.24, 0.00],    # 9 spine3
    [0.11, 0.02, 0.10],    # 10 l foot
    [-0.11, 0.02, 0.10],   # 11 r foot
    [0.00, 1.45, -0.03],   # 12 neck
    [0.08, 1.37, -0.02],   # 13 l collar
    [-0.08, 1.37, -0.02],  # 14 r collar
    [0.00, 1.55, 0.01],    # 15 head
    [0.19, 1.40, -0.03],   # 16 l shoulder
    [-0.19, 1.40, -0.03],  # 17 r shoulder
    [0.45, 1.40, -0.04],   # 18 l elbow
    [-0.45, 1.40, -0.04],  # 19 r elbow
    [0.70, 1.40, -0.04],   # 20 l wrist
    [-0.70, 1.40, -0.04],  # 21 r wrist
    [0.79, 1.40, -0.04],   # 22 l hand
    [-0.79, 1.40, -0.04],  # 23 r hand
], dtype=np.float64)

_BONE_RADIUS = np.array([
    0.13, 0.08, 0.08, 0.13, 0.055, 0.055, 0.13, 0.04, 0.04, 0.14, 0.035, 0.035,
    0.06, 0.07, 0.07, 0.095, 0.05, 0.05, 0.04, 0.04, 0.03, 0.03, 0.03, 0.03,
], dtype=np.float64)


@dataclass
class SyntheticModel:
    """Host-side arrays of a SMPL-shaped model (float32 unless noted)."""
    v_template: np.ndarray      # [V,3]
    shapedirs: np.ndarray       # [V,3,NB]
    posedirs: np.ndarray        # [V,3,P]  (smplx pkl layout; the layer reshapes to [P,3V])
    J_regressor: np.ndarray     # [J,V]
    weights: np.ndarray         # [V,J]
    parents: np.ndarray         # [J] int64, parents[0] = -1
    faces: np.ndarray           # [NF,3] int64
    extra_joint_ids: np.ndarray = field(default_factory=lambda: EXTRA_JOINT_VERTEX_IDS.copy())

    @property
    def kintree_table(self) -> np.ndarray:
        kt = np.zeros((2, self.parents.shape[0]), dtype=np.int64)
        kt[0] = self.parents
        kt[0, 0] = 4294967295  # the pkl stores uint32(-1); smplx overwrites it with -1
        kt[1] = np.arange(self.parents.shape[0])
        return kt

    def posedirs_2d(self) -> np.ndarray:
        """[P, 3V] exactly as smplx registers it (body_models.py:255-259)."""
        p = self.posedirs.shape[-1]
        return np.ascontiguousarray(np.reshape(self.posedirs, [-1, p]).T)

    def checksum(self) -> float:
        acc = 0.0
        for a in (self.v_template, self.shapedirs, self.posedirs, self.J_regressor, self.weights):
            acc += float(np.abs(a.astype(np.float64)).sum())
        return acc


def _point_segment_dist2(p: np.ndarray, a: np.ndarray, b: np.ndarray) -> np.ndarray:
    ab = b - a
    denom = float(ab @ ab)
    if denom < 1e-12:
        return ((p - a) ** 2).sum(-1)
    t = np.clip(((p - a) @ ab) / denom, 0.0, 1.0)
    proj = a + t[:, None] * ab
    return ((p - proj) ** 2).sum(-1)


def make_model(weights: str = "dense", seed: int = 0, num_verts: int = NUM_VERTS) -> SyntheticModel:
    """Build the synthetic model.

    weights: "dense"  -> random normalised 24-way skinning weights (BASELINE.json wording),
             "top4"   -> SMPL-like: the 4 nearest bones, renormalised (<= 4 nnz per vertex).
    """
    rng = np.random.default_rng(seed)
    V, J = num_verts, NUM_JOINTS
    parents = SMPL_PARENTS.copy()
    joints = _REST_JOINTS

    # bones: joint j owns the segment from itself to the mean of its children (or a short stub)
    ends = np.zeros_like(joints)
    for j in range(J):
        ch = np.where(parents == j)[0]
        if len(ch) > 0:
            ends[j] = joints[ch].mean(0)
        else:
            d = joints[j] - joints[parents[j]]
            ends[j] = joints[j] + 0.8 * d
    length = np.linalg.norm(ends - joints, axis=1) + 2 * _BONE_RADIUS
    share = length * _BONE_RADIUS
    share = share / share.sum()
    owner = rng.choice(J, size=V, p=share)
    # every bone gets at least a handful of vertices
    owner[: J * 8] = np.repeat(np.arange(J), 8)

    # points on capsule surfaces
    t = rng.uniform(-0.15, 1.15, size=V)
    axis_pt = joints[owner] + np.clip(t, 0, 1)[:, None] * (ends[owner] - joints[owner])
    dirs = rng.normal(size=(V, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    rad = _BONE_RADIUS[owner] * rng.uniform(0.9, 1.1, size=V)
    v_template = axis_pt + dirs * rad[:, None]

    # distance of every vertex to every bone
    d2 = np.stack([_point_segment_dist2(v_template, joints[j], ends[j]) for j in range(J)], axis=1)
    sigma = 0.06
    logits = -d2 / (sigma * sigma)
    logits -= logits.max(1, keepdims=True)
    w_geo = np.exp(logits)
    w_geo /= w_geo.sum(1, keepdims=True)
    if weights == "top4":
        order = np.argsort(-w_geo, axis=1)
        w = np.zeros_like(w_geo)
        rows = np.arange(V)[:, None]
        w[rows, order[:, :4]] = w_geo[rows, order[:, :4]]
        w /= w.sum(1, keepdims=True)
    elif weights == "dense":
        # "random normalised skinning weights": geometric prior blended with a random
        # positive 24-vector so every joint has non-zero influence on every vertex
        r = rng.uniform(0.0, 1.0, size=(V, J)) ** 4
        w = 0.7 * w_geo + 0.3 * r / r.sum(1, keepdims=True)
        w /= w.sum(1, keepdims=True)
    else:
        raise ValueError("weights must be 'dense' or 'top4'")

    # joint regressor: each joint = softmax-weighted mean of its 32 nearest vertices
    J_reg = np.zeros((J, V))
    for j in range(J):
        dj = ((v_template - joints[j]) ** 2).sum(1)
        nn = np.argsort(dj)[:32]
        wj = np.exp(-dj[nn] / (0.05 ** 2))
        J_reg[j, nn] = wj / wj.sum()

    shapedirs = rng.normal(0.0, 0.01, size=(V, 3, NUM_BETAS))
    posedirs = rng.normal(0.0, 0.001, size=(V, 3, NUM_POSE_BASIS))

    # faces are only needed for smplx's faces_tensor buffer; a fan over consecutive vertices
    nf = 2 * V - 4
    base = np.arange(nf)
    faces = np.stack([base % V, (base + 1) % V, (base + 2) % V], axis=1).astype(np.int64)

    return SyntheticModel(
        v_template=v_template.astype(np.float32),
        shapedirs=shapedirs.astype(np.float32),
        posedirs=posedirs.astype(np.float32),
        J_regressor=J_reg.astype(np.float32),
        weights=w.astype(np.float32),
        parents=parents,
        faces=faces,
    )


# ----------------------------------------------------------------------------------------------
# rotations (host-side numpy helpers for workload generation only)
# ----------------------------------------------------------------------------------------------
def axis_angle_to_matrix_np(aa: np.ndarray) -> np.ndarray:
    """Rodrigues, float64, [...,3] -> [...,3,3]."""
    aa = np.asarray(aa, dtype=np.float64)
    theta = np.linalg.norm(aa, axis=-1, keepdims=True)
    k = aa / np.maximum(theta, 1e-12)
    K = np.zeros(aa.shape[:-1] + (3, 3))
    K[..., 0, 1] = -k[..., 2]
    K[..., 0, 2] = k[..., 1]
    K[..., 1, 0] = k[..., 2]
    K[..., 1, 2] = -k[..., 0]
    K[..., 2, 0] = -k[..., 1]
    K[..., 2, 1] = k[..., 0]
    s = np.sin(theta)[..., None]
    c = np.cos(theta)[..., None]
    return np.eye(3) + s * K + (1 - c) * (K @ K)


def _smooth_series(rng: np.random.Generator, n: int, dims: tuple, scale: float, n_modes: int = 4) -> np.ndarray:
    """Band-limited random series [n, *dims] with amplitude <= ~scale."""
    t = np.linspace(0.0, 1.0, n)[(...,) + (None,) * len(dims)]
    out = np.zeros((n,) + dims)
    for m in range(1, n_modes + 1):
        amp = rng.normal(size=dims) / m
        ph = rng.uniform(0, 2 * np.pi, size=dims)
        out += amp * np.sin(2 * np.pi * m * t * (1 + n / 600.0) + ph)
    out *= scale / max(1e-9, np.abs(out).max())
    return out


@dataclass
class Capture:
    """One synthetic capture: ground truth, unlabeled markers and a video-pose prior."""
    markers: np.ndarray          # [F,M,3] float32, dropped = exact zeros, permuted per frame
    gt_pose_aa: np.ndarray       # [F,24,3] float32 axis-angle (root first)
    gt_betas: np.ndarray         # [1,10]
    gt_trans: np.ndarray         # [F,3]
    prior_pose_body: np.ndarray  # [F,23,3,3] float32 rotation matrices (o_pose_body)
    prior_root_orient: np.ndarray  # [F,1,3,3]
    prior_trans: np.ndarray      # [F,3]
    prior_betas: np.ndarray      # [1,10] (o_betas)
    img_mask: np.ndarray         # [F] float32 (1 = video prior available)
    marker_vertex_ids: np.ndarray  # [M_real] int64 ground-truth attachment vertices
    n_ghost: int = 0


def lbs_numpy(model: SyntheticModel, betas: np.ndarray, rot: np.ndarray, trans: np.ndarray,
              vertex_ids: Optional[np.ndarray] = None) -> np.ndarray:
    """float64 host LBS used only to place synthetic markers (not an oracle, not a product path)."""
    F = rot.shape[0]
    vt = model.v_template.astype(np.float64)
    sd = model.shapedirs.astype(np.float64)
    pd = model.posedirs.astype(np.float64)
    v_shaped = vt + sd @ betas.reshape(-1).astype(np.float64)
    J = model.J_regressor.astype(np.float64) @ v_shaped
    ids = np.arange(vt.shape[0]) if vertex_ids is None else np.asarray(vertex_ids)
    pf = (rot[:, 1:] - np.eye(3)).reshape(F, -1)
    v_posed = v_shaped[ids][None] + np.einsum("fp,vkp->fvk", pf, pd[ids])
    G = np.zeros((F, NUM_JOINTS, 4, 4))
    for j in range(NUM_JOINTS):
        L = np.zeros((F, 4, 4))
        L[:, :3, :3] = rot[:, j]
        L[:, 3, 3] = 1.0
        p = model.parents[j]
        L[:, :3, 3] = J[j] - (J[p] if p >= 0 else 0.0)
        G[:, j] = L if p < 0 else G[:, p] @ L
    A = G.copy()
    A[:, :, :3, 3] -= np.einsum("fjab,jb->fja", G[:, :, :3, :3], J)
    W = model.weights.astype(np.float64)[ids]
    T = np.einsum("vj,fjab->fvab", W, A)
    out = np.einsum("fvab,fvb->fva", T[..., :3, :3], v_posed) + T[..., :3, 3]
    return out + trans[:, None, :]


def make_capture(model: SyntheticModel, num_frames: int, num_markers: int = 41, seed: int = 1,
                 noise_std: float = 0.001, n_ghost: int = 0, dropout_p: float = 0.0,
                 dropout_block: int = 10, permute: bool = True, prior_pose_noise: float = 0.1,
                 prior_trans_err: float = 0.2, prior_yaw_err: float = 0.3,
                 max_angle: float = 0.6, chunk: int = 2048) -> Capture:
    """Config 1/3/4/5 workload generator."""
    rng = np.random.default_rng(seed)
    F = num_frames
    aa = _smooth_series(rng, F, (NUM_JOINTS, 3), max_angle)
    aa[:, 0] *= 0.5
    trans = np.cumsum(rng.normal(0, 0.004, size=(F, 3)), axis=0)
    trans += _smooth_series(rng, F, (3,), 0.5)
    betas = rng.normal(0, 1, size=(1, NUM_BETAS))
    rot = axis_angle_to_matrix_np(aa)

    # marker attachment: fixed vertices spread over the body
    stride = model.v_template.shape[0] // num_markers
    vids = (np.arange(num_markers) * stride + rng.integers(0, stride, size=num_markers)).astype(np.int64)

    M_total = num_markers + n_ghost
    markers = np.zeros((F, M_total, 3), dtype=np.float64)
    centroid = model.v_template.astype(np.float64).mean(0)
    for s in range(0, F, chunk):
        e = min(F, s + chunk)
        pts = lbs_numpy(model, betas, rot[s:e], trans[s:e], vertex_ids=vids)
        # 9.5 mm along an outward direction (posed direction from the posed body centroid)
        ctr = lbs_numpy(model, betas, rot[s:e], trans[s:e], vertex_ids=np.arange(0, model.v_template.shape[0], 64)).mean(1)
        d = pts - ctr[:, None, :]
        d /= np.maximum(np.linalg.norm(d, axis=-1, keepdims=True), 1e-9)
        markers[s:e, :num_markers] = pts + MARKER_DISTANCE * d
    del centroid
    markers[:, :num_markers] += rng.normal(0, noise_std, size=(F, num_markers, 3))

    if n_ghost > 0:
        lo = markers[:, :num_markers].min(1) - 0.3
        hi = markers[:, :num_markers].max(1) + 0.3
        for s in range(0, F, 50):
            e = min(F, s + 50)
            u = rng.uniform(size=(1, n_ghost, 3))
            markers[s:e, num_markers:] = lo[s:e, None] + u * (hi[s:e, None] - lo[s:e, None])

    if dropout_p > 0:
        nb = (F + dropout_block - 1) // dropout_block
        drop = rng.uniform(size=(nb, M_total)) < dropout_p
        drop = np.repeat(drop, dropout_block, axis=0)[:F]
        markers[drop] = 0.0

    if permute:
        perm = np.argsort(rng.uniform(size=(F, M_total)), axis=1)
        markers = np.take_along_axis(markers, perm[..., None], axis=1)

    # synthetic video-pose prior (HMR-like): noisy pose, yaw error, translation error
    prior_aa = aa + rng.normal(0, prior_pose_noise, size=aa.shape)
    prior_rot = axis_angle_to_matrix_np(prior_aa)
    yaw = prior_yaw_err * rng.normal()
    Ry = axis_angle_to_matrix_np(np.array([0.0, yaw, 0.0]))
    prior_root = Ry @ prior_rot[:, :1]
    prior_trans = trans + prior_trans_err * rng.normal(size=(1, 3)) + rng.normal(0, 0.01, size=(F, 3))
    prior_betas = betas + rng.normal(0, 0.5, size=betas.shape)

    f32 = np.float32
    return Capture(
        markers=markers.astype(f32), gt_pose_aa=aa.astype(f32), gt_betas=betas.astype(f32),
        gt_trans=trans.astype(f32), prior_pose_body=prior_rot[:, 1:].astype(f32),
        prior_root_orient=prior_root.astype(f32), prior_trans=prior_trans.astype(f32),
        prior_betas=prior_betas.astype(f32), img_mask=np.ones(F, dtype=f32),
        marker_vertex_ids=vids, n_ghost=n_ghost,
    )


def make_lbs_microbench(num_frames: int = 4096, seed: int = 2) -> Dict[str, np.ndarray]:
    """Config 2 inputs: random rotations within 0.8 rad, per-frame betas, transl, upstream grads."""
    rng = np.random.default_rng(seed)
    F = num_frames
    axis = rng.normal(size=(F, NUM_JOINTS, 3))
    axis /= np.linalg.norm(axis, axis=-1, keepdims=True)
    ang = rng.uniform(0, 0.8, size=(F, NUM_JOINTS, 1))
    aa = axis * ang
    rot = axis_angle_to_matrix_np(aa)
    f32 = np.float32
    return {
        "pose_aa": aa.astype(f32),
        "rot": rot.astype(f32),
        "betas": rng.normal(0, 1, size=(F, NUM_BETAS)).astype(f32),
        "transl": rng.normal(0, 1, size=(F, 3)).astype(f32),
        "g_vertices": rng.normal(0, 1, size=(F, NUM_VERTS, 3)).astype(f32),
        "g_joints": rng.normal(0, 1, size=(F, NUM_JOINTS + len(EXTRA_JOINT_VERTEX_IDS), 3)).astype(f32),
    }


def make_phalp_like(num_frames: int = 40, seed: int = 17, gaps=((0, 3), (11, 16), (25, 26), (37, 40))) -> Dict:
    """A PHALP-tracker-shaped per-frame dictionary (the .pkl the reference reads through ImgSmpl,
    src/video_mocap/img_smpl/img_smpl.py:13-137) with synthetic content: frames inside ``gaps`` carry no track."""
    rng = np.random.default_rng(seed)
    aa = _smooth_series(rng, num_frames, (NUM_JOINTS, 3), 0.6)
    rot = axis_angle_to_matrix_np(aa).astype(np.float32)
    out = {}
    for f in range(num_frames):
        missing = any(a <= f < b for a, b in gaps)
        j3 = rng.normal(0, 0.5, size=(45, 3)).astype(np.float32) + np.array([0.0, 0.0, 3.0], dtype=np.float32)
        j2 = (rng.uniform(0.2, 0.8, size=(45, 2)) + 0.001 * f).astype(np.float32)
        rec = {"tracked_ids": [] if missing else [1], "smpl": [], "3d_joints": [], "2d_joints": [], "camera_bbox": [],
               "center": [], "scale": [], "size": []}
        if not missing:
            rec["smpl"] = [{"global_orient": rot[f, :1], "body_pose": rot[f, 1:], "betas": rng.normal(0, 1, 10).astype(np.float32)}]
            rec["3d_joints"] = [j3]
            rec["2d_joints"] = [j2.reshape(-1)]
            rec["camera_bbox"] = [rng.normal(0, 1, 3).astype(np.float32)]
            rec["center"] = [rng.uniform(100, 500, 2).astype(np.float32)]
            rec["scale"] = [np.float32(rng.uniform(100, 300))]
            rec["size"] = [np.array([720.0, 1280.0], dtype=np.float32)]
        out["frame_%05d.jpg" % f] = rec
    return out
