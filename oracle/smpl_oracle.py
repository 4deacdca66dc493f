"""Oracle restatement of the smplx SMPL layer (TEST INFRASTRUCTURE, CPU, fp32).

PINNED against the reference's own code: ``tests/test_oracle_smplx.py`` runs the vendored
``/root/reference/third_party/smplx`` on the same inputs (when that tree is present) and
requires bit-identical vertices/joints/gradients, and ``tests/golden/smplx_*.npz`` holds
outputs produced by that real reference (``oracle/gen_golden.py``).

Each function names the reference lines it restates (paths relative to
``/root/reference/third_party/smplx/smplx``).
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import numpy as np
import torch


class OracleOutput(NamedTuple):
    vertices: torch.Tensor   # [B,V,3]
    joints: torch.Tensor     # [B,J+E,3]
    A: torch.Tensor          # [B,J,4,4] relative transforms (for FK-kernel parity)
    posed_joints: torch.Tensor  # [B,J,3] before transl / extra joints
    v_posed: torch.Tensor    # [B,V,3]


def shape_offsets(betas: torch.Tensor, shapedirs: torch.Tensor) -> torch.Tensor:
    """lbs.py:275-296 -- displacement[b,v,k] = sum_l betas[b,l] * shapedirs[v,k,l]."""
    return torch.einsum("bl,mkl->bmk", [betas, shapedirs])


def regress_joints(J_regressor: torch.Tensor, verts: torch.Tensor) -> torch.Tensor:
    """lbs.py:255-272 -- J[b,j,k] = sum_v J_regressor[j,v] * verts[b,v,k]."""
    return torch.einsum("bik,ji->bjk", [verts, J_regressor])


def rodrigues(rot_vecs: torch.Tensor) -> torch.Tensor:
    """lbs.py:299-333 -- axis-angle [N,3] -> [N,3,3]; note the +1e-8 INSIDE the norm."""
    n = rot_vecs.shape[0]
    dtype = rot_vecs.dtype
    angle = torch.norm(rot_vecs + 1e-8, dim=1, keepdim=True)
    axis = rot_vecs / angle
    c = torch.cos(angle).unsqueeze(1)
    s = torch.sin(angle).unsqueeze(1)
    ax, ay, az = torch.split(axis, 1, dim=1)
    z = torch.zeros((n, 1), dtype=dtype, device=rot_vecs.device)
    K = torch.cat([z, -az, ay, az, z, -ax, -ay, ax, z], dim=1).view(n, 3, 3)
    eye = torch.eye(3, dtype=dtype, device=rot_vecs.device).unsqueeze(0)
    return eye + s * K + (1 - c) * torch.bmm(K, K)


def rigid_chain(rot_mats: torch.Tensor, joints: torch.Tensor, parents: torch.Tensor):
    """lbs.py:336-405 -- local [R t;0 1], chain along the tree, remove the rest-pose joint."""
    B, J = joints.shape[:2]
    jt = joints.unsqueeze(-1)                      # [B,J,3,1]
    rel = jt.clone()
    rel[:, 1:] -= jt[:, parents[1:]]
    R = rot_mats.reshape(-1, 3, 3)
    t = rel.reshape(-1, 3, 1)
    local = torch.cat([torch.nn.functional.pad(R, [0, 0, 0, 1]),
                       torch.nn.functional.pad(t, [0, 0, 0, 1], value=1)], dim=2).reshape(-1, J, 4, 4)
    chain = [local[:, 0]]
    for i in range(1, J):
        chain.append(torch.matmul(chain[int(parents[i])], local[:, i]))
    G = torch.stack(chain, dim=1)
    posed = G[:, :, :3, 3]
    j_h = torch.nn.functional.pad(jt, [0, 0, 0, 1])
    A = G - torch.nn.functional.pad(torch.matmul(G, j_h), [3, 0, 0, 0, 0, 0, 0, 0])
    return posed, A


class OracleSMPL:
    """SMPL.forward (body_models.py:315-399) + lbs (lbs.py:156-252) + extra joints
    (vertex_joint_selector.py:73-80), as plain functions of tensors."""

    def __init__(self, v_template, shapedirs, posedirs_2d, J_regressor, weights, parents, extra_joint_ids):
        t = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float32)
        self.v_template = t(v_template)          # [V,3]
        self.shapedirs = t(shapedirs)            # [V,3,NB]
        self.posedirs = t(posedirs_2d)           # [P,3V]
        self.J_regressor = t(J_regressor)        # [J,V]
        self.lbs_weights = t(weights)            # [V,J]
        self.parents = torch.as_tensor(np.asarray(parents), dtype=torch.long).clone()
        self.parents[0] = -1
        self.extra_joint_ids = torch.as_tensor(np.asarray(extra_joint_ids), dtype=torch.long)

    def to(self, device) -> "OracleSMPL":
        """Same restatement with its constants on ``device`` (bench.py's stock-PyTorch-on-GPU comparison bar;
        ``parents`` stays on the host, it only drives the Python loop of the chain)."""
        for name in ("v_template", "shapedirs", "posedirs", "J_regressor", "lbs_weights", "extra_joint_ids"):
            setattr(self, name, getattr(self, name).to(device))
        return self

    @classmethod
    def from_model(cls, m) -> "OracleSMPL":
        return cls(m.v_template, m.shapedirs, m.posedirs_2d(), m.J_regressor, m.weights, m.parents,
                   m.extra_joint_ids)

    def forward(self, betas: torch.Tensor, body_pose: torch.Tensor, global_orient: torch.Tensor,
                transl: Optional[torch.Tensor] = None, pose2rot: bool = True) -> OracleOutput:
        full_pose = torch.cat([global_orient, body_pose], dim=1)                 # body_models.py:369
        B = max(betas.shape[0], global_orient.shape[0], body_pose.shape[0])      # :371-376
        if betas.shape[0] != B:
            betas = betas.expand(int(B / betas.shape[0]), -1)
        dtype = betas.dtype
        v_shaped = self.v_template + shape_offsets(betas, self.shapedirs)        # lbs.py:209
        J = regress_joints(self.J_regressor, v_shaped)                           # lbs.py:213
        eye = torch.eye(3, dtype=dtype, device=betas.device)
        if pose2rot:                                                             # lbs.py:218-225
            R = rodrigues(full_pose.reshape(-1, 3)).view(B, -1, 3, 3)
            feat = (R[:, 1:, :, :] - eye).view(B, -1)
            offs = torch.matmul(feat, self.posedirs).view(B, -1, 3)
        else:                                                                    # lbs.py:226-231
            feat = full_pose[:, 1:].view(B, -1, 3, 3) - eye
            R = full_pose.view(B, -1, 3, 3)
            offs = torch.matmul(feat.view(B, -1), self.posedirs).view(B, -1, 3)
        v_posed = offs + v_shaped                                                # lbs.py:233
        posed_joints, A = rigid_chain(R, J, self.parents)                        # lbs.py:235
        nj = self.J_regressor.shape[0]
        W = self.lbs_weights.unsqueeze(0).expand(B, -1, -1)                      # lbs.py:239
        T = torch.matmul(W, A.view(B, nj, 16)).view(B, -1, 4, 4)                 # lbs.py:242
        ones = torch.ones([B, v_posed.shape[1], 1], dtype=dtype, device=v_posed.device)
        v_h = torch.cat([v_posed, ones], dim=2)
        v_out = torch.matmul(T, v_h.unsqueeze(-1))[:, :, :3, 0]                  # lbs.py:247-250
        extra = torch.index_select(v_out, 1, self.extra_joint_ids)               # vertex_joint_selector.py:74
        joints = torch.cat([posed_joints, extra], dim=1)
        vertices = v_out
        if transl is not None:                                                   # body_models.py:387-390
            joints = joints + transl.unsqueeze(1)
            vertices = vertices + transl.unsqueeze(1)
        return OracleOutput(vertices, joints, A, posed_joints, v_posed)

    # SmplInference.forward, src/video_mocap/utils/smpl.py:28-49
    def smpl_inference(self, poses, betas, root_orient, trans):
        if betas.shape[1] != 10:
            raise ValueError("Betas array must have 10 beta values")
        out = self.forward(betas=betas, body_pose=poses, global_orient=root_orient, transl=trans, pose2rot=False)
        return {"joints": out.joints, "vertices": out.vertices}
