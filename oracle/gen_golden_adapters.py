"""Golden vectors of the video-prior adapter, produced by the REFERENCE'S OWN ImgSmpl class (TEST INFRASTRUCTURE).
    python oracle/gen_golden_adapters.py     (needs /root/reference; writes tests/golden/img_smpl.npz)
src/video_mocap/img_smpl/img_smpl.py imports pytorch3d and roma (neither installable here): the class and
utils/img_smpl_utils.py are compiled unmodified from source with those names bound to scipy-based stand-ins
(matrix <-> quaternion in pytorch3d's WXYZ order, roma's unit-quaternion slerp = scipy's Slerp between two keys)."""
import ast
import os
import sys

import numpy as np
import torch
from scipy.spatial.transform import Rotation, Slerp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import uuo_mocap_b200  # noqa: E402,F401
from uuo_mocap_b200 import synthetic  # noqa: E402

REF = "/root/reference/src/video_mocap"


def matrix_to_quaternion(m):
    q = Rotation.from_matrix(m.reshape(-1, 3, 3).double().numpy()).as_quat()          # xyzw
    return torch.from_numpy(q[:, [3, 0, 1, 2]]).reshape(m.shape[:-2] + (4,)).to(m.dtype)


def quaternion_to_matrix(q):
    r = Rotation.from_quat(q.reshape(-1, 4)[:, [1, 2, 3, 0]].double().numpy()).as_matrix()
    return torch.from_numpy(r).reshape(q.shape[:-1] + (3, 3)).to(q.dtype)


def matrix_to_axis_angle(m):
    v = Rotation.from_matrix(m.reshape(-1, 3, 3).double().numpy()).as_rotvec()
    return torch.from_numpy(v).reshape(m.shape[:-2] + (3,)).to(m.dtype)


class _RomaUtils:
    @staticmethod
    def unitquat_slerp(q0, q1, steps):
        # the reference hands pytorch3d's WXYZ quaternions to roma (XYZW): the great arc between the two 4-vectors does
        # not depend on the component order, so read them as WXYZ rotations and return WXYZ
        a = float(steps.reshape(-1)[0])
        out = []
        for x, y in zip(q0.reshape(-1, 4).double().numpy(), q1.reshape(-1, 4).double().numpy()):
            rr = Rotation.from_quat(np.stack([x[[1, 2, 3, 0]], y[[1, 2, 3, 0]]]))
            out.append(Slerp([0.0, 1.0], rr)([a]).as_quat()[0][[3, 0, 1, 2]])
        return torch.from_numpy(np.stack(out)).reshape(q0.shape).to(q0.dtype)


class _Roma:
    utils = _RomaUtils


def take(path, names, assigns=()):
    tree = ast.parse(open(path).read())
    nodes = [n for n in tree.body if (isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names)
             or (isinstance(n, ast.Assign) and any(getattr(t, "id", None) in assigns for t in n.targets))]
    return compile(ast.Module(body=nodes, type_ignores=[]), path, "exec")


ns = {"np": np, "torch": torch, "os": os, "roma": _Roma, "matrix_to_axis_angle": matrix_to_axis_angle,
      "matrix_to_quaternion": matrix_to_quaternion, "quaternion_to_matrix": quaternion_to_matrix}
exec(take(REF + "/utils/img_smpl_utils.py", ["get_foot_contacts"], ["joints_2d_labels", "joints_3d_labels"]), ns)
exec(take(REF + "/img_smpl/img_smpl.py", ["ImgSmpl"]), ns)
data = synthetic.make_phalp_like(40, seed=17)
ref = ns["ImgSmpl"](data, 30.0)
smpl = ref.get_smpl()
out = {"trans": ref.trans.numpy(), "root_orient": ref.root_orient.numpy(), "hmr_root_orient": ref.hmr_root_orient.numpy(),
       "pose_body": ref.pose_body.numpy(), "betas": ref.betas.numpy(), "img_mask": ref.img_mask.numpy(),
       "camera_bbox": ref.camera_bbox.numpy(), "center": ref.center.numpy(), "scale": ref.scale.numpy(),
       "size": ref.size.numpy(), "foot_contacts": ref.foot_contacts.numpy(), "poses_aa": smpl["poses"],
       "smpl_betas": smpl["betas"], "smpl_trans": smpl["trans"]}
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "img_smpl.npz"), **out)
print({k: v.shape for k, v in out.items()}, int(ref.img_mask.sum()))
