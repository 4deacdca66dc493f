"""Per-phase cycle breakdown of the presorted correspondence kernel (CTA 0).  Needs a library built with
    UUO_EXTRA_NVCC_FLAGS=-DNP_TIMING python -m uuo_mocap_b200.build --force
then: python tools/nn_phases.py [frames] [gt|prior] [mask] [hint]      (prior: body at the video prior, where the fit starts;
mask: weight-0 markers skipped through uuo_nn_argmin_presorted_weighted)"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import uuo_mocap_b200  # noqa: F401
from uuo_mocap_b200 import _lib, ops, synthetic
from uuo_mocap_b200.smpl import SMPL, _as_struct

F = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
model = synthetic.make_model("top4")
smpl = SMPL(_as_struct(model), device="cuda:0")
twin = smpl.visiting_twin()
cap = synthetic.make_capture(model, F, num_markers=41, seed=3, n_ghost=4, dropout_p=0.05, noise_std=0.002)
if "prior" in sys.argv[2:]:
    rot = torch.cat([torch.from_numpy(cap.prior_root_orient), torch.from_numpy(cap.prior_pose_body)], 1).cuda()
    betas, transl = torch.from_numpy(cap.prior_betas).cuda(), torch.from_numpy(cap.prior_trans).cuda()
else:
    rot = torch.from_numpy(synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)).cuda()
    betas, transl = torch.from_numpy(cap.gt_betas).cuda(), torch.from_numpy(cap.gt_trans).cuda()
out = twin(betas=betas, body_pose=rot[:, 1:], global_orient=rot[:, :1], transl=transl, pose2rot=False)
v = out.vertices.detach().contiguous()
mk = torch.from_numpy(cap.markers).cuda()
wts = ops.get_marker_mask(mk) if "mask" in sys.argv[2:] else None
hint = None
if "hint" in sys.argv[2:]:      # the previous iteration's matches: here the matches of the same vertices moved by 2 mm
    hint = ops.nn_argmin_presorted(mk, v + 0.002 * torch.randn_like(v[:, :1]), twin.visit_ids, want_pos=True)[2]
L = C.CDLL(_lib.LIB_PATH)
for _ in range(3):
    ops.nn_argmin_presorted(mk, v, twin.visit_ids, weights=wts, hint=hint)
torch.cuda.synchronize()
if not hasattr(L, "uuo_debug_nn_phases"):      # normal build: just run the kernel (e.g. under ncu)
    ops.nn_argmin_presorted(mk, v, twin.visit_ids, weights=wts, hint=hint)
    torch.cuda.synchronize()
    sys.exit(0)
L.uuo_debug_nn_phases(None, 1)
ops.nn_argmin_presorted(mk, v, twin.visit_ids, weights=wts, hint=hint)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 8)()
L.uuo_debug_nn_phases(buf, 0)
n = max(1, buf[7])
names = ["wait frame", "tile boxes", "round boxes", "markers (warp 0)", "waiting for other warps", "-"]
tot = sum(buf[i] for i in range(6))
for i, nm in enumerate(names):
    print("%-24s %8.0f cycles/frame  %5.1f %%" % (nm, buf[i] / n, 100.0 * buf[i] / tot))
print("frames on CTA 0: %d, cycles/frame %.0f, extra tiles/frame %.1f" % (n, tot / n, buf[6] / n))
