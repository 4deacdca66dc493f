"""Timing of the pieces of compute_nearest_points(use_mean) on a synthetic capture: forward on the visiting-order twin,
candidate vertices, candidate sums, bounded accumulation, unbounded accumulation (for comparison), and how many
(marker, vertex) entries the bounded kernel cut.   python tools/nearest_points_phases.py [frames] [permute 0|1]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import uuo_mocap_b200  # noqa: E402,F401
from uuo_mocap_b200 import ops, synthetic  # noqa: E402
from uuo_mocap_b200.smpl import SMPL, _as_struct  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
PERMUTE = (sys.argv[2] != "0") if len(sys.argv) > 2 else True     # marker columns shuffled per frame (bench default)
dev = torch.device("cuda", 0)
model = synthetic.make_model("top4")
smpl = SMPL(_as_struct(model), device=dev)
twin = smpl.visiting_twin()
cap = synthetic.make_capture(model, F, num_markers=41, seed=5, n_ghost=4, dropout_p=0.05, noise_std=0.002, permute=PERMUTE)
mk = torch.from_numpy(cap.markers).to(dev)
rot = torch.from_numpy(synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)).to(dev)
kw = dict(betas=torch.from_numpy(cap.gt_betas[:1]).to(dev), body_pose=rot[:, 1:], global_orient=rot[:, :1],
          transl=torch.from_numpy(cap.gt_trans).to(dev), pose2rot=False)


def timed(name, fn, n=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    print("%-28s %8.3f ms" % (name, e0.elapsed_time(e1) / n))
    return out


for which, m in (("twin", twin), ("model", smpl)):
    with torch.no_grad():
        v = m(**kw).vertices.contiguous()
    cand, _ = ops.mean_distance_argmin(mk[:1].contiguous(), v[:1].contiguous())
    sums = timed(which + ": candidate sums", lambda: ops.mean_distance_candidate_sums(mk, v, cand))
    bound = ops.mean_bound_from_sums(sums, F)
    acc_b = timed(which + ": bounded accumulate", lambda: ops.mean_distance_accumulate_bounded(mk, v, bound))
    _, acc = timed(which + ": full accumulate", lambda: ops.mean_distance_argmin(mk, v, finalize=False), n=1)
    cut = (acc_b < acc).float().mean().item()
    same = torch.equal(ops.mean_argmin(acc_b, F), ops.mean_argmin(acc, F))
    frac = (acc_b / acc.clamp_min(1e-30)).mean().item()
    print("%s: entries cut %.3f, mean fraction of the frames evaluated %.3f, argmin equal %s, bound/F (m) %s" % (
        which, cut, frac, same, (bound[:6] / F).cpu().numpy().round(4)))
    print("   min_v acc/F (m)", (acc.min(dim=1)[0][:6] / F).cpu().numpy().round(4), "frames with the marker at the origin",
          (mk.abs().sum(-1) == 0).float().mean(0)[:6].cpu().numpy().round(3))
    d2, _ = ops.nn_argmin_presorted(mk, v, m.visit_ids) if which == "twin" else ops.nn_argmin(mk[:2000].contiguous(), v[:2000].contiguous())
    print("   mean per-frame NN distance (m)", d2.sqrt().mean(0)[:6].cpu().numpy().round(4))
    del v, acc, acc_b
