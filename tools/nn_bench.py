"""Correspondence kernels on posed bodies (config 3 inputs): exhaustive vs box-pruned, CUDA-event timed.
    python tools/nn_bench.py [frames]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import uuo_mocap_b200  # noqa: F401
from uuo_mocap_b200 import ops, synthetic
from uuo_mocap_b200.smpl import SMPL, _as_struct

F = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
for kind in ("top4", "dense"):
    model = synthetic.make_model(kind)
    smpl = SMPL(_as_struct(model), device="cuda:0")
    cap = synthetic.make_capture(model, F, num_markers=41, seed=3, n_ghost=4, dropout_p=0.05, noise_std=0.002)
    rot = torch.from_numpy(synthetic.axis_angle_to_matrix_np(cap.gt_pose_aa.astype(np.float64)).astype(np.float32)).cuda()
    out = smpl(betas=torch.from_numpy(cap.gt_betas).cuda(), body_pose=rot[:, 1:], global_orient=rot[:, :1],
               transl=torch.from_numpy(cap.gt_trans).cuda(), pose2rot=False)
    v = out.vertices.detach().contiguous()
    mk = torch.from_numpy(cap.markers).cuda()
    twin = smpl.visiting_twin()
    ids = twin.visit_ids
    vs = v[:, ids.long()].contiguous()
    res = {}
    for name, kw in (("exhaustive", {}), ("pruned", {"model": smpl.handle}), ("presorted", None)):
        fn = (lambda: ops.nn_argmin(mk, v, **kw)) if kw is not None else (lambda: ops.nn_argmin_presorted(mk, vs, ids))
        for _ in range(3):
            d, i = fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            d, i = fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = (e0.elapsed_time(e1) / 20, d, i)
    same = all(torch.equal(res["exhaustive"][1], res[k][1]) and torch.equal(res["exhaustive"][2], res[k][2])
               for k in ("pruned", "presorted"))
    byt = F * (6890 * 12 + 45 * 20)
    print("%s F=%d: exhaustive %.3f ms, pruned (gather) %.3f ms, presorted %.3f ms (%.0f GB/s algorithmic), identical=%s"
          % (kind, F, res["exhaustive"][0], res["pruned"][0], res["presorted"][0], byt / res["presorted"][0] / 1e6, same))
