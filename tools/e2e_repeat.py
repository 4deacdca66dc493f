"""Repeatability of the end-to-end call of bench.py's e2e leg (fit100k, one GPU): N un-synchronised calls in a row, per
call the wall time, the caching allocator's new segments (cudaMalloc calls) and what the cyclic GC collected.
    python tools/e2e_repeat.py [frames] [iters] [calls] [gc: 0|1]"""
import gc
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import uuo_mocap_b200  # noqa: E402,F401
from uuo_mocap_b200 import fit as fitm, synthetic  # noqa: E402
from uuo_mocap_b200.dist import Shard  # noqa: E402
from uuo_mocap_b200.smpl import SMPL, _as_struct  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
N = int(sys.argv[3]) if len(sys.argv) > 3 else 8
GC = (sys.argv[4] != "0") if len(sys.argv) > 4 else False
dev = torch.device("cuda", 0)
model = synthetic.make_model("top4")
smpl = SMPL(_as_struct(model), device=dev)
cap = synthetic.make_capture(model, F, num_markers=41, seed=5, n_ghost=4, dropout_p=0.05, noise_std=0.002)
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
h = {"markers": T(cap.markers), "pose": T(cap.prior_pose_body), "root": T(cap.prior_root_orient), "trans": T(cap.prior_trans)}
o_betas = torch.from_numpy(cap.prior_betas)
cfg = fitm.default_config(num_iters=K, optimizer="adam", lr=0.01)
cfg["stages"]["chamfer"]["losses"]["trans_vel"] = 1.0
cfg["stages"]["marker"]["losses"]["temporal"] = 1.0
shard = Shard()


def once():
    mk = h["markers"].to(dev, non_blocking=True)
    pose = h["pose"].to(dev, non_blocking=True)
    opose = pose.clone()
    rt, tr = h["root"].to(dev, non_blocking=True), h["trans"].to(dev, non_blocking=True)
    bt = o_betas.reshape(1, -1).to(dev)
    p = fitm.FitProblem(smpl, mk, opose, o_betas, shard, 1)
    fitm.optim_chamfer(mk, pose, opose, bt, o_betas, rt, tr, None, None, smpl, cfg, problem=p)
    vi = fitm.compute_nearest_points(mk, pose, bt, rt, tr, smpl, config=cfg, shard=shard, as_one_hot=False, problem=p)
    fitm.optim_markers(mk, pose, opose, bt, o_betas, rt, tr, vi, None, smpl, cfg, problem=p)
    out = [x.to("cpu") for x in (pose, rt, tr, bt)]
    torch.cuda.current_stream().synchronize()
    return out


gc.disable() if not GC else None
for i in range(N):
    if GC:
        gc.collect()
    torch.cuda.synchronize()
    s0 = torch.cuda.memory_stats()
    t0 = time.perf_counter()
    once()
    dt = time.perf_counter() - t0
    s1 = torch.cuda.memory_stats()
    print("call %d  %.1f ms  new segments %d  reserved %.1f GB  allocated after %.1f GB  gc counts %s" % (
        i, dt * 1e3, s1["segment.all.allocated"] - s0["segment.all.allocated"], s1["reserved_bytes.all.current"] / 1e9,
        s1["allocated_bytes.all.current"] / 1e9, gc.get_count()))
    if not GC and i == N // 2:
        print("  gc.collect() ->", gc.collect(), "objects")
