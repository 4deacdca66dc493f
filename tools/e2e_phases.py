"""Where the end-to-end time of one fit call goes (bench.py's e2e leg, fit100k): every phase bracketed by a
synchronize.  python tools/e2e_phases.py [frames] [iters]   (or under torchrun for the sharded case)."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import uuo_mocap_b200  # noqa: E402,F401
from uuo_mocap_b200 import fit as fitm, synthetic  # noqa: E402
from uuo_mocap_b200.dist import Shard, frame_range  # noqa: E402
from uuo_mocap_b200.smpl import SMPL, _as_struct  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
model = synthetic.make_model("top4")
smpl = SMPL(_as_struct(model), device=dev)
cap = synthetic.make_capture(model, F, num_markers=41, seed=5, n_ghost=4, dropout_p=0.05, noise_std=0.002)
lo, hi = frame_range(F, rank, world)
shard = Shard.from_env(peer_device=dev) if world > 1 else Shard()
T = lambda a: torch.from_numpy(np.ascontiguousarray(a[lo:hi])).pin_memory()
h = {"markers": T(cap.markers), "pose": T(cap.prior_pose_body), "root": T(cap.prior_root_orient), "trans": T(cap.prior_trans)}
o_betas = torch.from_numpy(cap.prior_betas)
cfg = fitm.default_config(num_iters=K, optimizer="adam", lr=0.01)
cfg["stages"]["chamfer"]["losses"]["trans_vel"] = 1.0
cfg["stages"]["marker"]["losses"]["temporal"] = 1.0


def once(report):
    marks = []

    def mark(name):
        torch.cuda.synchronize()
        marks.append((name, time.perf_counter()))
    mark("start")
    mk = h["markers"].to(dev, non_blocking=True)
    pose = h["pose"].to(dev, non_blocking=True)
    opose = pose.clone()
    rt, tr = h["root"].to(dev, non_blocking=True), h["trans"].to(dev, non_blocking=True)
    bt = o_betas.reshape(1, -1).to(dev)
    mark("h2d")
    p = fitm.FitProblem(smpl, mk, opose, o_betas, shard, 1)
    mark("FitProblem")
    fitm.optim_chamfer(mk, pose, opose, bt, o_betas, rt, tr, None, None, smpl, cfg, problem=p)
    mark("optim_chamfer")
    vi = fitm.compute_nearest_points(mk, pose, bt, rt, tr, smpl, config=cfg, shard=shard, as_one_hot=False, problem=p)
    mark("compute_nearest_points")
    fitm.optim_markers(mk, pose, opose, bt, o_betas, rt, tr, vi, None, smpl, cfg, problem=p)
    mark("optim_markers")
    out = [x.to("cpu") for x in (pose, rt, tr, bt)]
    mark("d2h")
    if report and rank == 0:
        for (a, ta), (b, tb) in zip(marks[:-1], marks[1:]):
            print("%-24s %8.2f ms" % (b, (tb - ta) * 1e3))
        print("%-24s %8.2f ms" % ("total", (marks[-1][1] - marks[0][1]) * 1e3))
    return out


once(False)
once(True)
once(True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
