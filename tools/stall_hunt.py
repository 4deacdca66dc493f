import os, sys, time
import numpy as np, torch
sys.path.insert(0, '/root/repo')
import uuo_mocap_b200
from uuo_mocap_b200 import fit as fitm, synthetic, _lib
from uuo_mocap_b200.dist import Shard
from uuo_mocap_b200.smpl import SMPL, _as_struct
F, K = 100000, 10
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
import uuo_mocap_b200.fit as F_
orig_run_adam = F_.run_adam
sub = []
def run_adam_timed(*a, **k):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = orig_run_adam(*a, **k)
    torch.cuda.synchronize(); sub.append(round((time.perf_counter() - t0) * 1e3, 1))
    return r
F_.run_adam = run_adam_timed
def once():
    marks = []
    def mark(n):
        torch.cuda.synchronize(); marks.append((n, time.perf_counter()))
    mark("s")
    mk = h["markers"].to(dev, non_blocking=True); pose = h["pose"].to(dev, non_blocking=True); opose = pose.clone()
    rt, tr = h["root"].to(dev, non_blocking=True), h["trans"].to(dev, non_blocking=True)
    bt = o_betas.reshape(1, -1).to(dev)
    mark("h2d")
    p = fitm.FitProblem(smpl, mk, opose, o_betas, shard, 1); mark("prob")
    fitm.optim_chamfer(mk, pose, opose, bt, o_betas, rt, tr, None, None, smpl, cfg, problem=p); mark("chamfer")
    vi = fitm.compute_nearest_points(mk, pose, bt, rt, tr, smpl, config=cfg, shard=shard, as_one_hot=False, problem=p); mark("nearest")
    fitm.optim_markers(mk, pose, opose, bt, o_betas, rt, tr, vi, None, smpl, cfg, problem=p); mark("markers")
    return {b[0]: round((b[1] - a[1]) * 1e3, 1) for a, b in zip(marks[:-1], marks[1:])}
for i in range(int(os.environ.get("CALLS", "14"))):
    sub.clear()
    ph = once()
    tot = sum(ph.values())
    print(i, round(tot, 1), ph if tot > 260 or i < 2 else "", "run_adam:", list(sub))
