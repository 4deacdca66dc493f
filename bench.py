#!/usr/bin/env python
"""bench.py -- frame-iterations/s of SMPL marker fitting (BASELINE.json `metric`) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload fit100k|fit5k|mht|lbs] [--impl ours|reference]

Default workload = BASELINE.json configs[4] ("fit100k"): ONE 100 000-frame capture, 41 markers + 4 ghosts, noise,
dropout, synthetic video-pose prior, video_mocap.yaml loss weights plus the trans_vel / temporal terms, cut into
contiguous frame ranges over the N ranks (strong scaling; per iteration the ranks exchange the 1-/2-frame halo and one
[d betas | loss] sum over NVLink peer memory).  One "step" = one fitting iteration over every frame = closure (SMPL
forward, correspondence or marker loss, priors, temporal term, full backward) + fused Adam update; the K timed steps
are ceil(K/2) chamfer-stage iterations followed by floor(K/2) marker-stage iterations (the two stages of the yaml).
  value : parameters / markers / prior resident in HBM, C-ABI calls replayed from one CUDA graph per stage,
          CUDA-event timed between barriers, max over ranks
  e2e   : the reference's stage functions (optim_chamfer -> compute_nearest_points -> optim_markers) on pinned HOST
          tensors: H2D of the capture, the same number of iterations, D2H of the fitted parameters, all inside the
          timed region
  roofline : dominant kernel of the step, timed live with CUDA events inside the library (a separate eager pass)
  parity   : 64 sampled frames of the first closure evaluation checked against the CPU oracle (vertices, indices, grads)
  cpu_baseline : the oracle closures + torch.optim.Adam on the host cores, bounded sample of the same capture
`--impl reference` times that CPU path alone on the same config with the same JSON contract.
`--workload lbs` is BASELINE.json configs[1] (lbs() forward+backward microbench, weak scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

# Libraries (NCCL's version banner, torch warnings) may write to fd 1; the contract is ONE JSON line on stdout.
# Keep the real stdout aside, point fd 1 at stderr for everything else, and emit the line through _emit().
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def _emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


METRIC = "frame_iters_per_sec"
UNIT = "frame-iters/s"
V, J, NE, NB, KC = 6890, 24, 21, 10, 217


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": float(d["hbm_gbs"]), "bf16": float(d["bf16_tflops"]), "src": "measured"}
    return {"hbm": 6650.0, "bf16": 1590.0, "src": "fallback"}


class ClockSampler:
    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_oracle_rate(frames: int, reps: int, threads: int):
    """Oracle lbs() forward+backward on the host cores: (frame-iters/s, seconds per rep)."""
    import uuo_mocap_b200  # noqa: F401
    from uuo_mocap_b200 import synthetic
    from oracle.smpl_oracle import OracleSMPL
    torch.set_num_threads(threads)
    orc = OracleSMPL.from_model(synthetic.make_model("dense"))
    mb = synthetic.make_lbs_microbench(frames, seed=2)
    gv, gj = torch.from_numpy(mb["g_vertices"]), torch.from_numpy(mb["g_joints"])
    times = []
    for it in range(reps + 1):
        rot = torch.from_numpy(mb["rot"]).requires_grad_(True)
        betas = torch.from_numpy(mb["betas"]).requires_grad_(True)
        transl = torch.from_numpy(mb["transl"]).requires_grad_(True)
        t0 = time.perf_counter()
        o = orc.forward(betas, rot[:, 1:], rot[:, :1], transl, pose2rot=False)
        torch.autograd.backward([o.vertices, o.joints], [gv, gj])      # upstream gradients are inputs of the workload
        dt = time.perf_counter() - t0
        if it > 0:
            times.append(dt)
    med = float(np.median(times))
    return frames / med, med


def torch_gpu_oracle_rate(frames: int, reps: int, dev, seed: int):
    """SURVEY.md section 8(d) "composition" bar: the oracle's lbs() forward+backward (the reference's own sequence of
    torch ops) on the B200 through stock PyTorch CUDA kernels (ATen + cuBLAS, fp32, TF32 off like the reference).
    A reported comparison only -- nothing of the product path runs here.  -> (frame-iters/s, ms per step)."""
    import uuo_mocap_b200  # noqa: F401
    from uuo_mocap_b200 import synthetic
    from oracle.smpl_oracle import OracleSMPL
    orc = OracleSMPL.from_model(synthetic.make_model("dense")).to(dev)
    mb = synthetic.make_lbs_microbench(frames, seed=seed)
    T = lambda k: torch.from_numpy(mb[k]).to(dev)
    rot0, betas0, tr0, gv, gj = T("rot"), T("betas"), T("transl"), T("g_vertices"), T("g_joints")

    def step():
        rot, betas, transl = (x.clone().requires_grad_(True) for x in (rot0, betas0, tr0))
        o = orc.forward(betas, rot[:, 1:], rot[:, :1], transl, pose2rot=False)
        torch.autograd.backward([o.vertices, o.joints], [gv, gj])
    for _ in range(3):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return frames / (ms * 1e-3), ms


def run_reference_lbs(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.ref_frames
    # one "step" = one bounded sample of the workload
    rates = []
    t_all0 = time.perf_counter()
    rate, sec = cpu_oracle_rate(sample, max(1, args.steps + args.warmup - 1), threads)
    wall = time.perf_counter() - t_all0
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD % args.frames, "frames_per_gpu": args.frames,
                   "parallelism": "frame-range shards x%d" % world,
                   "l2": "inputs larger than L2 (>= 680 MB streamed per step)"},
        "note": "reference smplx/PyTorch lbs() fwd+bwd on host cores (oracle port, bit-identical to the vendored smplx); "
                "each step runs %d frames" % sample,
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%d of %d frames per step, torch threads=%d" % (sample, args.frames, threads)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
    }
    _emit(line)


# ------------------------------------------------------------------------------------------------
# fitting-loop workloads (BASELINE.json configs[4] "fit100k" = the default, configs[2] "fit5k", configs[3] "mht")
# ------------------------------------------------------------------------------------------------
FIT_TRAFFIC_KEYS = ("nn_argmin_presorted", "rows_forward", "rows_backward", "adam_step")
FIT_WORKLOADS = {
    # name: (description, total frames, hypotheses, scaling, temporal terms)
    "fit5k": ("configs[2]: video_mocap.yaml fit, 5000 frames per GPU, 41 markers + 4 ghosts, noise 2 mm, dropout 5 %, "
              "synthetic video-pose prior", 5000, 1, "weak", False),
    "mht": ("configs[3]: mht_rotation.yaml, 64 yaw hypotheses x 2000 frames stacked along the frame axis and sharded by "
            "hypothesis", 2000, 64, "strong", False),
    "fit100k": ("configs[4]: 100k-frame long-capture marker fit sharded by frame range, trans_vel + temporal terms on "
                "(1- and 2-frame halo at shard boundaries)", 100000, 1, "strong", True),
}
LOSS_CH = {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0}      # config/video_mocap.yaml:19-31
LOSS_MK = {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0}             # config/video_mocap.yaml:41-50


def fit_losses(temporal):
    lc, lm = dict(LOSS_CH), dict(LOSS_MK)
    if temporal:
        lc["trans_vel"] = 1.0
        lm["temporal"] = 1.0
    return lc, lm


def fit_config(args, world):
    """The `config` object of the JSON line: identical on the `ours` and `reference` arms (what differs between the
    arms -- graph replay, the CPU arm's frame sample -- is reported outside it, in `run` / `cpu_baseline.sample`)."""
    desc, Ftot, H, scaling, temporal = FIT_WORKLOADS[args.workload]
    if args.frames:
        Ftot = args.frames
    lc, lm = fit_losses(temporal)
    per = "per GPU" if scaling == "weak" else "in total"
    return {"workload": desc, "name": args.workload, "frames": Ftot, "frames_are": per, "hypotheses": H,
            "markers": 45 if args.workload != "mht" else 41,
            "model": "synthetic SMPL-shaped (6890 vertices, 24 joints, 10 betas, 207 posedirs, top-4 skinning weights)",
            "losses": {"chamfer": lc, "marker": lm}, "optimizer": "Adam lr 0.01 (fused on the GPU arm, torch.optim.Adam on "
            "the CPU arm)", "step": "one fitting iteration over all frames (closure + Adam); K steps = ceil(K/2) "
            "chamfer-stage + floor(K/2) marker-stage iterations", "parallelism": "frame-range shards x%d" % world,
            "l2": "per-iteration working set (vertices of this rank's frames, >= 400 MB) exceeds L2: streamed through HBM"}


def make_fit_capture(args, rank, world, model):
    """-> (capture, lo, hi, frames_global, H): this rank's slice [lo, hi) of the synthetic capture."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.dist import frame_range
    desc, Ftot, H, scaling, temporal = FIT_WORKLOADS[args.workload]
    if args.frames:
        Ftot = args.frames
    n_ghost, drop = (0, 0.0) if args.workload == "mht" else (4, 0.05)
    if scaling == "weak":            # independent sequences: no cross-rank term
        cap = synthetic.make_capture(model, Ftot, num_markers=41, seed=3 + rank, n_ghost=n_ghost, dropout_p=drop,
                                     noise_std=0.002)
        return cap, 0, Ftot, Ftot * world, H
    if H > 1:                        # hypotheses are independent; sharded by hypothesis
        cap = synthetic.make_capture(model, Ftot, num_markers=41, seed=4, noise_std=0.002)
        return cap, 0, Ftot, Ftot * H, H
    cap = synthetic.make_capture(model, Ftot, num_markers=41, seed=5, n_ghost=n_ghost, dropout_p=drop, noise_std=0.002)
    lo, hi = frame_range(Ftot, rank, world)
    return cap, lo, hi, Ftot, H


def cpu_fit_steps(cap, model, frames, temporal, n_ch, n_mk, warm, device=None):
    """Oracle chamfer / marker closures (forward + backward) + torch.optim.Adam on the first `frames` frames of the
    capture, on the host cores (or, with `device`, the same torch ops through stock PyTorch CUDA kernels: the SURVEY.md
    8(d) "composition" bar).  `warm` untimed iterations per stage, then n_ch + n_mk timed ones.
    -> dict(seconds, chamfer_s, marker_s per iteration)."""
    from oracle import closures
    from oracle.smpl_oracle import OracleSMPL
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    orc = OracleSMPL.from_model(model)
    on_gpu = device is not None
    if on_gpu:
        orc = orc.to(device)
    sync = (lambda: torch.cuda.synchronize(device)) if on_gpu else (lambda: None)
    frames = min(frames, cap.markers.shape[0])
    sl = slice(0, frames)
    where = device if on_gpu else "cpu"
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a[sl])).to(where)
    mk, opose, obetas = T(cap.markers), T(cap.prior_pose_body), torch.from_numpy(cap.prior_betas).to(where)
    root, trans = T(cap.prior_root_orient), T(cap.prior_trans)
    vidx = torch.from_numpy(np.resize(cap.marker_vertex_ids, mk.shape[1]).astype(np.int64)).to(mk.device)
    lc, lm = fit_losses(temporal)
    res = {}
    for stage, n in (("chamfer", n_ch), ("marker", n_mk)):
        pose = opose.clone().requires_grad_(True)
        betas = obetas.clone().requires_grad_(True)
        tr = trans.clone().requires_grad_(True)
        if stage == "chamfer":
            z = torch.zeros(frames, 1, 1, device=mk.device, requires_grad=True)
            params = [tr, z, betas, pose]
            fn = lambda: closures.chamfer_stage_loss(orc, mk, pose, opose, betas, obetas, root, tr, z, lc)
        else:
            rt = root.clone().requires_grad_(True)
            params = [pose, betas, rt, tr]
            fn = lambda: closures.marker_stage_loss(orc, mk, pose, opose, betas, obetas, rt, tr, vidx, lm)
        opt = torch.optim.Adam(params, lr=0.01)
        ts = []
        for i in range(warm + n):
            sync()
            t0 = time.perf_counter()
            opt.zero_grad()
            fn().backward()
            opt.step()
            sync()
            if i >= warm:
                ts.append(time.perf_counter() - t0)
        res[stage + "_s"] = float(np.mean(ts)) if ts else 0.0
        res[stage + "_total_s"] = float(np.sum(ts))
    res["seconds"] = res["chamfer_total_s"] + res["marker_total_s"]
    res["frames"] = frames
    res["threads"] = threads
    return res


def run_reference_fit(args, rank, world):
    """`--impl reference`: the reference's CPU implementation of the same fitting iteration (oracle port of
    optimization.py:187-275 / :329-394 + torch.optim.Adam) on the host cores with all threads, same config; each step
    runs on a bounded contiguous sample of the capture (the full 100 000 frames would take ~50 s per step)."""
    if rank != 0:
        return
    import uuo_mocap_b200  # noqa: F401
    from uuo_mocap_b200 import synthetic
    desc, Ftot, H, scaling, temporal = FIT_WORKLOADS[args.workload]
    model = synthetic.make_model("top4")
    sample = args.ref_frames
    n_ghost, drop = (0, 0.0) if args.workload == "mht" else (4, 0.05)
    seed = {"fit5k": 3, "mht": 4, "fit100k": 5}[args.workload]
    # a capture of `sample` frames from the same generator and seed (same statistics: motion bandwidth, noise, dropout,
    # ghosts; the per-frame cost of the CPU path does not depend on which frames are drawn)
    cap = synthetic.make_capture(model, sample, num_markers=41, seed=seed, n_ghost=n_ghost, dropout_p=drop, noise_std=0.002)
    K = args.steps
    n_ch, n_mk = (K + 1) // 2, K // 2
    t0 = time.perf_counter()
    r = cpu_fit_steps(cap, model, sample, temporal, n_ch, n_mk, max(1, args.warmup // 2))
    wall = time.perf_counter() - t0
    rate = sample * K / r["seconds"]
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": K,
        "warmup": args.warmup, "ms_per_step": r["seconds"] / K * 1e3, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": fit_config(args, world),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": r["threads"], "kind": "port",
                         "sample": "each step = one iteration over a %d-frame contiguous sample of the capture "
                                   "(chamfer %.3f s, marker %.3f s per iteration), torch threads=%d"
                                   % (sample, r["chamfer_s"], r["marker_s"], r["threads"])},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
        "note": "reference path on host cores: oracle port of the stage closures (vendored-smplx lbs + restated pytorch3d "
                "chamfer) + torch.optim.Adam; the Python reference itself cannot be installed here (pytorch3d absent)",
    }
    _emit(line)


def fit_parity_sample(prob, pb, root_d, cap, lo, model, n_sample=64, check=True):
    """Self-check of the bench run: ONE chamfer-closure evaluation (full_chamfer term only, so the
    per-frame gradients do not depend on the other frames) over all of this rank's frames on the CUDA path, then
    `n_sample` frames spread over the range are checked against the CPU oracle run on just those frames:
    vertices (1e-5 relative, per frame), correspondence indices (bit-exact on identical vertices), d loss / d pose,
    d trans, d yaw (1e-4 relative).  Frames are independent in this term, so the sample is a fair check of the whole."""
    from uuo_mocap_b200 import fit as fitm
    from oracle import closures, p3d_oracle as p3d
    from oracle.smpl_oracle import OracleSMPL
    F = prob.F
    pb.grad.zero_()
    fitm.chamfer_closure(prob, pb, root_d, {"full_chamfer": 10.0})
    torch.cuda.synchronize()
    if not check:                       # the closure holds the shards' exchange: every rank runs it, rank 0 checks
        return None
    sel = np.unique(np.linspace(0, F - 1, n_sample).round().astype(np.int64))
    st = torch.from_numpy(sel).to(prob.dev)
    ids = prob.twin.visit_ids.long()                     # row r of the twin's vertex array = original vertex ids[r]
    v_gpu = torch.empty((len(sel), prob.mh.V, 3), device=prob.dev)
    v_gpu[:, ids] = prob.vertices[st]
    v_gpu = v_gpu.cpu()
    idx_gpu = prob.idx[st].cpu().long()
    mask = prob.mask[st].cpu()
    g_gpu = {k: pb.gviews[k][st].cpu().numpy() for k in ("pose_body", "trans", "z_angle")}
    pose = pb.views["pose_body"][st].cpu().clone().requires_grad_(True)
    trans = pb.views["trans"][st].cpu().clone().requires_grad_(True)
    z = pb.views["z_angle"][st].cpu().clone().reshape(-1, 1, 1).requires_grad_(True)
    G = prob.groups                                       # stacked hypotheses: one betas row and one normaliser per group
    betas_all = pb.views["betas"].cpu().clone().reshape(G, -1)
    grp = sel // (F // G)
    root = root_d[st].cpu()
    mk = prob.markers[st].cpu()
    orc = OracleSMPL.from_model(model)
    torch.set_num_threads(os.cpu_count() or 1)
    # the oracle normalises by the weight sum of the frames it is given, the run by the group's global one
    scale = torch.ones(len(sel))
    v_ref = torch.empty((len(sel), prob.mh.V, 3))
    for g in np.unique(grp):
        w = torch.from_numpy(np.where(grp == g)[0])
        bg = betas_all[g:g + 1]
        loss = closures.chamfer_stage_loss(orc, mk[w], pose[w], pose[w].detach(), bg, bg, root[w], trans[w], z[w],
                                           {"full_chamfer": 10.0})
        loss.backward()
        scale[w] = float(mask[w].sum()) / prob.wsum_global
        with torch.no_grad():
            zr = p3d.compute_root_orient_z(z[w]) @ root[w]
            v_ref[w] = orc.smpl_inference(p3d.normalize_rot(pose[w]), bg.repeat(len(w), 1), p3d.normalize_rot(zr),
                                          trans[w])["vertices"]
    per_frame = lambda a, b: float((np.abs(a - b).reshape(len(sel), -1).max(1)
                                    / np.maximum(np.abs(b).reshape(len(sel), -1).max(1), 1e-30)).max())
    v_err = per_frame(v_gpu.numpy(), v_ref.numpy())
    _, idx_ref = p3d.knn_k1(mk, v_gpu)                   # the oracle's argmin on the SAME vertices: must agree exactly
    on = mask != 0
    idx_ok = bool(torch.equal(idx_gpu[on], idx_ref[on]))
    rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
    sc = scale.numpy()
    g_err = max(rel(g_gpu["pose_body"], pose.grad.numpy() * sc.reshape(-1, 1, 1, 1)),
                rel(g_gpu["trans"], trans.grad.numpy() * sc.reshape(-1, 1)),
                rel(g_gpu["z_angle"], z.grad.numpy().reshape(-1) * sc))
    return {"frames_checked": int(len(sel)), "of_frames": int(F), "vertices_rel_err_per_frame": v_err, "vertices_tol": 1e-5,
            "idx_bit_exact": idx_ok, "markers_compared": int(on.sum()), "grad_rel_err": g_err, "grad_tol": 1e-4,
            "ok": bool(v_err < 1e-5 and idx_ok and g_err < 1e-4),
            "against": "oracle/ (CPU restatement pinned on the vendored smplx; pytorch3d ops restated)"}


def run_fit(args, rank, world, local, dev, light=False):
    """K steps = ceil(K/2) chamfer-stage iterations (full lbs() forward -> correspondence -> row-sparse backward ->
    priors [-> trans_vel halo] -> shard sum -> Adam) then floor(K/2) marker-stage iterations (row-sparse forward /
    backward [-> temporal halo] -> shard sum -> Adam) over this rank's frames.  `light`: device-timed numbers only
    (used for the `other_workloads` entries of the default line), returns the dict instead of printing it."""
    import uuo_mocap_b200  # noqa: F401
    from uuo_mocap_b200 import synthetic, ops, fit as fitm
    from uuo_mocap_b200.dist import Shard, frame_range
    from uuo_mocap_b200.smpl import SMPL, _as_struct

    desc, _, _, scaling, temporal = FIT_WORKLOADS[args.workload]
    model = synthetic.make_model("top4")
    smpl = SMPL(_as_struct(model), device=dev)
    cap, lo, hi, frames_global, H = make_fit_capture(args, rank, world, model)
    Ftot = cap.markers.shape[0]
    sharded = scaling == "strong" and H == 1 and world > 1
    shard_fit = Shard.from_env(peer_device=dev) if sharded else Shard()
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a[lo:hi]))
    markers, o_pose, o_betas = T(cap.markers), T(cap.prior_pose_body), torch.from_numpy(cap.prior_betas)
    root, trans = T(cap.prior_root_orient), T(cap.prior_trans)
    if H > 1:                               # this rank's hypotheses, stacked along the frame axis
        h0, h1 = frame_range(H, rank, world)
        Hl = h1 - h0
        ang = torch.arange(h0, h1, dtype=torch.float32) * (2 * np.pi / H)
        rep = lambda t: t.repeat((Hl,) + (1,) * (t.dim() - 1)).contiguous()
        root = (fitm.rot_z(ang.view(Hl, 1, 1, 1)) @ root.unsqueeze(0)).reshape(Hl * Ftot, 1, 3, 3)
        markers, o_pose, trans = rep(markers), rep(o_pose), rep(trans)
        groups = Hl
    else:
        groups = 1
    F, M = markers.shape[0], markers.shape[1]
    prob = fitm.FitProblem(smpl, markers, o_pose, o_betas, shard_fit, groups)
    betas0 = o_betas.reshape(1, -1).repeat(groups, 1)
    pb_ch = fitm.ParamBlock.build({"trans": trans, "z_angle": torch.zeros(F), "betas": betas0, "pose_body": o_pose}, dev)
    pb_mk = fitm.ParamBlock.build({"pose_body": o_pose, "betas": betas0, "root_orient": root, "trans": trans}, dev)
    root_d = root.to(dev).contiguous()
    loss_ch, loss_mk = fit_losses(temporal)
    vidx = torch.from_numpy(np.resize(cap.marker_vertex_ids, M).astype(np.int32)).to(dev)
    if groups > 1:  # every hypothesis carries its own correspondences (fit.fit_hypotheses): indices [groups, M]
        vidx = vidx.reshape(1, M).expand(groups, M).contiguous()
    state = {k: (torch.zeros_like(pb.flat), torch.zeros_like(pb.flat)) for k, pb in (("ch", pb_ch), ("mk", pb_mk))}
    counters = {k: torch.zeros(1, dtype=torch.int32, device=dev) for k in ("ch", "mk")}
    graphs = {}
    use_graph = shard_fit.graph_safe and not args.no_graph

    parity = None
    if not light and not args.no_parity:
        parity = fit_parity_sample(prob, pb_ch, root_d, cap, lo, model, check=rank == 0)   # every rank runs the closure
        prob.pos.fill_(-1)

    def step(stage):
        pb = pb_ch if stage == "ch" else pb_mk
        if stage == "ch":
            fitm.chamfer_closure(prob, pb, root_d, loss_ch)
        else:
            fitm.marker_closure(prob, pb, vidx, loss_mk)
        ops.adam_step_dev(pb.flat, pb.grad, state[stage][0], state[stage][1], 0.01, counters[stage])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_n(stage, n, replay):
        for _ in range(n):
            if replay:
                graphs[stage].replay()
            else:
                step(stage)

    K = args.steps
    n_ch, n_mk = (K + 1) // 2, K // 2
    for stage in ("ch", "mk"):
        for _ in range(args.warmup):
            step(stage)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    ev = lambda: torch.cuda.Event(enable_timing=True)

    # ---- pass 1 (eager, the library's per-kernel events on, one stage after the other): the per-iteration kernel times
    # behind `roofline` (a kernel may be launched several times per iteration, e.g. once per frame chunk: times are summed)
    def profiled(stage, n):
        ops.profile_enable(True)
        a, b = ev(), ev()
        barrier()
        a.record(); run_n(stage, n, False); b.record()
        barrier()
        pr = ops.profile_read()
        ops.profile_enable(False)
        return {k: (v[0] / max(n, 1), v[1] / max(n, 1)) for k, v in (pr or {}).items()}, a.elapsed_time(b)
    prof_ch, pms_ch = profiled("ch", n_ch)
    prof_mk, pms_mk = profiled("mk", n_mk)
    prof_ms = pms_ch + pms_mk
    # ms per step-pair of stages summed per kernel, launches per iteration
    prof = {}
    for pr, n in ((prof_ch, n_ch), (prof_mk, n_mk)):
        for k, (ms_it, cnt_it) in pr.items():
            t, c = prof.get(k, (0.0, 0.0))
            prof[k] = (t + ms_it * n, c + cnt_it * n)

    # ---- pass 2 (the timed K steps): one captured iteration per stage replayed (eager when graphs are off)
    if use_graph:
        for stage in ("ch", "mk"):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local" if world > 1 else "global"):
                step(stage)
            graphs[stage] = g
        run_n("ch", 1, True); run_n("mk", 1, True)
    e0, e1, e2 = ev(), ev(), ev()
    barrier()
    e0.record(); run_n("ch", n_ch, use_graph); e1.record(); run_n("mk", n_mk, use_graph); e2.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e2), e0.elapsed_time(e1), e1.elapsed_time(e2)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_all, ms_ch, ms_mk = (float(x) for x in t.tolist())
    value = frames_global * K / (ms_all * 1e-3)

    def stage_rates():
        return {"chamfer": {"iters": n_ch, "ms_per_iter": ms_ch / max(n_ch, 1),
                            "frame_iters_per_sec": frames_global * n_ch / (ms_ch * 1e-3) if n_ch else None},
                "marker": {"iters": n_mk, "ms_per_iter": ms_mk / max(n_mk, 1),
                           "frame_iters_per_sec": frames_global * n_mk / (ms_mk * 1e-3) if n_mk else None}}

    kern = {"chamfer_iteration": {k: round(v[0], 4) for k, v in prof_ch.items()},
            "marker_iteration": {k: round(v[0], 4) for k, v in prof_mk.items()}}
    if light:
        if sampler:
            sampler.stop()
        return {"value": value, "unit": UNIT, "ms_per_step": ms_all / K, "steps": K, "frames_this_rank": F,
                "stages": stage_rates(), "ms_per_kernel": kern, "cuda_graph": bool(use_graph)}

    # ---- e2e: the reference's stage functions with HOST tensors in and out (optim_chamfer -> compute_nearest_points
    # -> optim_markers), the same n_ch / n_mk Adam iterations, copies inside the timed region
    cfg = fitm.default_config(num_iters=n_ch, optimizer="adam", lr=0.01)
    cfg["stages"]["marker"]["num_iters"] = n_mk
    cfg["stages"]["chamfer"]["losses"], cfg["stages"]["marker"]["losses"] = dict(loss_ch), dict(loss_mk)
    pin = lambda x: x.clone().pin_memory()
    h = {"markers": pin(markers), "pose": pin(o_pose), "root": pin(root), "trans": pin(trans), "betas": pin(betas0)}
    h2d = sum(x.numel() * 4 for x in h.values())
    d2h = sum(h[k].numel() * 4 for k in ("pose", "root", "trans", "betas"))
    h_out = {k: torch.empty_like(h[k]).pin_memory() for k in ("pose", "root", "trans", "betas")}
    del prob, pb_ch, pb_mk, state, graphs
    torch.cuda.empty_cache()

    def e2e_once(marks=None):
        def mark(name):
            if marks is not None:                 # diagnostic call only: a synchronize between the phases
                torch.cuda.synchronize()
                marks.append((name, time.perf_counter()))
        mark("start")
        mk = h["markers"].to(dev, non_blocking=True)
        pose = h["pose"].to(dev, non_blocking=True)
        opose = pose.clone()                      # the prior pose starts as the same tensor (multimodal.py): copied on the device
        rt, tr, bt = (h[k].to(dev, non_blocking=True) for k in ("root", "trans", "betas"))
        mark("h2d")
        p = fitm.FitProblem(smpl, mk, opose, o_betas, shard_fit, groups)
        mark("FitProblem")
        fitm.optim_chamfer(mk, pose, opose, bt, o_betas, rt, tr, None, None, smpl, cfg, problem=p, groups=groups)
        mark("optim_chamfer")
        if groups == 1:
            vi = fitm.compute_nearest_points(mk, pose, bt, rt, tr, smpl, config=cfg, shard=shard_fit, as_one_hot=False,
                                                 problem=p)
        else:
            vi = vidx
        mark("compute_nearest_points")
        fitm.optim_markers(mk, pose, opose, bt, o_betas, rt, tr, vi, None, smpl, cfg, problem=p, groups=groups)
        mark("optim_markers")
        for k, x in (("pose", pose), ("root", rt), ("trans", tr), ("betas", bt)):     # fitted parameters -> pinned host
            h_out[k].copy_(x, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        mark("d2h")
        return h_out

    e2e_once()
    e2e_runs = []
    for _ in range(3):            # three timed calls, the fastest is reported (the host side of a call is exposed to jitter)
        barrier()
        t0 = time.perf_counter()
        e2e_once()
        barrier()
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        ranks = [float(t.item())]
        if world > 1:
            every = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(every, t)
            ranks = [float(x.item()) for x in every]
        e2e_runs.append(ranks)
    e2e_ranks = min(e2e_runs, key=max)
    e2e_s = max(e2e_ranks)
    e2e_value = frames_global * K / e2e_s
    marks = []
    e2e_once(marks)                               # one more call, NOT timed above: where the end-to-end time goes
    barrier()
    e2e_phases = {b[0]: round((b[1] - a[1]) * 1e3, 2) for a, b in zip(marks[:-1], marks[1:])}
    clocks = sampler.stop() if sampler else None

    line = None
    if rank == 0:
        peaks = _peaks()
        Vn = V
        nflat = F * 211 + 10 * groups
        algo = {   # algorithmic bytes (or flops) per launch over this rank's F frames
            "vposed_gemm": ("tensor", 2.0 * F * KC * 3 * Vn),
            "skin_forward": ("hbm", F * (2 * 3 * Vn * 4 + J * 48 + 12)),
            "nn_argmin_tiles": ("hbm", F * (3 * Vn * 4 + M * 12 + M * 8)),
            "nn_argmin_presorted": ("hbm", F * (3 * Vn * 4 + M * 12 + M * 12)),
            "nn_argmin": ("hbm", F * (3 * Vn * 4 + M * 12 + M * 8)),
            "rows_forward": ("hbm", F * (KC * 4 + J * 48 + M * 12)),
            "rows_backward": ("hbm", F * (KC * 4 + J * 48 + M * 12 + 292 * 4 + 224 * 4)),
            "fk_forward": ("hbm", F * (916 + KC * 4 + J * 48 + J * 12)),
            "fk_backward": ("hbm", F * (2 * 916 + 292 * 4 + 224 * 4)),
            "adam_step": ("hbm", nflat * 4 * 7),
        }
        tp = os.path.join(ROOT, "profiles", "traffic.json")       # ncu DRAM bytes per launch, captured at F = 5000
        traffic = json.load(open(tp)) if os.path.exists(tp) else {}
        roofline = None
        if prof:
            dom = max(prof.items(), key=lambda kv: kv[1][0])[0]
            roofline = {"kernel": dom, "share_of_step": prof[dom][0] / prof_ms, "launches": int(round(prof[dom][1])),
                        "ms_per_kernel": kern, "traffic": None,
                        "note": "ms_per_kernel = device time of the kernel per ITERATION of the stage (all its launches summed)"}
            table = {}
            for k, (bound, work) in algo.items():
                per_it = (prof_ch.get(k) or prof_mk.get(k))       # time per iteration over this rank's F frames
                if per_it and per_it[0] > 0:
                    dur = per_it[0] * 1e-3
                    if bound == "hbm":
                        ach, peak, unit = work / dur / 1e9, peaks["hbm"], "GB/s"
                    else:
                        ach, peak, unit = work / dur / 1e12, peaks["bf16"] / 2.0, "TFLOP/s"
                    table[k] = {"bound": bound, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak}
            if dom in table:
                roofline.update(table[dom])
                roofline["peak_source"] = peaks["src"] + (" (TF32 dense = bf16/2; logical flops, the split issues 2-3 MMAs per product)"
                                                          if table[dom]["bound"] == "tensor" else "")
                # ncu captures (dram bytes per launch) exist for F = 5000 (flat keys) and F = 100000 (the default at N = 1)
                by_size = traffic.get("F%d" % F, {})
                roofline["traffic"] = by_size.get(dom, traffic.get(dom) if F == 5000 and dom in FIT_TRAFFIC_KEYS else None)
            roofline["per_kernel"] = table
        cpu = tg = None
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_fit_steps(cap, model, args.ref_frames, temporal, 2, 2, 1)
            cpu = {"value": r["frames"] * 4 / r["seconds"], "unit": UNIT, "cores": r["threads"], "kind": "port",
                   "sample": "first %d frames of the capture, 2 chamfer + 2 marker iterations (closure fwd+bwd + torch Adam) "
                             "after 1 warm-up each, torch threads=%d" % (r["frames"], r["threads"]),
                   "chamfer_s": r["chamfer_s"], "marker_s": r["marker_s"]}
            try:                              # a reported bar, never a reason to lose the bench line
                torch.cuda.empty_cache()
                r = cpu_fit_steps(cap, model, 2048, temporal, 2, 2, 1, device=dev)
                tg = {"value": r["frames"] * 4 / r["seconds"], "unit": UNIT, "frames": r["frames"],
                      "kind": "oracle chamfer + marker closures (fwd+bwd) + torch Adam through stock PyTorch CUDA ops (fp32, TF32 "
                              "off), frame chunks of 256", "chamfer_s": r["chamfer_s"], "marker_s": r["marker_s"]}
            except Exception as e:
                tg = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:160])}
        launches = int(round(sum(v[1] for v in (prof or {}).values())))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": ms_all / K, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": fit_config(args, world),
            "run": {"frames_this_rank": F, "frames_global": frames_global, "cuda_graph": bool(use_graph),
                    "exchange": ("peer-memory kernels (uuo_comm_*)" if shard_fit.comm is not None else
                                 ("torch.distributed" if shard_fit.world > 1 else "none (one shard)"))},
            "stages": stage_rates(), "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d // K, "d2h_bytes_per_step": d2h // K,
                    "seconds": e2e_s, "seconds_per_rank": e2e_ranks,
                    "seconds_of_the_timed_calls": [max(r) for r in e2e_runs], "phase_ms_rank0": e2e_phases,
                    "note": "optim_chamfer -> compute_nearest_points -> optim_markers on pinned host tensors, %d + %d Adam "
                            "iterations, one H2D of this rank's capture and one D2H of its fitted parameters per call" % (n_ch, n_mk)},
            "gpu_launches": launches, "roofline": roofline, "parity": parity, "cpu_baseline": cpu, "torch_gpu_baseline": tg,
        }
    return line


WORKLOAD = "configs[1]: SMPL lbs() forward+backward microbench, %d frames per GPU, synthetic SMPL-shaped model (dense-24 weights)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--frames", type=int, default=0, help="frames (lbs: per GPU, default 4096; fit workloads: their own)")
    ap.add_argument("--workload", default="fit100k", choices=["lbs"] + sorted(FIT_WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-frames", type=int, default=0,
                    help="frames per step of the CPU arm (default: 1024 for the fit workloads, all 4096 for lbs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="fit workloads: skip the oracle self-check of the first closure")
    ap.add_argument("--no-extras", action="store_true", help="skip the short runs of the other workloads (N = 1 only)")
    ap.add_argument("--e2e-sync", action="store_true", help="lbs e2e: one synchronous call per step instead of the double-buffered loop")
    ap.add_argument("--no-graph", action="store_true", help="fit workloads: time eager launches instead of graph replays")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if args.workload == "lbs":
            args.frames = args.frames or 4096
            args.ref_frames = args.ref_frames or args.frames      # every step runs ALL the frames of the config
            run_reference_lbs(args, rank, world)
        else:
            args.ref_frames = args.ref_frames or 1024
            run_reference_fit(args, rank, world)
        return
    args.ref_frames = args.ref_frames or (256 if args.workload == "lbs" else 1024)

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    if args.workload == "lbs":
        line = run_lbs(args, rank, world, local, dev)
    else:
        line = run_fit(args, rank, world, local, dev)
        if rank == 0 and world == 1 and not args.no_extras:
            # the other configs of BASELINE.json, short device-timed runs carried as extra keys of the same line
            import copy
            other = {}
            for name, steps in (("fit5k", 20), ("lbs", 20)):
                a2 = copy.copy(args)
                a2.workload, a2.steps, a2.frames, a2.warmup = name, steps, 0, 3
                try:
                    torch.cuda.empty_cache()
                    other[name] = (run_lbs if name == "lbs" else run_fit)(a2, rank, world, local, dev, light=True)
                except Exception as e:       # never a reason to lose the headline line
                    other[name] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:160])}
            line["other_workloads"] = other
    if rank == 0 and line is not None:
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_lbs(args, rank, world, local, dev, light=False):
    """BASELINE.json configs[1]: lbs() forward + backward over F frames per GPU (weak scaling, no data-path collective)."""
    args.frames = args.frames or 4096
    import uuo_mocap_b200  # noqa: F401
    from uuo_mocap_b200 import synthetic, ops, _lib
    from uuo_mocap_b200._lib import check, ptr, stream_ptr
    from uuo_mocap_b200.smpl import SMPL, _as_struct, lbs_apply

    F = args.frames
    smpl = SMPL(_as_struct(synthetic.make_model("dense")), device=dev)
    mh = smpl.handle
    mb = synthetic.make_lbs_microbench(F, seed=2 + rank)
    T = lambda k: torch.from_numpy(mb[k]).to(dev)
    rot, betas, transl, gv, gj = T("rot"), T("betas"), T("transl"), T("g_vertices"), T("g_joints")
    go, bp = rot[:, :1].contiguous(), rot[:, 1:].contiguous()
    vertices = torch.empty((F, V, 3), device=dev)
    joints = torch.empty((F, J + NE, 3), device=dev)
    ws = mh.workspace(F)
    g_betas, g_go, g_bp, g_tr = torch.empty_like(betas), torch.empty_like(go), torch.empty_like(bp), torch.empty_like(transl)
    L = _lib.lib()
    st = stream_ptr(dev)

    def step_device():
        check(L.uuo_lbs_forward(mh.handle, ptr(betas), F, ptr(go), ptr(bp), 0, None, ptr(transl), F, ptr(vertices),
                                ptr(joints), ptr(ws), st))
        check(L.uuo_lbs_backward(mh.handle, ptr(betas), F, ptr(go), ptr(bp), 0, None, F, ptr(gv), ptr(gj), ptr(ws),
                                 ptr(g_betas), ptr(g_go), ptr(g_bp), None, ptr(g_tr), st))
    # per step (profiles/r1i_ncu_summary.txt): fk_forward, pack_transforms, vposed gemm, skin fwd, extra_joints |
    # skin bwd (d v_posed), extras_vposed, skin bwd (d A), extras_dA, gcoef gemm, fk_backward, copy_gbetas
    launches_per_step = 12

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    ops.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)                          # pass 1: per-kernel events on (feeds roofline.ms_per_launch)
    prof = ops.profile_read()
    ops.profile_enable(False)
    barrier()
    e0.record()
    for _ in range(args.steps):                       # pass 2: the timed K steps, events off
        step_device()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = F * world * args.steps / (ms_max * 1e-3)
    if light:
        if sampler:
            sampler.stop()
        return {"value": value, "unit": UNIT, "ms_per_step": ms_max / args.steps, "steps": args.steps, "frames_per_gpu": F,
                "ms_per_launch": {k: round(v[0] / max(v[1], 1), 4) for k, v in prof.items()},
                "step_hbm_roofline_frac": (168272.0 * F * args.steps / (ms_max * 1e-3) / 1e9) / _peaks()["hbm"],
                "workload": WORKLOAD % F}

    # ---- e2e through the public API with host buffers ----
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    h_go, h_bp, h_betas, h_tr = pin(mb["rot"][:, :1]), pin(mb["rot"][:, 1:]), pin(mb["betas"]), pin(mb["transl"])
    h_out = [torch.empty_like(x).pin_memory() for x in (h_go, h_bp, h_betas, h_tr)]
    h_loss = torch.empty(1).pin_memory()
    h2d = sum(x.numel() * 4 for x in (h_go, h_bp, h_betas, h_tr))
    d2h = h2d + 4

    hosts = (h_go, h_bp, h_betas, h_tr)
    if args.e2e_sync:
        def step_e2e(i):
            d = [x.to(dev, non_blocking=True).requires_grad_(True) for x in hosts]
            o = smpl(betas=d[2], body_pose=d[1], global_orient=d[0], transl=d[3], pose2rot=False)
            # the upstream gradients g_vertices / g_joints are inputs of the workload:
            # they go straight into autograd, no [F,V,3] elementwise "loss" kernels of torch's in the timed region
            torch.autograd.backward([o.vertices, o.joints], [gv, gj])
            for dst, src in zip(h_out, d):
                dst.copy_(src.grad, non_blocking=True)
            h_loss.copy_((o.joints.detach() * gj).sum().reshape(1), non_blocking=True)   # the step's scalar result
            torch.cuda.current_stream().synchronize()   # the caller reads it

        def drain():
            pass
    else:
        # the loop a caller with a stream of host batches writes: the copies of step i+1 (in) and step i-1 (out) run on a
        # copy stream under the kernels of step i; every step still copies its own inputs in and its own results out
        cs = torch.cuda.Stream(device=dev)
        d_in = [[torch.empty(x.shape, device=dev).requires_grad_(True) for x in hosts] for _ in range(2)]
        h_outs = [[torch.empty_like(x).pin_memory() for x in hosts] for _ in range(2)]
        h_losses = [torch.empty(1).pin_memory() for _ in range(2)]
        ev = lambda: [torch.cuda.Event(), torch.cuda.Event()]
        in_ready, comp_done, out_done = ev(), ev(), ev()
        keep = [None, None]

        def issue_h2d(slot):
            with torch.cuda.stream(cs), torch.no_grad():
                for dst, src in zip(d_in[slot], hosts):
                    dst.copy_(src, non_blocking=True)
                in_ready[slot].record(cs)

        def step_e2e(i):
            slot = i & 1
            cur = torch.cuda.current_stream()
            cur.wait_event(in_ready[slot])
            issue_h2d(slot ^ 1)                       # next step's inputs (its buffers were released by out_done below)
            d = d_in[slot]
            for t_ in d:
                t_.grad = None
            o = smpl(betas=d[2], body_pose=d[1], global_orient=d[0], transl=d[3], pose2rot=False)
            torch.autograd.backward([o.vertices, o.joints], [gv, gj])
            loss = (o.joints.detach() * gj).sum().reshape(1)
            comp_done[slot].record(cur)
            keep[slot] = [t_.grad for t_ in d] + [loss]
            with torch.cuda.stream(cs):
                cs.wait_event(comp_done[slot])
                for dst, g_ in zip(h_outs[slot], keep[slot][:4]):
                    dst.copy_(g_, non_blocking=True)
                h_losses[slot].copy_(loss, non_blocking=True)
                out_done[slot].record(cs)
            if i > 0:
                out_done[slot ^ 1].synchronize()      # the caller reads the PREVIOUS step's gradients and scalar

        def drain():
            torch.cuda.current_stream().wait_stream(cs)

        issue_h2d(0)

    n_e2e = 0
    for _ in range(args.warmup):
        step_e2e(n_e2e)
        n_e2e += 1
    drain()
    barrier()
    if not args.e2e_sync:
        issue_h2d(n_e2e & 1)                          # inside the timed region: K steps = K copies in, K copies out
    e0.record()
    for _ in range(args.steps):
        step_e2e(n_e2e)
        n_e2e += 1
    drain()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = F * world * args.steps / (float(t.item()) * 1e-3)
    clocks = sampler.stop() if sampler else None   # sampled across both timed regions

    if rank == 0:
        peaks = _peaks()
        # algorithmic work per launch of each kernel (DESIGN.md "Kernels")
        gemm_flops = 2.0 * F * KC * 3 * V
        algo = {
            "vposed_gemm": ("tensor", gemm_flops),
            "gcoef_gemm": ("tensor", gemm_flops),
            "skin_forward": ("hbm", F * (2 * 3 * V * 4 + J * 48 + 12)),
            "skin_backward_vposed": ("hbm", F * (2 * 3 * V * 4 + J * 48)),
            "skin_backward_A": ("hbm", F * (2 * 3 * V * 4 + J * 48 + 12)),
            "fk_forward": ("hbm", F * (916 + KC * 4 + J * 48 + J * 12)),
            "fk_backward": ("hbm", F * (2 * 916 + 292 * 4 * 4 + 224 * 4 * 8)),
        }
        dom = max(prof.items(), key=lambda kv: kv[1][0])[0] if prof else None
        roofline = None
        kernels = {k: round(v[0] / max(v[1], 1), 4) for k, v in prof.items()}
        if dom in algo:
            bound, work = algo[dom]
            dur = prof[dom][0] / max(prof[dom][1], 1) * 1e-3
            traffic = None
            tp = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tp):
                traffic = json.load(open(tp)).get(dom) if F == 4096 else None   # captured at F = 4096
            if bound == "hbm":
                ach, peak, unit = work / dur / 1e9, peaks["hbm"], "GB/s"
            else:
                ach, peak, unit = work / dur / 1e12, peaks["bf16"] / 2.0, "TFLOP/s"
            roofline = {"kernel": dom, "bound": bound, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak,
                        "traffic": traffic, "peak_source": peaks["src"] + (" (TF32 dense = bf16/2)" if bound == "tensor" else ""),
                        "share_of_step": prof[dom][0] / ms, "ms_per_launch": kernels}
        # whole-step HBM roofline: 168 272 algorithmic bytes per frame-iteration
        step_frac = (168272.0 * F * args.steps / (ms_max * 1e-3) / 1e9) / peaks["hbm"]
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, sec = cpu_oracle_rate(args.ref_frames, 3, threads)
            cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "%d of %d frames, 3 timed reps after 1 warm-up, torch threads=%d" % (args.ref_frames, F, threads)}
        tg = None
        if world == 1 and not args.no_cpu_baseline:
            del vertices, joints, ws
            torch.cuda.empty_cache()
            try:
                rate, tms = torch_gpu_oracle_rate(F, 10, dev, 2 + rank)
                tg = {"value": rate, "unit": UNIT, "ms_per_step": tms, "frames": F,
                      "kind": "oracle lbs() forward+backward through stock PyTorch CUDA ops (fp32, TF32 off), same inputs"}
            except Exception as e:   # a reported bar, never a reason to lose the bench line
                tg = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:120])}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD % F, "frames_per_gpu": F, "parallelism": "frame-range shards x%d" % world,
                       "l2": "inputs larger than L2 (>= 680 MB streamed per step)"},
            "clocks": clocks, "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                                      "mode": "one synchronous call per step" if args.e2e_sync else
                                      "double-buffered: copies of step i+1 (in) / i-1 (out) on a copy stream under the kernels of step i"},
            "gpu_launches": launches_per_step * args.steps, "roofline": roofline,
            "step_hbm_roofline_frac": step_frac, "cpu_baseline": cpu, "torch_gpu_baseline": tg,
            "note": "value: direct C-ABI calls, events off; roofline.ms_per_launch from a separate pass of the same K steps with "
                    "the library's per-kernel CUDA events on; e2e: SMPL.forward + autograd (parameters in, parameter "
                    "gradients + a scalar out; vertices / joints stay on the device, they feed the backward)",
        }
        return line
    return None


if __name__ == "__main__":
    main()
