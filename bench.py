#!/usr/bin/env python
"""bench.py -- frame-iterations/s of the SMPL lbs() forward+backward hot path (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--frames F] [--impl ours|reference]

One "step" = one pass of lbs() forward + backward over a batch of F synthetic frames per GPU
(weak scaling: frames are independent, ranks never communicate on the data path).
  value : device-resident inputs, direct C-ABI calls, CUDA-event timed, max over ranks
  e2e   : the public API (SMPL.forward + autograd backward) with pinned HOST parameters copied in and
          gradients + loss copied out inside the timed region
  roofline : dominant kernel of the step, timed live with CUDA events inside the library
  cpu_baseline : the oracle (bit-identical restatement of the vendored smplx) on the host cores, bounded sample
`--impl reference` times that CPU path alone with the same JSON contract.
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
        ((o.vertices * gv).sum() + (o.joints * gj).sum()).backward()
        dt = time.perf_counter() - t0
        if it > 0:
            times.append(dt)
    med = float(np.median(times))
    return frames / med, med


def run_reference(args, rank, world):
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
                   "note": "reference smplx/PyTorch lbs() fwd+bwd on host cores (oracle port, bit-identical to the "
                           "vendored smplx); each step is a %d-frame sample of the workload" % sample},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%d of %d frames per step, torch threads=%d" % (sample, args.frames, threads)},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
    }
    print(json.dumps(line), flush=True)


WORKLOAD = "configs[1]: SMPL lbs() forward+backward microbench, %d frames per GPU, synthetic SMPL-shaped model (dense-24 weights)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=4096, help="frames per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-frames", type=int, default=256)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

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
    launches_per_step = 10   # fk, transform pack, gemm, skin, extra-joint gather | skin-bwd x2, gemm^T, fk-bwd, beta-grad copy

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
    ms = e0.elapsed_time(e1)
    prof = ops.profile_read()
    ops.profile_enable(False)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = F * world * args.steps / (ms_max * 1e-3)

    # ---- e2e through the public API with host buffers ----
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    h_go, h_bp, h_betas, h_tr = pin(mb["rot"][:, :1]), pin(mb["rot"][:, 1:]), pin(mb["betas"]), pin(mb["transl"])
    h_out = [torch.empty_like(x).pin_memory() for x in (h_go, h_bp, h_betas, h_tr)]
    h_loss = torch.empty(1).pin_memory()
    h2d = sum(x.numel() * 4 for x in (h_go, h_bp, h_betas, h_tr))
    d2h = h2d + 4

    def step_e2e():
        d = [x.to(dev, non_blocking=True).requires_grad_(True) for x in (h_go, h_bp, h_betas, h_tr)]
        o = smpl(betas=d[2], body_pose=d[1], global_orient=d[0], transl=d[3], pose2rot=False)
        loss = (o.vertices * gv).sum() + (o.joints * gj).sum()
        loss.backward()
        for dst, src in zip(h_out, d):
            dst.copy_(src.grad, non_blocking=True)
        h_loss.copy_(loss.detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller reads the loss

    for _ in range(args.warmup):
        step_e2e()
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_e2e()
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
                traffic = json.load(open(tp)).get(dom)
            if bound == "hbm":
                ach, peak, unit = work / dur / 1e9, peaks["hbm"], "GB/s"
            else:
                ach, peak, unit = work / dur / 1e12, peaks["bf16"] / 2.0, "TFLOP/s"
            roofline = {"kernel": dom, "bound": bound, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak,
                        "traffic": traffic, "peak_source": peaks["src"] + (" (TF32 dense = bf16/2)" if bound == "tensor" else ""),
                        "share_of_step": prof[dom][0] / ms, "ms_per_launch": kernels}
        # whole-step HBM roofline: 168 272 algorithmic bytes per frame-iteration
        step_frac = (168272.0 * F * args.steps / (ms * 1e-3) / 1e9) / peaks["hbm"]
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, sec = cpu_oracle_rate(args.ref_frames, 3, threads)
            cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "%d of %d frames, 3 timed reps after 1 warm-up, torch threads=%d" % (args.ref_frames, F, threads)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD % F, "frames_per_gpu": F, "parallelism": "frame-range shards x%d" % world,
                       "l2": "inputs larger than L2 (>= 680 MB streamed per step)"},
            "clocks": clocks, "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches_per_step * args.steps, "roofline": roofline,
            "step_hbm_roofline_frac": step_frac, "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
