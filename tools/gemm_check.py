"""GPU check of the tcgen05 3xTF32 GEMMs against the exact-fp32 SIMT kernels (UUO_GEMM_SIMT=1) through the
SMPL layer: same inputs, compare vertices and every gradient; then per-kernel timings.

    UUO_EXTRA_NVCC_FLAGS=-DUUO_DEBUG_SIMT python -m uuo_mocap_b200.build --force     # the release build has no SIMT path
    python tools/gemm_check.py [F ...]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import uuo_mocap_b200  # noqa: F401
from uuo_mocap_b200 import synthetic
from uuo_mocap_b200.smpl import SMPL, _as_struct


def run(smpl, mb, simt):
    os.environ["UUO_GEMM_SIMT"] = "1" if simt else "0"
    os.environ["UUO_SKIN_SIMT"] = "1" if simt else "0"
    rot = torch.from_numpy(mb["rot"]).cuda().requires_grad_(True)
    betas = torch.from_numpy(mb["betas"]).cuda().requires_grad_(True)
    transl = torch.from_numpy(mb["transl"]).cuda().requires_grad_(True)
    o = smpl(betas=betas, body_pose=rot[:, 1:], global_orient=rot[:, :1], transl=transl, pose2rot=False)
    gv = torch.from_numpy(mb["g_vertices"]).cuda()
    gj = torch.from_numpy(mb["g_joints"]).cuda()
    ((o.vertices * gv).sum() + (o.joints * gj).sum()).backward()
    torch.cuda.synchronize()
    return [t.detach().cpu().numpy() for t in (o.vertices, o.joints, rot.grad, betas.grad, transl.grad)]


def main():
    Fs = [int(a) for a in sys.argv[1:]] or [8, 300, 4096]
    model = synthetic.make_model("dense")
    smpl = SMPL(_as_struct(model), device="cuda:0")
    ok = True
    for F in Fs:
        mb = synthetic.make_lbs_microbench(F, seed=2)
        ref = run(smpl, mb, True)
        t0 = time.time()
        out = run(smpl, mb, False)
        names = ["vertices", "joints", "g_rot", "g_betas", "g_transl"]
        tol = [1e-5, 1e-5, 1e-4, 1e-4, 1e-4]
        for n, a, b, t in zip(names, out, ref, tol):
            e = float(np.abs(a - b).max() / np.abs(b).max())
            bad = not (e < t)
            ok &= not bad
            print(f"F={F:6d} {n:10s} rel err tensor-core vs simt = {e:.3e} {'FAIL' if bad else 'ok'}", flush=True)
    print("GEMM_CHECK", "PASS" if ok else "FAIL")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
