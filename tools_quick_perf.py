"""Scratch timing of the LBS fwd+bwd at F=4096 with per-kernel events (not the bench)."""
import sys, time, torch
import uuo_mocap_b200
from uuo_mocap_b200 import synthetic, ops, _lib
from uuo_mocap_b200.smpl import SMPL, _as_struct, lbs_apply
F = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
smpl = SMPL(_as_struct(synthetic.make_model("dense")), device="cuda:0")
mb = synthetic.make_lbs_microbench(F, seed=2)
T = lambda k: torch.from_numpy(mb[k]).cuda()
rot, betas, transl, gv, gj = T("rot"), T("betas"), T("transl"), T("g_vertices"), T("g_joints")
go, bp = rot[:, :1].contiguous().requires_grad_(True), rot[:, 1:].contiguous().requires_grad_(True)
betas.requires_grad_(True); transl.requires_grad_(True)
def step():
    v, j = lbs_apply(smpl.handle, betas, go, bp, transl)
    torch.autograd.backward([v, j], [gv, gj])
for _ in range(3): step()
torch.cuda.synchronize()
ops.profile_enable(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
N = 10
for _ in range(N): step()
e1.record(); torch.cuda.synchronize()
print("ms/step", e0.elapsed_time(e1) / N, "frame-iters/s", F * N / (e0.elapsed_time(e1) * 1e-3))
for k, (ms, c) in ops.profile_read().items(): print("%-24s %8.3f ms/launch  x%d" % (k, ms / max(c, 1), c))
ops.profile_enable(False)
