"""Parity of the CUDA SMPL layer (through the C-ABI) against the oracle and the smplx golden vectors.

Tolerances (BASELINE.json north_star): vertices / joints 1e-5 relative, gradients 1e-4 relative
(relative to the largest magnitude of the reference tensor).
"""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL_OUT = 1e-5
TOL_GRAD = 1e-4


@pytest.fixture(scope="module")
def layers(get_model):
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    from oracle.smpl_oracle import OracleSMPL
    out = {}
    for kind in ("dense", "top4"):
        m = get_model(kind)
        out[kind] = (SMPL(_as_struct(m), device="cuda:0"), OracleSMPL.from_model(m))
    return out


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda().requires_grad_(True)


@pytest.mark.parametrize("kind", ["dense", "top4"])
@pytest.mark.parametrize("tag,pose2rot", [("rot", False), ("aa", True)])
def test_against_smplx_golden(layers, golden, kind, tag, pose2rot):
    """tests/golden/smplx_*.npz are outputs of the reference's own vendored smplx."""
    smpl, _ = layers[kind]
    g = golden(f"smplx_{kind}.npz")
    go, bp = _cuda(g[f"{tag}_global_orient"]), _cuda(g[f"{tag}_body_pose"])
    betas, transl = _cuda(g[f"{tag}_betas"]), _cuda(g[f"{tag}_transl"])
    o = smpl(betas=betas, body_pose=bp, global_orient=go, transl=transl, pose2rot=pose2rot)
    assert rel_err(o.vertices.detach().cpu(), g[f"{tag}_vertices"]) < TOL_OUT
    assert rel_err(o.joints.detach().cpu(), g[f"{tag}_joints"]) < TOL_OUT
    gv, gj = torch.from_numpy(g["g_vertices"]).cuda(), torch.from_numpy(g["g_joints"]).cuda()
    ((o.vertices * gv).sum() + (o.joints * gj).sum()).backward()
    for t, key in ((go, "g_global_orient"), (bp, "g_body_pose"), (betas, "g_betas"), (transl, "g_transl")):
        ref = g[f"{tag}_{key}"]
        assert rel_err(t.grad.cpu().numpy().reshape(ref.shape), ref) < TOL_GRAD, key


def test_shared_betas_golden(layers, golden):
    from uuo_mocap_b200 import synthetic
    smpl, _ = layers["dense"]
    g = golden("smplx_dense.npz")
    mb = synthetic.make_lbs_microbench(g["rot_betas"].shape[0], seed=11)
    rot = torch.from_numpy(mb["rot"]).cuda()
    o = smpl(betas=torch.from_numpy(g["rot_betas"][:1]).cuda(), body_pose=rot[:, 1:], global_orient=rot[:, :1],
             transl=torch.from_numpy(g["rot_transl"]).cuda(), pose2rot=False)
    assert rel_err(o.vertices.cpu(), g["shared_vertices"]) < TOL_OUT
    assert rel_err(o.joints.cpu(), g["shared_joints"]) < TOL_OUT


@pytest.mark.parametrize("kind", ["dense", "top4"])
@pytest.mark.parametrize("F", [1, 37, 130])
@pytest.mark.parametrize("mode", ["rot", "aa", "raw6d_yaw"])
@pytest.mark.parametrize("shared_betas", [False, True])
def test_against_oracle(layers, kind, F, mode, shared_betas):
    """Seeded inputs, ragged frame counts (not multiples of any tile), every pose mode."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.smpl import lbs_apply, POSE_RAW6D
    from oracle import p3d_oracle as p3d
    smpl, orc = layers[kind]
    mb = synthetic.make_lbs_microbench(F, seed=100 + F)
    rng = np.random.default_rng(F)
    betas_np = mb["betas"][:1] if shared_betas else mb["betas"]
    yaw_np = rng.normal(0, 0.5, size=(F, 1, 1)).astype(np.float32)
    gv, gj = torch.from_numpy(mb["g_vertices"]), torch.from_numpy(mb["g_joints"])

    def run(dev):
        T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev).requires_grad_(True)
        betas, transl = T(betas_np), T(mb["transl"])
        leaves = [betas, transl]
        if mode == "aa":
            go, bp = T(mb["pose_aa"][:, :1]), T(mb["pose_aa"][:, 1:])
            leaves += [go, bp]
            if dev == "cpu":
                o = orc.forward(betas, bp, go, transl, pose2rot=True)
                v, j = o.vertices, o.joints
            else:
                o = smpl(betas=betas, body_pose=bp.reshape(F, -1), global_orient=go.reshape(F, -1), transl=transl,
                         pose2rot=True)
                v, j = o.vertices, o.joints
        elif mode == "rot":
            go, bp = T(mb["rot"][:, :1]), T(mb["rot"][:, 1:])
            leaves += [go, bp]
            if dev == "cpu":
                o = orc.forward(betas, bp, go, transl, pose2rot=False)
            else:
                o = smpl(betas=betas, body_pose=bp, global_orient=go, transl=transl, pose2rot=False)
            v, j = o.vertices, o.joints
        else:
            noise = np.random.default_rng(7).normal(0, 0.05, size=mb["rot"].shape).astype(np.float32)
            raw = mb["rot"] + noise                       # free 3x3 parameters, not orthonormal
            go, bp = T(raw[:, :1]), T(raw[:, 1:])
            yaw = T(yaw_np)
            leaves += [go, bp, yaw]
            if dev == "cpu":                              # optimization.py:192-202
                z_root = p3d.compute_root_orient_z(yaw) @ go
                o = orc.forward(betas, p3d.normalize_rot(bp), p3d.normalize_rot(z_root), transl, pose2rot=False)
                v, j = o.vertices, o.joints
            else:
                v, j = lbs_apply(smpl.handle, betas, go, bp, transl, yaw, POSE_RAW6D)
        ((v * gv.to(dev)).sum() + (j * gj.to(dev)).sum()).backward()
        return v.detach().cpu().numpy(), j.detach().cpu().numpy(), [t.grad.cpu().numpy() for t in leaves]

    v_ref, j_ref, g_ref = run("cpu")
    v, j, g = run("cuda:0")
    assert rel_err(v, v_ref) < TOL_OUT
    assert rel_err(j, j_ref) < TOL_OUT
    for a, b in zip(g, g_ref):
        assert a.shape == b.shape
        assert rel_err(a, b) < TOL_GRAD


def test_partial_grads_and_errors(layers):
    """Only joints used (g_vertices NULL), transl absent; error behaviour of SmplInference."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.smpl import SmplInference
    smpl, orc = layers["top4"]
    F = 5
    mb = synthetic.make_lbs_microbench(F, seed=5)
    res = []
    for dev in ("cpu", "cuda:0"):
        rot = torch.from_numpy(mb["rot"]).to(dev).requires_grad_(True)
        betas = torch.from_numpy(mb["betas"]).to(dev).requires_grad_(True)
        if dev == "cpu":
            j = orc.forward(betas, rot[:, 1:], rot[:, :1], None, pose2rot=False).joints
        else:
            j = smpl(betas=betas, body_pose=rot[:, 1:], global_orient=rot[:, :1], pose2rot=False).joints
        (j ** 2).sum().backward()
        res.append((j.detach().cpu().numpy(), rot.grad.cpu().numpy(), betas.grad.cpu().numpy()))
    for a, b in zip(res[1], res[0]):
        assert rel_err(a, b) < TOL_GRAD
    inf = SmplInference("cuda:0", model=smpl)
    with pytest.raises(ValueError, match="Betas array must have 10 beta values"):
        inf(poses=torch.zeros(1, 23, 3, 3).cuda(), betas=torch.zeros(1, 9).cuda(),
            root_orient=torch.zeros(1, 1, 3, 3).cuda(), trans=torch.zeros(1, 3).cuda())
    assert inf.get_lbs_weights().shape == (6890, 24)
    assert inf.smpl.parents.shape == (24,)


def test_unaligned_caller_tensors(layers):
    """Caller tensors are only 4-byte aligned in general (views into larger buffers): the vectorised / async-copy
    paths must give bit-identical results for aligned and shifted storage."""
    from uuo_mocap_b200 import synthetic
    smpl, _ = layers["dense"]
    F = 37
    mb = synthetic.make_lbs_microbench(F, seed=21)

    def shifted(a, k):
        flat = torch.zeros(a.size + 8, device="cuda")
        flat[k:k + a.size] = torch.from_numpy(a.reshape(-1)).cuda()
        return flat[k:k + a.size].view(*a.shape)

    outs = []
    for k in (0, 1, 3):
        rot = shifted(mb["rot"], k).requires_grad_(True)
        betas = shifted(mb["betas"], k).requires_grad_(True)
        transl = shifted(mb["transl"], k).requires_grad_(True)
        o = smpl(betas=betas, body_pose=rot[:, 1:], global_orient=rot[:, :1], transl=transl, pose2rot=False)
        gv, gj = shifted(mb["g_vertices"], k), shifted(mb["g_joints"], k)
        grads = torch.autograd.grad([o.vertices, o.joints], [rot, betas, transl], grad_outputs=[gv, gj])
        outs.append([t.detach().cpu() for t in (o.vertices, o.joints) + tuple(grads)])
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            assert torch.equal(a, b)


def test_betas_rows_per_group(layers):
    """betas_rows = G (a divisor of F): G stacked hypothesis groups of F/G consecutive frames, one betas row each (the
    batched MHT loop, multimodal.py:462-606) == the same rows expanded to one per frame; d betas summed per group."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.smpl import lbs_apply
    smpl, _ = layers["top4"]
    F, G = 12, 3
    mb = synthetic.make_lbs_microbench(F, seed=17)
    rot = torch.from_numpy(mb["rot"]).cuda()
    tr = torch.from_numpy(mb["transl"]).cuda()
    gv, gj = torch.from_numpy(mb["g_vertices"]).cuda(), torch.from_numpy(mb["g_joints"]).cuda()
    bG = torch.from_numpy(mb["betas"][:G]).cuda().requires_grad_(True)
    bF = bG.detach().repeat_interleave(F // G, dim=0).requires_grad_(True)
    outs = []
    for b in (bG, bF):
        v, j = lbs_apply(smpl.handle, b, rot[:, :1].contiguous(), rot[:, 1:].contiguous(), tr)
        torch.autograd.backward([v, j], [gv, gj])
        outs.append((v.detach(), j.detach(), b.grad))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    want = outs[1][2].view(G, F // G, -1).sum(1)
    assert rel_err(outs[0][2].cpu().numpy(), want.cpu().numpy()) < 1e-6
    with pytest.raises(ValueError):
        lbs_apply(smpl.handle, torch.zeros(5, 10, device="cuda"), rot[:, :1].contiguous(), rot[:, 1:].contiguous(), tr)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices in one process")
def test_two_devices_in_one_process(get_model):
    """Launch configuration (SM count, > 48 KB dynamic-smem opt-in) is cached per DEVICE (csrc/devcfg.cu): a model on
    cuda:1 created after cuda:0 was used in the same process gives the same results."""
    from uuo_mocap_b200 import ops, synthetic
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    m = get_model("top4")
    mb = synthetic.make_lbs_microbench(6, seed=4)
    cap = synthetic.make_capture(m, 6, num_markers=41, seed=9)
    outs = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        smpl = SMPL(_as_struct(m), device=dev)
        rot = torch.from_numpy(mb["rot"]).to(dev).requires_grad_(True)
        o = smpl(betas=torch.from_numpy(mb["betas"]).to(dev), body_pose=rot[:, 1:], global_orient=rot[:, :1],
                 transl=torch.from_numpy(mb["transl"]).to(dev), pose2rot=False)
        (o.vertices * torch.from_numpy(mb["g_vertices"]).to(dev)).sum().backward()
        d, i = ops.nn_argmin(torch.from_numpy(cap.markers).to(dev), o.vertices.detach().contiguous(), smpl.handle)
        outs.append((o.vertices.detach().cpu(), rot.grad.cpu(), d.cpu(), i.cpu()))
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            assert torch.equal(a, b)
