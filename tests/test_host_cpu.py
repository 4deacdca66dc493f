"""CPU-side checks: the C-ABI library loads and exports every symbol include/uuo_mocap.h declares (no compute
calls), config loading, frame-range sharding and the halo exchange (gloo, world_size 2)."""
import os
import re
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from uuo_mocap_b200 import _lib, build
    build.build_lib()
    hdr = open(os.path.join(ROOT, "include", "uuo_mocap.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(uuo_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    l = _lib.lib()
    for name in declared:
        assert getattr(l, name) is not None
    assert l.uuo_version() >= 100
    assert l.uuo_last_error() is not None


def test_no_cpu_fallback():
    """The product path refuses CPU devices / tensors instead of silently computing elsewhere."""
    from uuo_mocap_b200 import _lib, ops, synthetic
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    with pytest.raises(_lib.UuoError):
        SMPL(_as_struct(synthetic.make_model("top4")), device="cpu")
    with pytest.raises(_lib.UuoError):
        ops.nn_argmin(torch.zeros(1, 2, 3), torch.zeros(1, 5, 3))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "uuo-mocap_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), fn


def test_load_config_parent_merge(tmp_path):
    from uuo_mocap_b200.fit import load_config, default_config
    (tmp_path / "config").mkdir()
    base = tmp_path / "config" / "base.yaml"
    base.write_text("name: base\nparent:\nnum_root_orient_angles: 4\nstages:\n  chamfer:\n    num_iters: 10\n    losses:\n      full_chamfer: 10.0\n")
    child = tmp_path / "config" / "child.yaml"
    child.write_text("name: child\nparent: %s\nnum_root_orient_angles: 1\nstages:\n  chamfer:\n    num_iters: 3\n" % base)
    cfg = load_config(str(child))
    assert cfg["name"] == "child" and cfg["num_root_orient_angles"] == 1
    assert cfg["stages"]["chamfer"] == {"num_iters": 3, "losses": {"full_chamfer": 10.0}}
    d = default_config()
    assert d["stages"]["chamfer"]["losses"] == {"full_chamfer": 10.0, "reg_pose_body": 1.0, "reg_betas": 1.0}
    assert d["stages"]["marker"]["losses"] == {"marker": 1.0, "reg_pose_body": 0.1, "reg_betas": 1.0}


def test_frame_range_partitions():
    from uuo_mocap_b200.dist import frame_range
    for F in (1, 7, 300, 100000):
        for w in (1, 2, 4, 8):
            r = [frame_range(F, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == F
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


def _temporal_ref(x):
    """optimization.py:368-374 as written."""
    x = x.clone().requires_grad_(True)
    vel = x[2:] - 2 * x[1:-1] - x[:-2]
    loss = torch.nn.functional.mse_loss(vel, torch.zeros_like(vel))
    loss.backward()
    return loss.detach(), x.grad


def _temporal_shard(x, prev, nxt, F_global):
    """What uuo_temporal_accumulate computes for one shard, restated with torch for the gloo test."""
    rows = [t for t in (prev, x, nxt) if t is not None]
    off = 0 if prev is None else prev.shape[0]
    full = torch.cat(rows, 0).clone().requires_grad_(True)
    D = x.shape[1]
    F = x.shape[0]
    e = full[2:] - 2 * full[1:-1] - full[:-2]                     # e_t for t = -off .. ; own terms: t in [0,F)
    t_idx = torch.arange(e.shape[0]) - off
    own = (t_idx >= 0) & (t_idx < F)
    scale = 1.0 / ((F_global - 2) * D)
    # gradient of ALL terms touching own rows, loss of own terms only
    (scale * (e ** 2).sum()).backward()
    return scale * (e[own] ** 2).sum().detach(), full.grad[off:off + F]


def _worker(rank, world, port, F, D, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import uuo_mocap_b200  # noqa: F401
    from uuo_mocap_b200.dist import Shard, exchange_halo
    torch.manual_seed(0)
    x_all = torch.randn(F, D)
    sh = Shard.from_env()
    a, b = sh.range(F)
    x = x_all[a:b].contiguous()
    prev, nxt = exchange_halo(x, 2, sh)
    if rank > 0:
        assert torch.equal(prev, x_all[a - 2:a])
    else:
        assert prev is None
    if rank < world - 1:
        assert torch.equal(nxt, x_all[b:b + 2])
    else:
        assert nxt is None
    # the persistent-buffer form the closures use (fit.chamfer_closure / marker_closure), with a column offset
    from uuo_mocap_b200.dist import exchange_halo_into, halo_buffers
    pbuf, nbuf = halo_buffers(2, D + 3, sh, "cpu")
    exchange_halo_into(x, 2, sh, pbuf, nbuf, col=3)
    assert (pbuf is None) == (rank == 0) and (nbuf is None) == (rank == world - 1)
    if pbuf is not None:
        assert torch.equal(pbuf[:, 3:], prev) and float(pbuf[:, :3].abs().sum()) == 0.0
    if nbuf is not None:
        assert torch.equal(nbuf[:, 3:], nxt)
    assert not sh.graph_safe                     # torch.distributed transport: the iteration cannot be graph-captured
    loss, g = _temporal_shard(x, prev, nxt, F)
    if torch.cuda.is_available():                # with a GPU the shard's terms come from the real kernel
        from uuo_mocap_b200 import _lib
        from uuo_mocap_b200._lib import check, ptr, stream_ptr
        xc = x.cuda()
        pc, nc = (None if t is None else t.cuda().contiguous() for t in (prev, nxt))
        lk, gk = torch.zeros(1, device="cuda"), torch.zeros_like(xc)
        check(_lib.lib().uuo_temporal_accumulate(ptr(xc), ptr(pc), ptr(nc), xc.shape[0], D, 1.0 / ((F - 2) * D), ptr(lk),
                                                 ptr(gk), stream_ptr(xc.device)))
        assert torch.allclose(gk.cpu(), g, rtol=1e-5, atol=1e-7) and abs(float(lk) - float(loss)) < 1e-6
        loss, g = lk.cpu()[0], gk.cpu()
    red = torch.zeros(16)
    red[15] = loss
    red[:3] = float(rank + 1)
    sh.iter_all_reduce_sum_(red)
    cnt = sh.all_reduce_sum(torch.tensor([float(b - a)], dtype=torch.float64))
    ref_loss, ref_g = _temporal_ref(x_all)
    ok = (abs(float(red[15]) - float(ref_loss)) < 1e-6 * abs(float(ref_loss)) and
          torch.allclose(g, ref_g[a:b], rtol=1e-5, atol=1e-7) and int(cnt.item()) == F and
          float(red[0]) == sum(range(1, world + 1)))
    out[rank] = bool(ok)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_halo_exchange_and_sharded_temporal_gloo(world):
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, 11, 5, out), nprocs=world, join=True)
    assert all(out.get(r) for r in range(world)), dict(out)


def test_sub_hierarchies_match_reference_golden():
    """Candidate sub-trees of the part fits == the reference's get_sub_hierachies (utils/smpl_utils.py:106-160),
    same sets AND same order; the golden file was produced by the reference's own function
    (oracle/gen_golden_parts.py)."""
    import json
    import os
    from uuo_mocap_b200.parts import sub_hierarchies
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "sub_hierarchies.json")))
    for k, ref in g["subtrees"].items():
        assert sub_hierarchies(g["parents"], int(k)) == ref, k
    assert sub_hierarchies(g["parents"], 99) == [sorted(range(24))]


def test_stageii_writer_roundtrip(tmp_path):
    """test/test.py:115-143: keys, shapes, per-stage file names; axis-angle vs scipy (angle in [0, pi])."""
    from scipy.spatial.transform import Rotation
    from uuo_mocap_b200 import stageii
    F = 7
    rot = Rotation.random(F * 24, random_state=11)
    R = torch.from_numpy(rot.as_matrix().astype(np.float32)).reshape(F, 24, 3, 3)
    R[0, 0] = torch.eye(3)                                        # identity: the small-angle branch
    R[1, 0] = torch.tensor([[-1.0, 0, 0], [0, -1.0, 0], [0, 0, 1.0]])   # a half turn
    aa = stageii.matrix_to_axis_angle(R)
    back = Rotation.from_rotvec(aa.reshape(-1, 3).numpy().astype(np.float64)).as_matrix().reshape(F, 24, 3, 3)
    assert np.abs(back - R.numpy()).max() < 2e-6
    assert float(aa.norm(dim=-1).max()) <= np.pi + 1e-5 and float(aa[0, 0].abs().max()) == 0.0
    ref = rot.as_rotvec().reshape(F, 24, 3)
    ok = np.ones((F, 24), bool)
    ok[0, 0] = ok[1, 0] = False
    assert np.abs(aa.numpy()[ok] - ref[ok]).max() < 2e-5
    name = str(tmp_path / "take_stageii.npz")
    mk = np.random.default_rng(0).normal(size=(F, 41, 3)).astype(np.float32)
    stages = {"chamfer": {"trans": np.zeros((F, 3), np.float32), "betas": np.ones((1, 10), np.float32),
                          "root_orient": R[:, :1].numpy(), "pose_body": R[:, 1:].numpy()}}
    out = stageii.save_stageii(name, torch.arange(10.0).reshape(1, 10), torch.ones(F, 3), R[:, :1], R[:, 1:], 120, mk,
                               stages=stages)
    z = stageii.load_stageii(name)
    assert sorted(z) == ["betas", "gender", "mocap_frame_rate", "mocap_markers", "poses", "trans"]
    assert z["betas"].shape == (10,) and z["trans"].shape == (F, 3) and z["poses"].shape == (F, 72)
    assert str(z["gender"]) == "neutral" and int(z["mocap_frame_rate"]) == 120 and np.array_equal(z["mocap_markers"], mk)
    assert np.array_equal(z["poses"], out["poses"])
    zs = stageii.load_stageii(str(tmp_path / "take_stageii.chamfer.npz"))
    assert zs["betas"].shape == (1, 10) and np.array_equal(zs["poses"], z["poses"]) and float(zs["trans"].sum()) == 0.0


@pytest.mark.parametrize("img_freq,mocap_freq", [(30.0, 120.0), (30.0, 100.0), (60.0, 50.0), (30.0, 30.0)])
def test_resample_prior_matches_reference_loop(img_freq, mocap_freq):
    """multimodal.py:144-182 restated frame by frame, rotations through scipy's Slerp (shortest great arc, what
    roma.utils.unitquat_slerp traces whatever the component order): frame count, the repeated tail, linear terms."""
    from scipy.spatial.transform import Rotation, Slerp
    from uuo_mocap_b200 import stageii
    F, J = 9, 24
    rng = np.random.default_rng(5)
    rot = Rotation.random(F * J, random_state=3).as_matrix().reshape(F, J, 3, 3)
    rot[3] = rot[2]                                               # identical neighbours: the small-angle branch
    R = torch.from_numpy(rot.astype(np.float32))
    trans = torch.from_numpy(rng.normal(size=(F, 3)).astype(np.float32))
    foot = torch.from_numpy(rng.random(size=(F, 2)).astype(np.float32))
    t, ro, pb, fc = stageii.resample_prior(trans, R[:, :1], R[:, 1:], foot, img_freq, mocap_freq)
    if img_freq == mocap_freq:
        assert t is trans and fc is foot and ro.shape[0] == F
        return
    n_new = round(F * (mocap_freq / img_freq))
    assert t.shape == (n_new, 3) and ro.shape == (n_new, 1, 3, 3) and pb.shape == (n_new, J - 1, 3, 3) and fc.shape == (n_new, 2)
    got = torch.cat((ro, pb), dim=1).numpy()
    for i in range(n_new):
        frame = int(i * (img_freq / mocap_freq))
        alpha = i * (img_freq / mocap_freq) - frame
        if frame + 1 < F:
            assert np.abs(t[i].numpy() - (trans[frame + 1].numpy() * alpha + trans[frame].numpy() * (1 - alpha))).max() < 1e-6
            assert np.abs(fc[i].numpy() - (foot[frame + 1].numpy() * alpha + foot[frame].numpy() * (1 - alpha))).max() < 1e-6
            for j in range(J):
                key = Rotation.from_matrix(np.stack((rot[frame, j], rot[frame + 1, j])))
                ref = Slerp([0.0, 1.0], key)([alpha]).as_matrix()[0]
                assert np.abs(got[i, j] - ref).max() < 5e-6, (i, j)
        else:
            assert torch.equal(t[i], trans[frame]) and torch.equal(fc[i], foot[frame])
            assert np.array_equal(got[i], rot[frame].astype(np.float32))


def test_lbfgs_is_refused_across_shards():
    """torch's L-BFGS decides from the rank-local gradient: not defined across frame-range shards (fit._optimise)."""
    from uuo_mocap_b200.dist import Shard
    from uuo_mocap_b200.fit import _optimise
    with pytest.raises(NotImplementedError):
        _optimise(None, lambda: None, {"num_iters": 5}, {"name": "lbfgs"}, 0.1, None, Shard(0, 2))
    with pytest.raises(NotImplementedError):       # no name = the reference's optimiser = L-BFGS
        _optimise(None, lambda: None, {"num_iters": 5}, {}, 0.1, None, Shard(1, 2))


def test_bench_reference_arm_line(tmp_path):
    """`bench.py --impl reference` (the driver's CPU arm) prints ONE JSON line with the contract's keys, on the same
    `config` object as the GPU arm (bench.fit_config), from a bounded frame sample.  Tiny sample here: seconds."""
    import json
    import subprocess
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                          "--ref-frames", "12"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "frame_iters_per_sec" and d["unit"] == "frame-iters/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and "12-frame" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    sys.path.insert(0, ROOT)
    import argparse
    import importlib
    saved = (os.dup(1),)
    try:
        bench = importlib.import_module("bench")       # (bench redirects fd 1 to stderr at import: restore it afterwards)
    finally:
        os.dup2(saved[0], 1)
        os.close(saved[0])
    want = bench.fit_config(argparse.Namespace(workload="fit100k", frames=0), 1)
    assert d["config"] == want and "configs[4]" in d["config"]["workload"]


def test_rot_z_apply_equals_matmul():
    """fit.rot_z_apply (elementwise Rz @ m, used for the yaw write-back and the stacked hypotheses) == rot_z(angle) @ m."""
    from uuo_mocap_b200 import fit
    g = torch.Generator().manual_seed(3)
    z, R = torch.randn(7, 1, 1, generator=g), torch.randn(7, 1, 3, 3, generator=g)
    assert torch.allclose(fit.rot_z_apply(z, R), fit.rot_z(z) @ R, atol=1e-6)
    ang = torch.randn(3, generator=g)
    a = fit.rot_z(ang.view(3, 1, 1, 1)) @ R.unsqueeze(0)
    b = fit.rot_z_apply(ang.view(3, 1, 1, 1), R.unsqueeze(0))
    assert a.shape == b.shape and torch.allclose(a, b, atol=1e-6)


def test_mean_bound_covers_the_in_order_fp32_sum():
    """ops.mean_bound_from_sums: the fp32 bound built from an fp64 sum is >= the in-order fp32 sum of the same
    non-negative terms (what uuo_mean_dist_accumulate_bounded compares its running sums with)."""
    from uuo_mocap_b200 import ops
    rng = np.random.default_rng(0)
    for n in (1, 7, 1000, 200000):
        x = np.abs(rng.normal(0.3, 0.2, size=(5, n))).astype(np.float32)
        s32 = np.zeros(5, dtype=np.float32)
        for k in range(n if n <= 1000 else 0):
            s32 = s32 + x[:, k]
        if n > 1000:
            s32 = np.cumsum(x, axis=1, dtype=np.float32)[:, -1]          # sequential fp32 accumulation
        s64 = torch.from_numpy(x.astype(np.float64).sum(1))
        b = ops.mean_bound_from_sums(s64, n).numpy()
        rel = 2.0 ** -10 + n * 2.0 ** -23                                 # the margin mean_bound_from_sums adds
        assert b.dtype == np.float32 and (b >= s32).all() and (b <= s64.numpy() * (1 + 1.01 * rel) + 1e-6).all()
