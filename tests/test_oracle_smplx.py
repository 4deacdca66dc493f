"""The oracle's SMPL restatement against the reference's own smplx (pinning the oracle).

* golden: tests/golden/smplx_*.npz were produced by the unmodified vendored smplx
  (oracle/gen_golden.py) -> the oracle must reproduce them bit for bit on the CPU.
* live: when /root/reference is present (build container) the vendored smplx is executed
  again on fresh random inputs and compared bit for bit.
"""
import os
import sys

import numpy as np
import pytest
import torch

from oracle.smpl_oracle import OracleSMPL

REF_SMPLX = "/root/reference/third_party/smplx"


def _run_oracle(smpl, g, tag, pose2rot):
    F = g[f"{tag}_betas"].shape[0]
    t = lambda k: torch.from_numpy(g[f"{tag}_{k}"]).clone().requires_grad_(True)
    go, bp, betas, transl = t("global_orient"), t("body_pose"), t("betas"), t("transl")
    if pose2rot:
        o = smpl.forward(betas, bp.view(F, 23, 3), go.view(F, 1, 3), transl, pose2rot=True)
    else:
        o = smpl.forward(betas, bp, go, transl, pose2rot=False)
    loss = (o.vertices * torch.from_numpy(g["g_vertices"])).sum() + (o.joints * torch.from_numpy(g["g_joints"])).sum()
    loss.backward()
    return o, go, bp, betas, transl


@pytest.mark.parametrize("kind", ["dense", "top4"])
@pytest.mark.parametrize("tag,pose2rot", [("rot", False), ("aa", True)])
def test_oracle_reproduces_smplx_golden(get_model, golden, kind, tag, pose2rot):
    model = get_model(kind)
    g = golden(f"smplx_{kind}.npz")
    assert abs(float(g["model_checksum"]) - model.checksum()) < 1e-6 * model.checksum(), \
        "synthetic model differs from the one the golden vectors were made with"
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    smpl = OracleSMPL.from_model(model)
    o, go, bp, betas, transl = _run_oracle(smpl, g, tag, pose2rot)
    # same torch ops, same machine class: 1e-6 relative covers BLAS thread-count differences
    for got, key in ((o.vertices, "vertices"), (o.joints, "joints")):
        ref = g[f"{tag}_{key}"]
        assert np.abs(got.detach().numpy() - ref).max() <= 1e-6 * np.abs(ref).max()
    for got, key in ((go.grad, "g_global_orient"), (bp.grad, "g_body_pose"), (betas.grad, "g_betas"),
                     (transl.grad, "g_transl")):
        ref = g[f"{tag}_{key}"]
        assert np.abs(got.numpy().reshape(ref.shape) - ref).max() <= 2e-5 * np.abs(ref).max(), key


def test_oracle_shared_betas_golden(get_model, golden):
    model = get_model("dense")
    g = golden("smplx_dense.npz")
    smpl = OracleSMPL.from_model(model)
    T = lambda k: torch.from_numpy(g[k])
    # rot_* inputs of frame 0 were perturbed; the shared-betas vectors used the clean rotations
    from uuo_mocap_b200 import synthetic
    mb = synthetic.make_lbs_microbench(g["rot_betas"].shape[0], seed=11)
    rot = torch.from_numpy(mb["rot"])
    o = smpl.forward(T("rot_betas")[:1], rot[:, 1:], rot[:, :1], T("rot_transl"), pose2rot=False)
    assert np.abs(o.vertices.numpy() - g["shared_vertices"]).max() <= 1e-6 * np.abs(g["shared_vertices"]).max()
    assert np.abs(o.joints.numpy() - g["shared_joints"]).max() <= 1e-6 * np.abs(g["shared_joints"]).max()


@pytest.mark.skipif(not os.path.isdir(REF_SMPLX), reason="reference tree not present (GPU box)")
@pytest.mark.parametrize("pose2rot", [False, True])
def test_oracle_matches_live_smplx(get_model, pose2rot):
    from oracle.gen_golden import reference_smpl
    from uuo_mocap_b200 import synthetic
    model = get_model("dense")
    ref = reference_smpl(model)
    smpl = OracleSMPL.from_model(model)
    mb = synthetic.make_lbs_microbench(5, seed=77)
    betas, transl = torch.from_numpy(mb["betas"]), torch.from_numpy(mb["transl"])
    if pose2rot:
        aa = torch.from_numpy(mb["pose_aa"])
        r = ref(betas=betas, body_pose=aa[:, 1:].reshape(5, 69), global_orient=aa[:, 0], transl=transl, pose2rot=True)
        o = smpl.forward(betas, aa[:, 1:], aa[:, :1], transl, pose2rot=True)
    else:
        rot = torch.from_numpy(mb["rot"])
        r = ref(betas=betas, body_pose=rot[:, 1:], global_orient=rot[:, :1], transl=transl, pose2rot=False)
        o = smpl.forward(betas, rot[:, 1:], rot[:, :1], transl, pose2rot=False)
    assert torch.equal(r.vertices, o.vertices)
    assert torch.equal(r.joints, o.joints)
