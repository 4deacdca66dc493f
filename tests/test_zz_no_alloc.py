"""uuo_lbs_forward_static (include/uuo_mocap.h): the lbs() forward that allocates nothing -- its workspace is a file-scope
device array of UUO_STATIC_WORKSPACE_BYTES.  Called through the C-ABI and compared with uuo_lbs_forward on a
caller-allocated workspace (the same kernels: the outputs must agree far inside the 1e-5 contract of test_lbs_gpu.py).
Collected last on purpose."""
import numpy as np
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL_OUT = 1e-5               # vertices / joints, relative (test_lbs_gpu.py)
STATIC_BYTES = 8 << 20       # UUO_STATIC_WORKSPACE_BYTES


def _inputs(model, F, seed, shared_betas):
    from uuo_mocap_b200 import synthetic
    rng = np.random.default_rng(seed)
    aa = rng.normal(0, 0.4, size=(F, 24, 3))
    rot = synthetic.axis_angle_to_matrix_np(aa).astype(np.float32)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a.astype(np.float32))).cuda()
    return dict(betas=T(rng.normal(0, 1, size=(1 if shared_betas else F, 10))), go=T(rot[:, :1]), bp=T(rot[:, 1:]),
                transl=T(rng.normal(0, 1, size=(F, 3))))


def _forward(mh, x, F, static):
    from uuo_mocap_b200 import _lib
    from uuo_mocap_b200._lib import ptr, stream_ptr
    from uuo_mocap_b200.smpl import POSE_ROTMAT
    dev = mh.device
    vertices = torch.empty((F, mh.V, 3), dtype=torch.float32, device=dev)
    joints = torch.empty((F, mh.J + mh.NE, 3), dtype=torch.float32, device=dev)
    L = _lib.lib()
    with torch.cuda.device(dev):
        if static:
            before = torch.cuda.memory_allocated(dev)
            rc = L.uuo_lbs_forward_static(mh.handle, ptr(x["betas"]), x["betas"].shape[0], ptr(x["go"]), ptr(x["bp"]),
                                          POSE_ROTMAT, None, ptr(x["transl"]), F, ptr(vertices), ptr(joints),
                                          stream_ptr(dev))
            assert torch.cuda.memory_allocated(dev) == before
        else:
            ws = mh.workspace(F)
            rc = L.uuo_lbs_forward(mh.handle, ptr(x["betas"]), x["betas"].shape[0], ptr(x["go"]), ptr(x["bp"]),
                                   POSE_ROTMAT, None, ptr(x["transl"]), F, ptr(vertices), ptr(joints), ptr(ws),
                                   stream_ptr(dev))
    assert rc == 0, L.uuo_last_error().decode()
    torch.cuda.synchronize(dev)
    return vertices, joints


def test_lbs_forward_static_matches_lbs_forward(get_model):
    """Two calls in a row on the shared static workspace (different frame counts, betas layouts and inputs, so the second
    call runs over the first one's leftovers), each against uuo_lbs_forward on a fresh workspace."""
    from uuo_mocap_b200.smpl import SMPL, _as_struct
    smpl = SMPL(_as_struct(get_model("top4")), device="cuda:0")
    mh = smpl.handle
    for F, seed, shared in ((33, 1, False), (7, 2, True), (33, 3, True)):
        x = _inputs(get_model("top4"), F, seed, shared)
        v_ref, j_ref = _forward(mh, x, F, static=False)
        v, j = _forward(mh, x, F, static=True)
        assert torch.isfinite(v).all() and torch.isfinite(j).all()
        assert rel_err(v.cpu(), v_ref.cpu()) <= TOL_OUT and rel_err(j.cpu(), j_ref.cpu()) <= TOL_OUT
        assert torch.equal(v, v_ref) and torch.equal(j, j_ref)        # the same kernels on the same inputs


def test_lbs_forward_static_rejects_what_does_not_fit(get_model):
    from uuo_mocap_b200 import _lib
    from uuo_mocap_b200._lib import ptr, stream_ptr
    from uuo_mocap_b200.smpl import POSE_ROTMAT, SMPL, _as_struct
    smpl = SMPL(_as_struct(get_model("top4")), device="cuda:0")
    mh = smpl.handle
    L = _lib.lib()
    F = 1
    while L.uuo_lbs_workspace_bytes(mh.handle, F) <= STATIC_BYTES:
        F += 1
    assert F > 32                                   # the stated limit: about 35 frames of an SMPL-sized model
    x = _inputs(get_model("top4"), F, 4, True)
    out = torch.empty((F, mh.V, 3), device="cuda:0")
    args = lambda n: (mh.handle, ptr(x["betas"]), 1, ptr(x["go"]), ptr(x["bp"]), POSE_ROTMAT, None, ptr(x["transl"]), n,
                      ptr(out), None, stream_ptr(mh.device))
    with torch.cuda.device(mh.device):
        assert L.uuo_lbs_forward_static(*args(F)) == -1
        msg = L.uuo_last_error().decode()
        assert "static workspace" in msg and str(STATIC_BYTES) in msg and ("F = %d" % F) in msg
        assert L.uuo_lbs_forward_static(*args(F - 1)) == 0          # the largest size that fits still runs
        v_ref, _ = _forward(mh, {k: (t if k == "betas" else t[:F - 1].contiguous()) for k, t in x.items()}, F - 1, False)
    torch.cuda.synchronize()
    assert torch.equal(out[:F - 1], v_ref)
