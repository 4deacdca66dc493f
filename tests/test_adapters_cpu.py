"""N2 readers: the .c3d marker reader and the PHALP .pkl -> prior tensors adapter.  CPU only."""
import numpy as np
import pytest
import torch

from conftest import rel_err


@pytest.mark.parametrize("units,integer_scale", [("mm", None), ("m", None), ("cm", None), ("mm", 0.1)])
def test_c3d_round_trip_and_markers_class(tmp_path, units, integer_scale):
    """markers/markers.py:6-54: points come back in METRES [F,M,3], RATE as int, LABELS; invalid points are NaN."""
    from uuo_mocap_b200.adapters import Markers, read_c3d, write_c3d
    rng = np.random.default_rng(4)
    F, M = 37, 41
    scale = {"m": 1.0, "cm": 100.0, "mm": 1000.0}[units]
    pts_m = rng.normal(0, 0.6, size=(F, M, 3)).astype(np.float32)
    pts = pts_m * np.float32(scale)
    pts[3, 5] = np.nan
    pts[10:14, 0] = np.nan
    labels = ["MK%02d" % i for i in range(M)]
    fn = str(tmp_path / "take.c3d")
    write_c3d(fn, pts, 120.0, labels, units, integer_scale)
    raw = read_c3d(fn)
    assert raw["units"] == units and raw["rate"] == 120.0 and raw["labels"] == labels
    assert raw["points"].shape == (F, M, 3)
    mk = Markers(fn)
    got = mk.get_points()
    assert got.shape == (F, M, 3) and len(mk) == F and mk.get_num_markers() == M
    assert mk.get_frequency() == 120 and mk.get_labels() == labels
    bad = np.isnan(pts).any(-1)
    assert np.array_equal(np.isnan(got).any(-1), bad)
    tol = 1e-6 if integer_scale is None else 0.5 * integer_scale / scale + 1e-6
    assert np.nanmax(np.abs(got - pts_m)[~bad]) <= tol
    # the caller's nan_to_num (multimodal.py:188) turns invalid points into the (0,0,0) the fit masks out
    z = torch.nan_to_num(torch.from_numpy(got), nan=0.0)
    assert float(z[3, 5].abs().sum()) == 0.0


def test_c3d_rejects_other_files(tmp_path):
    from uuo_mocap_b200.adapters import read_c3d
    fn = tmp_path / "x.c3d"
    fn.write_bytes(b"\0" * 1024)
    with pytest.raises(ValueError):
        read_c3d(str(fn))


def test_img_smpl_matches_reference_class(golden):
    """ImgSmpl against the REFERENCE'S OWN class executed from source (oracle/gen_golden_adapters.py): masks, the
    pelvis translation, the corrected root, frames without a track copied / blended from their neighbours, camera
    boxes, foot contacts, the axis-angle export."""
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.adapters import ImgSmpl
    g = golden("img_smpl.npz")
    got = ImgSmpl(synthetic.make_phalp_like(40, seed=17), 30.0)
    assert np.array_equal(got.img_mask.numpy(), g["img_mask"])
    for k in ("trans", "betas", "camera_bbox", "center", "scale", "size"):
        assert rel_err(getattr(got, k).numpy(), g[k]) < 1e-6, k
    for k in ("root_orient", "hmr_root_orient", "pose_body"):
        assert np.abs(getattr(got, k).numpy() - g[k]).max() < 2e-6, k
    assert np.array_equal(got.foot_contacts.numpy(), g["foot_contacts"])
    s = got.get_smpl()
    assert np.abs(s["poses"] - g["poses_aa"]).max() < 1e-5
    assert rel_err(s["betas"], g["smpl_betas"]) < 1e-6 and rel_err(s["trans"], g["smpl_trans"]) < 1e-6


def test_phalp_pkl_loader(tmp_path):
    import joblib
    from uuo_mocap_b200 import synthetic
    from uuo_mocap_b200.adapters import ImgSmpl, load_phalp_pkl
    fn = str(tmp_path / "track.pkl")
    joblib.dump(synthetic.make_phalp_like(12, seed=3, gaps=((4, 6),)), fn)
    p = ImgSmpl(load_phalp_pkl(fn), 30.0)
    assert p.pose_body.shape == (12, 23, 3, 3) and int(p.img_mask.sum()) == 10
    R = p.pose_body.reshape(-1, 3, 3)
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand_as(R), atol=1e-5)
