"""Data formats either side of the fitting path: the marker capture (.c3d) and the video
prior (PHALP / HMR .pkl) readers.  The writer half (``*_stageii.npz``) and the frame-rate resampling live in stageii.py.

* ``Markers`` mirrors /root/reference/src/video_mocap/markers/markers.py:6-54.  The reference parses the file with
  ``ezc3d`` (not installable here); ``read_c3d`` is a numpy reader of the documented C3D layout (c3d.org "The C3D File
  Format": 512-byte blocks, header block, parameter section of group / parameter records, interleaved point + analog
  frames; Intel byte order, integer or floating-point storage).  Like ezc3d it reports a point whose 4th word
  (residual) is negative as NaN -- the caller's ``nan_to_num`` (multimodal.py:188) turns those into the (0,0,0)
  "dropped marker" the fit masks out.
* ``ImgSmpl`` mirrors /root/reference/src/video_mocap/img_smpl/img_smpl.py:13-150: per-frame PHALP records -> prior
  tensors (trans = the low-pelvis 3-D joint, root_orient = correction @ global_orient, pose_body, betas, img_mask),
  frames without a track filled from their neighbours (linear / quaternion slerp), camera boxes, foot contacts
  (utils/img_smpl_utils.py:55-95).  Pinned on the reference's own class executed from source (tests/golden/img_smpl.npz).
"""
from __future__ import annotations

import struct
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .stageii import matrix_to_axis_angle, matrix_to_quaternion, quaternion_to_matrix, unitquat_slerp

# ------------------------------------------------------------------------------------------------
# C3D
# ------------------------------------------------------------------------------------------------
_C3D_BLOCK = 512


def _c3d_parameters(buf: bytes, start: int) -> Dict[str, Dict[str, object]]:
    """Parameter section -> {GROUP: {PARAM: value}}.  Records: name length (negative = locked), id (negative = group),
    name, int16 offset to the next record, then for a parameter: element type (-1 char, 1 byte, 2 int16, 4 float),
    number of dimensions, dimensions, data, description; for a group: description."""
    n_blocks, proc = buf[start + 2], buf[start + 3]
    if proc != 84:
        raise ValueError("only Intel (little-endian, IEEE float) C3D files are supported (processor type %d)" % proc)
    end = start + max(1, n_blocks) * _C3D_BLOCK
    pos = start + 4
    groups: Dict[int, str] = {}
    params: List = []
    while pos + 2 <= min(end, len(buf)):
        nlen, gid = struct.unpack_from("<bb", buf, pos)
        nlen = abs(nlen)
        if nlen == 0:
            break
        name = buf[pos + 2:pos + 2 + nlen].decode("latin-1").strip().upper()
        off_pos = pos + 2 + nlen
        (offset,) = struct.unpack_from("<h", buf, off_pos)
        body = off_pos + 2
        if gid < 0:
            groups[-gid] = name
        else:
            etype, ndim = struct.unpack_from("<bB", buf, body)
            dims = list(buf[body + 2:body + 2 + ndim])
            count = int(np.prod(dims)) if ndim else 1
            data = body + 2 + ndim
            if etype == -1:
                raw = buf[data:data + count]
                if ndim <= 1:
                    val = raw.decode("latin-1").strip()
                else:                                   # [len, n, ...] column-major: n strings of dims[0] characters
                    L = dims[0]
                    val = [raw[i * L:(i + 1) * L].decode("latin-1").strip() for i in range(count // max(L, 1))]
            else:
                dt = {1: "<u1", 2: "<i2", 4: "<f4"}[abs(etype)]
                arr = np.frombuffer(buf, dtype=dt, count=count, offset=data)
                val = arr.reshape(dims[::-1]).T if ndim > 1 else (arr.copy() if ndim == 1 else arr[0])
            params.append((gid, name, val))
        if offset <= 0:
            break
        pos = off_pos + offset
    out: Dict[str, Dict[str, object]] = {g: {} for g in groups.values()}
    for gid, name, val in params:
        out.setdefault(groups.get(gid, "GROUP%d" % gid), {})[name] = val
    return out


def read_c3d(filename: str) -> Dict[str, object]:
    """-> {"points": [F,M,3] float32 in the file's units (NaN where the residual word is negative), "residuals" [F,M],
    "rate": POINT:RATE, "units": POINT:UNITS, "labels": POINT:LABELS (list of str, may be empty), "parameters": {...}}."""
    with open(filename, "rb") as fh:
        buf = fh.read()
    if len(buf) < _C3D_BLOCK or buf[1] != 0x50:
        raise ValueError("%s is not a C3D file (key byte 0x50 missing)" % filename)
    par_block = buf[0]
    n_points, n_analog, first, last, _gap = struct.unpack_from("<HHHHH", buf, 2)
    (scale,) = struct.unpack_from("<f", buf, 12)
    data_block, analog_per_frame = struct.unpack_from("<HH", buf, 16)
    (rate_hdr,) = struct.unpack_from("<f", buf, 20)
    par = _c3d_parameters(buf, (par_block - 1) * _C3D_BLOCK)
    point = par.get("POINT", {})
    n_points = int(point.get("USED", n_points))
    if "SCALE" in point:
        scale = float(point["SCALE"])
    if "DATA_START" in point:
        data_block = int(point["DATA_START"]) & 0xFFFF
    n_frames = last - first + 1
    if "FRAMES" in point:                                # files longer than 65535 frames keep the count here
        fr = point["FRAMES"]
        fr = int(np.asarray(fr).reshape(-1)[0])
        n_frames = fr + 65536 if fr < 0 else fr
    if last - first + 1 > 0 and n_frames in (0, 65535):
        n_frames = last - first + 1
    is_float = scale < 0
    word = 4 if is_float else 2
    frame_words = 4 * n_points + n_analog
    start = (data_block - 1) * _C3D_BLOCK
    avail = (len(buf) - start) // (frame_words * word) if frame_words else 0
    n_frames = min(n_frames, avail)
    raw = np.frombuffer(buf, dtype="<f4" if is_float else "<i2", count=n_frames * frame_words, offset=start)
    pts = raw.reshape(n_frames, frame_words)[:, :4 * n_points].reshape(n_frames, n_points, 4)
    if is_float:
        xyz = pts[..., :3].astype(np.float32)
        resid = pts[..., 3].astype(np.float32)
        # the 4th float packs (camera mask << 8 | residual) as an integer value; a negative value marks an invalid point
    else:
        xyz = pts[..., :3].astype(np.float32) * np.float32(abs(scale))
        resid = pts[..., 3].astype(np.float32)
    xyz = np.where((resid < 0)[..., None], np.float32(np.nan), xyz)
    labels = point.get("LABELS", [])
    if isinstance(labels, str):
        labels = [labels]
    units = point.get("UNITS", "mm")
    if isinstance(units, list):
        units = units[0] if units else "mm"
    rate = float(point.get("RATE", rate_hdr))
    return {"points": xyz, "residuals": resid, "rate": rate, "units": str(units), "labels": list(labels)[:n_points],
            "parameters": par, "first_frame": int(first)}


def write_c3d(filename: str, points: np.ndarray, rate: float, labels: Optional[Sequence[str]] = None, units: str = "mm",
              integer_scale: Optional[float] = None) -> None:
    """Minimal C3D writer (points only; used by the tests and to hand captures to other tools).  ``points`` [F,M,3] in
    ``units``; NaN rows are written as invalid (residual -1).  ``integer_scale``: store int16 words with that POINT:SCALE
    instead of floats."""
    pts = np.asarray(points, dtype=np.float32)
    F, M = pts.shape[:2]
    labels = list(labels) if labels is not None else ["M%03d" % i for i in range(M)]

    def rec_group(gid, name, desc=""):
        b = struct.pack("<bb", len(name), -gid) + name.encode()
        body = struct.pack("<B", len(desc)) + desc.encode()
        return b + struct.pack("<h", 2 + len(body)) + body

    def rec_param(gid, name, etype, dims, data: bytes):
        b = struct.pack("<bb", len(name), gid) + name.encode()
        body = struct.pack("<bB", etype, len(dims)) + bytes(dims) + data + struct.pack("<B", 0)
        return b + struct.pack("<h", 2 + len(body)) + body

    L = max(4, max(len(s) for s in labels))
    lab = b"".join(s.ljust(L).encode("latin-1") for s in labels)
    scale = -1.0 if integer_scale is None else float(integer_scale)
    recs = rec_group(1, "POINT", "3-D point parameters")
    recs += rec_param(1, "USED", 2, [], struct.pack("<h", M))
    recs += rec_param(1, "FRAMES", 2, [], struct.pack("<H", min(F, 65535)))
    recs += rec_param(1, "SCALE", 4, [], struct.pack("<f", scale))
    recs += rec_param(1, "RATE", 4, [], struct.pack("<f", float(rate)))
    recs += rec_param(1, "UNITS", -1, [len(units)], units.encode())
    recs += rec_param(1, "LABELS", -1, [L, M], lab)
    n_par_blocks = (4 + len(recs) + 64 + _C3D_BLOCK - 1) // _C3D_BLOCK
    data_block = 2 + n_par_blocks
    ds = rec_param(1, "DATA_START", 2, [], struct.pack("<H", data_block))
    recs += ds
    par = bytes([1, 0x50, n_par_blocks, 84]) + recs + struct.pack("<bb", 0, 0)
    par = par.ljust(n_par_blocks * _C3D_BLOCK, b"\0")
    hdr = bytearray(_C3D_BLOCK)
    hdr[0], hdr[1] = 2, 0x50
    struct.pack_into("<HHHHH", hdr, 2, M, 0, 1, min(F, 65535), 0)
    struct.pack_into("<f", hdr, 12, scale)
    struct.pack_into("<HH", hdr, 16, data_block, 0)
    struct.pack_into("<f", hdr, 20, float(rate))
    bad = np.isnan(pts).any(-1)
    resid = np.where(bad, -1.0, 0.0).astype(np.float32)
    xyz = np.where(bad[..., None], 0.0, pts).astype(np.float32)
    if integer_scale is None:
        data = np.concatenate([xyz, resid[..., None]], -1).astype("<f4").tobytes()
    else:
        q = np.concatenate([np.rint(xyz / np.float32(integer_scale)), resid[..., None]], -1)
        data = q.astype("<i2").tobytes()
    with open(filename, "wb") as fh:
        fh.write(bytes(hdr) + par + data)
        fh.write(b"\0" * ((-len(data)) % _C3D_BLOCK))


class Markers:
    """markers/markers.py:6-54: points in METRES [F,M,3], frequency, labels."""

    def __init__(self, filename: str, shuffle: bool = False):
        self.filename = filename
        self.shuffle = True                                                 # (sic) :9
        c = read_c3d(filename)
        self.units = c["units"]
        self.scale_factor = {"m": 1, "cm": 100, "mm": 1000}.get(self.units)  # :12-18
        if self.scale_factor is None:
            raise ValueError("POINT:UNITS must be m, cm or mm, got %r" % (self.units,))
        self.points = c["points"] / self.scale_factor                       # :20-21
        if shuffle:                                                         # :23-28
            tmp = np.zeros_like(self.points)
            for f in range(self.points.shape[0]):
                tmp[f] = np.ascontiguousarray(self.points[f, np.random.permutation(self.points.shape[1])])
            self.points = tmp
        self.freq = int(c["rate"])                                          # :30
        if c["labels"]:
            self.labels = c["labels"]                                       # :32-33

    def get_points(self):
        return self.points

    def set_points(self, points):
        self.points = points

    def get_labels(self):
        return self.labels

    def get_num_markers(self):
        return self.points.shape[1]

    def __len__(self):
        return self.points.shape[0]

    def get_duration(self):
        return self.freq * self.points.shape[0]                             # (sic) :50-51

    def get_frequency(self):
        return self.freq


# ------------------------------------------------------------------------------------------------
# PHALP / HMR video prior
# ------------------------------------------------------------------------------------------------
JOINTS_2D_LABELS = {  # utils/img_smpl_utils.py:4-50 (the entries this module uses)
    "pelvis_low": 8, "l_toe_in": 19, "l_toe_out": 20, "r_toe_in": 22, "r_toe_out": 23,
}
_CORRECTION = np.array([[1, 0, 0], [0, 0, 1], [0, -1, 0]], dtype=np.float32)   # img_smpl.py:40-44


def get_foot_contacts(joints_2d: np.ndarray, freq: float) -> np.ndarray:
    """utils/img_smpl_utils.py:55-95: a foot is in contact when both of its toe key points move slower than a threshold
    scaled by the person's 2-D extent.  joints_2d [F,45,2] -> [F,2] (left, right)."""
    min_x, max_x = np.min(joints_2d[:, :, 0], axis=1), np.max(joints_2d[:, :, 0], axis=1)
    min_y, max_y = np.min(joints_2d[:, :, 1], axis=1), np.max(joints_2d[:, :, 1], axis=1)
    extent = np.sqrt(((max_x - min_x) ** 2) + ((max_y - min_y) ** 2))
    extent = np.maximum(extent, np.ones_like(extent) * 0.01)
    threshold = np.repeat(np.expand_dims(0.0001 / extent, axis=-1), repeats=joints_2d.shape[1], axis=-1)
    vel = np.concatenate((np.zeros((1,) + joints_2d.shape[1:]), joints_2d[1:] - joints_2d[:-1]), axis=0) / freq
    contact = np.linalg.norm(vel, axis=-1) < threshold
    L = JOINTS_2D_LABELS
    out = np.ones((joints_2d.shape[0], 2))
    for g, grp in enumerate(([L["l_toe_in"], L["l_toe_out"]], [L["r_toe_in"], L["r_toe_out"]])):
        for j in grp:
            out[:, g] *= contact[:, j]
    return out


def load_phalp_pkl(filename: str) -> Dict:
    """The PHALP tracker's per-frame dictionary (test/test.py:88-96 loads it with joblib)."""
    import joblib
    return joblib.load(filename)


class ImgSmpl:
    """img_smpl/img_smpl.py:13-150 with the same attributes: trans [F,3], root_orient / hmr_root_orient [F,1,3,3],
    pose_body [F,23,3,3], betas [F,10], img_mask [F] bool, camera_bbox, center, scale, size, foot_contacts [F,2]."""

    def __init__(self, data: Dict, freq: float):
        self.data, self.freq = data, freq
        keys = sorted(data.keys())
        F = len(keys)
        trans = np.zeros((F, 3), np.float32)
        root = np.zeros((F, 1, 3, 3), np.float32)
        hmr_root = np.zeros((F, 1, 3, 3), np.float32)
        pose = np.zeros((F, 23, 3, 3), np.float32)
        betas = np.zeros((F, 10), np.float32)
        mask = np.zeros(F, bool)
        self.camera_bbox = np.zeros((F, 3), np.float32)
        self.center = np.zeros((F, 2), np.float32)
        self.size = np.zeros((F, 2), np.float32)
        self.scale = np.zeros((F, 1), np.float32)
        joints_2d = np.zeros((F, 45, 2))
        for i, k in enumerate(keys):                                                     # :33-49, :117-137
            d = data[k]
            if len(d["tracked_ids"]) > 0:
                mask[i] = True
                go = np.asarray(d["smpl"][0]["global_orient"], np.float32)
                hmr_root[i] = go
                trans[i] = np.asarray(d["3d_joints"][0], np.float32)[JOINTS_2D_LABELS["pelvis_low"]]
                root[i] = _CORRECTION @ go
                pose[i] = np.asarray(d["smpl"][0]["body_pose"], np.float32)
                betas[i] = np.asarray(d["smpl"][0]["betas"], np.float32)
            if len(d["camera_bbox"]) > 0:
                self.camera_bbox[i] = d["camera_bbox"][0]
                self.center[i] = d["center"][0]
                self.scale[i] = d["scale"][0]
                self.size[i] = d["size"][0]
            try:
                j2 = np.asarray(d["2d_joints"][0], np.float64).reshape(-1)
                n = min(45, j2.shape[0] // 2)
                joints_2d[i, :n] = j2[:2 * n].reshape(n, 2)
            except (IndexError, KeyError, TypeError, ValueError):
                pass
        T = torch.from_numpy
        trans, root, hmr_root, pose, betas = T(trans), T(root), T(hmr_root), T(pose), T(betas)
        self.img_mask = T(mask)
        # frames without a track (:52-95): copy the nearest valid frame at the ends, blend between the two neighbours
        valid = np.where(mask)[0]
        if valid.size and valid.size < F:
            miss = np.where(~mask)[0]
            pos = np.searchsorted(valid, miss)
            has_l, has_r = pos > 0, pos < valid.size
            left = valid[np.clip(pos - 1, 0, valid.size - 1)]
            right = valid[np.clip(pos, 0, valid.size - 1)]
            for arr in (hmr_root, trans, root, pose, betas):
                only_r, only_l = miss[~has_l], miss[~has_r]
                arr[only_r] = arr[right[~has_l]]
                arr[only_l] = arr[left[~has_r]]
            both = has_l & has_r
            if both.any():
                m, l, r = miss[both], left[both], right[both]
                alpha = torch.from_numpy(((m - l) / (r - l)).astype(np.float32))
                a1 = alpha.view(-1, 1)
                trans[m] = trans[l] * (1.0 - a1) + trans[r] * a1
                betas[m] = betas[l] * (1.0 - a1) + betas[r] * a1
                for arr in (hmr_root, root, pose):
                    q = unitquat_slerp(matrix_to_quaternion(arr[l]), matrix_to_quaternion(arr[r]), alpha.view(-1, 1, 1))
                    arr[m] = quaternion_to_matrix(q).to(arr.dtype)
        self.trans, self.root_orient, self.hmr_root_orient, self.pose_body, self.betas = trans, root, hmr_root, pose, betas
        self.camera_bbox, self.center = T(self.camera_bbox), T(self.center)
        self.scale, self.size = T(self.scale), T(self.size)
        self.foot_contacts = T(get_foot_contacts(joints_2d, freq).astype(np.float32))

    def get_smpl(self) -> Dict:
        """:139-150: the AMASS-style dict (axis-angle poses [F,72])."""
        poses = torch.flatten(matrix_to_axis_angle(torch.cat((self.root_orient, self.pose_body), dim=1)), start_dim=1)
        return {"betas": self.betas[0].detach().cpu().numpy(), "gender": np.array("neutral"), "mocap_frame_rate": self.freq,
                "poses": poses.detach().cpu().numpy(), "trans": self.trans.detach().cpu().numpy()}
