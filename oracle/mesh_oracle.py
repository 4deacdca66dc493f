"""TEST INFRASTRUCTURE ONLY.  CPU restatement of the marker-to-surface metric
(/root/reference/src/video_mocap/evaluation/metrics.py:26-46).  The reference calls igl.signed_distance, which is not
installable here (no network): PARITY UNPINNED for this function.  What it measures, mean |signed distance|, is the
unsigned point-to-triangle distance; it is restated in float64 numpy with a formulation independent of the CUDA
kernel's region walk: distance to the supporting plane when the projection falls inside the triangle, otherwise the
smallest of the three point-to-segment distances."""
import numpy as np


def _seg_dist2(p, a, b):
    ab = b - a
    t = np.einsum("...k,...k->...", p - a, ab) / np.maximum(np.einsum("...k,...k->...", ab, ab), 1e-300)
    t = np.clip(t, 0.0, 1.0)
    d = p - (a + t[..., None] * ab)
    return np.einsum("...k,...k->...", d, d)


def point_mesh_distance(points: np.ndarray, vertices: np.ndarray, faces: np.ndarray) -> np.ndarray:
    """points [F,M,3], vertices [F,V,3], faces [NF,3] -> dist [F,M] (float64)."""
    points, vertices = points.astype(np.float64), vertices.astype(np.float64)
    F, M, _ = points.shape
    out = np.zeros((F, M))
    for f in range(F):
        a, b, c = vertices[f][faces[:, 0]], vertices[f][faces[:, 1]], vertices[f][faces[:, 2]]     # [NF,3]
        p = points[f][:, None, :]                                                                   # [M,1,3]
        n = np.cross(b - a, c - a)
        nn = np.einsum("fk,fk->f", n, n)
        # barycentric coordinates of the projection
        ap = p - a[None]
        u = np.einsum("mfk,fk->mf", np.cross((b - a)[None], ap), n) / np.maximum(nn, 1e-300)        # weight of c
        v = np.einsum("mfk,fk->mf", np.cross(ap, (c - a)[None]), n) / np.maximum(nn, 1e-300)        # weight of b
        inside = (u >= 0) & (v >= 0) & (u + v <= 1) & (nn > 0)
        plane = np.einsum("mfk,fk->mf", ap, n) ** 2 / np.maximum(nn, 1e-300)
        edges = np.minimum(np.minimum(_seg_dist2(p, a[None], b[None]), _seg_dist2(p, b[None], c[None])),
                           _seg_dist2(p, c[None], a[None]))
        d2 = np.where(inside, plane, edges)
        out[f] = np.sqrt(d2.min(axis=1))
    return out
