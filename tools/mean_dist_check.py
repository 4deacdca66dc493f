"""Debug aid: uuo_mean_dist_accumulate vs numpy (bit level), mismatch pattern."""
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
import uuo_mocap_b200
from uuo_mocap_b200 import ops
for (F, M, V) in [(1, 1, 1), (2, 3, 48), (1, 13, 700), (2, 13, 700), (9, 13, 700), (16, 45, 480)]:
    rng = np.random.default_rng(3 + F)
    v = rng.normal(0, 0.5, size=(F, V, 3)).astype(np.float32)
    m = rng.normal(0, 0.5, size=(F, M, 3)).astype(np.float32)
    acc = np.zeros((M, V), dtype=np.float32)
    for f in range(F):
        acc += np.linalg.norm(v[f][None, :, :] - m[f][:, None, :], axis=-1)
    _, a = ops.mean_distance_argmin(torch.from_numpy(m).cuda(), torch.from_numpy(v).cuda(), finalize=False)
    a = a.cpu().numpy()
    bad = np.argwhere(a != acc)
    print((F, M, V), "mismatches", len(bad), "of", a.size, "max abs", float(np.abs(a - acc).max()))
    if len(bad):
        print("  first:", bad[:8].tolist(), a[tuple(bad[0])], acc[tuple(bad[0])])
        print("  by marker:", np.bincount(bad[:, 0], minlength=M).tolist())
