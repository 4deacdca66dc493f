"""CPU oracle for the UUO-Mocap SMPL marker-fitting hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, in plain PyTorch on the CPU, the reference's arithmetic for the
path named by BASELINE.json:north_star.  Only ``tests/``, ``__graft_entry__.smoke()`` and
the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the product
package (``uuo-mocap_b200/``) never does and fails loudly when its CUDA library is missing.

Pinning status (see DESIGN.md "Oracle"):
* ``smpl_oracle``  -- PINNED: checked bit-for-bit against the reference's own vendored
  ``smplx`` (``/root/reference/third_party/smplx``) executed in the build container, and the
  outputs of that real reference are committed as ``tests/golden/smplx_*.npz`` by
  ``oracle/gen_golden.py``.
* ``p3d_oracle``   -- pytorch3d is an un-vendored, un-pinned conda dependency of the
  reference (``/root/reference/install.sh:7``) and is not installable here, and the
  reference has no tests or golden vectors for these ops: PARITY UNPINNED for the
  pytorch3d-derived ops (chamfer/knn K=1, 6D Gram-Schmidt, axis-angle->matrix).  They are a
  restatement of the published pytorch3d 0.7.x algorithm, cross-checked against scipy and
  against the reference's own numpy ``closest_point`` (``src/video_mocap/utils/points.py:19-42``).
* ``closures``     -- follows ``src/video_mocap/optimization.py`` line by line on top of
  the two modules above.
"""
