"""uuo-mocap_b200: B200-native SMPL marker-fitting hot path (drop-in for UUO-Mocap's
``SmplInference`` / ``smplx.SMPL`` call and the ``video_mocap.yaml`` stage optimisers).

Host code is Python/PyTorch (device memory, streams, ``torch.distributed``); every kernel
on the path is hand-written sm_100a CUDA behind the C-ABI declared in ``include/uuo_mocap.h``
and loaded with ctypes (``_lib.py``).  There is no CPU fallback: using an op without the
built library raises.
"""
__version__ = "0.1.0"
