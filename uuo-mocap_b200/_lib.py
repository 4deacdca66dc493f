"""ctypes binding of libuuomocap.so (the C-ABI in include/uuo_mocap.h).

There is NO fallback: if the shared library is missing every op raises.  ``lib()`` loads the
in-tree build (``uuo-mocap_b200/lib/libuuomocap.so``); build it with ``python -m uuo_mocap_b200.build``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# UUO_MOCAP_LIB: another build of the same library (tools/ab.sh times two builds on one box); the default is the in-tree one
LIB_PATH = os.environ.get("UUO_MOCAP_LIB") or os.path.join(_HERE, "lib", "libuuomocap.so")

_P = C.c_void_p
_I = C.c_int
_F = C.c_float
_SZ = C.c_size_t

# name -> (restype, argtypes); must list every symbol include/uuo_mocap.h declares
SIGNATURES = {
    "uuo_last_error": (C.c_char_p, []),
    "uuo_version": (_I, []),
    "uuo_model_create": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, C.POINTER(_P)]),
    "uuo_model_destroy": (None, [_P]),
    "uuo_model_info": (_I, [_P, _I]),
    "uuo_lbs_workspace_bytes": (_SZ, [_P, _I]),
    "uuo_lbs_forward": (_I, [_P, _P, _I, _P, _P, _I, _P, _P, _I, _P, _P, _P, _P]),
    "uuo_lbs_forward_static": (_I, [_P, _P, _I, _P, _P, _I, _P, _P, _I, _P, _P, _P]),
    "uuo_lbs_backward": (_I, [_P, _P, _I, _P, _P, _I, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "uuo_lbs_backward_rows": (_I, [_P, _P, _I, _P, _P, _I, _P, _I, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "uuo_lbs_backward_rows_cached": (_I, [_P, _P, _I, _P, _P, _I, _P, _I, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "uuo_lbs_cache_vposed_rows": (_I, [_P, _I, _P, _I, _P, _P]),
    "uuo_lbs_forward_rows": (_I, [_P, _P, _I, _P, _P, _I, _P, _P, _I, _P, _I, _I, _P, _P, _P, _P]),
    "uuo_nn_argmin": (_I, [_P, _P, _I, _I, _I, _P, _P, _P]),
    "uuo_nn_argmin_ordered": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P]),
    "uuo_model_vertex_order": (_P, [_P]),
    "uuo_model_vertex_order_host": (_I, [_P, _P]),
    "uuo_nn_argmin_presorted": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P, _P]),
    "uuo_nn_argmin_presorted_fit": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P, _P, _P, _P]),
    "uuo_marker_mask": (_I, [_P, _I, _P, _P]),
    "uuo_chamfer_rows": (_I, [_P, _P, _P, _P, _I, _I, _I, _F, _P, _P, _P]),
    "uuo_marker_loss_rows": (_I, [_P, _P, _P, _I, _F, _F, _P, _P, _P]),
    "uuo_mean_dist_accumulate": (_I, [_P, _P, _P, _I, _I, _I, _P, _P]),
    "uuo_mean_argmin": (_I, [_P, _I, _I, _F, _P, _P]),
    "uuo_mean_dist_accumulate_bounded": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P]),
    "uuo_mean_dist_candidate_sums": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P]),
    "uuo_sqdiff_accumulate": (_I, [_P, _P, _SZ, _F, _P, _P, _P]),
    "uuo_temporal_accumulate": (_I, [_P, _P, _P, _I, _I, _F, _P, _P, _P]),
    "uuo_mask_vertices": (_I, [_P, _P, _I, _I, _I, _F, _P]),
    "uuo_marker_mean": (_I, [_P, _I, _I, _P, _P]),
    "uuo_transvel_accumulate": (_I, [_P, _P, _P, _P, _I, _F, _P, _P, _P]),
    "uuo_point_error_mean": (_I, [_P, _P, _I, _I, _P, _I, _I, _F, _P, _P, _P]),
    "uuo_similarity_transform": (_I, [_P, _P, _I, _I, _P, _P]),
    "uuo_point_mesh_distance": (_I, [_P, _P, _P, _I, _I, _I, _I, _P, _P]),
    "uuo_profile_enable": (_I, [_I]),
    "uuo_profile_read": (_I, [_I, C.c_char_p, _I, C.POINTER(_F), C.POINTER(_I)]),
    "uuo_comm_create": (_I, [_I, _I, _I, C.POINTER(_P)]),
    "uuo_comm_set_timeout": (_I, [_P, C.c_double]),
    "uuo_comm_handle_bytes": (_I, []),
    "uuo_comm_export": (_I, [_P, _P]),
    "uuo_comm_connect": (_I, [_P, _P]),
    "uuo_comm_connect_local": (_I, [_P, C.POINTER(_P)]),
    "uuo_comm_destroy": (None, [_P]),
    "uuo_comm_all_reduce_sum": (_I, [_P, _P, _I, _P]),
    "uuo_comm_halo_exchange": (_I, [_P, _P, _I, _I, _I, _P, _P, _I, _I, _P]),
    "uuo_comm_group_all_reduce_sum": (_I, [C.POINTER(_P), C.POINTER(_P), _I, _I, _P]),
    "uuo_comm_group_halo_exchange": (_I, [C.POINTER(_P), C.POINTER(_P), C.POINTER(_I), _I, _I, C.POINTER(_P), C.POINTER(_P),
                                          _I, _I, _I, _P]),
    "uuo_adam_step": (_I, [_P, _P, _P, _P, _SZ, _F, _F, _F, _F, _I, _P]),
    "uuo_adam_step_dev": (_I, [_P, _P, _P, _P, _SZ, _F, _F, _F, _F, _P, _P]),
}

_lib = None


class UuoError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise UuoError(
                "libuuomocap.so is not built (%s); run `python -m uuo_mocap_b200.build`. "
                "There is no CPU/PyTorch fallback for this path." % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int, exc=UuoError):
    if rc != 0:
        msg = lib().uuo_last_error().decode("utf-8", "replace")
        raise (ValueError if rc == -1 and exc is UuoError else exc)(msg or "libuuomocap error %d" % rc)


def ptr(t):
    """data pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
