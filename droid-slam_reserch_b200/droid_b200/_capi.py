"""ctypes binding of libdroid_b200.so (C-ABI declared in include/droid_b200.h).

There is deliberately no fallback: if the CUDA library is missing or a call fails, a RuntimeError is
raised.  Nothing under oracle/ is imported from here.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libdroid_b200.so")

F32, F16, F64 = 0, 1, 2
_lib = None

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

_SIGNATURES = {
    "db200_version": (_i, []),
    "db200_status_string": (C.c_char_p, [_i]),
    "db200_last_cuda_error": (C.c_char_p, []),
    "db200_launch_count": (C.c_longlong, []),
    "db200_profile_enable": (None, [_i]),
    "db200_profile_collect": (_i, [C.POINTER(C.c_double), C.POINTER(_i)]),
    "db200_debug_chol_trace": (_i, [_vp]),
    "db200_debug_fp64_peak_gflops": (C.c_double, [_vp]),
    "db200_corr_index_forward": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "db200_corr_index_backward": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "db200_corr_pyramid_lookup": (_i, [C.POINTER(_vp), _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "db200_corr_volume_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "db200_corr_volume_pyramid": (_i, [_vp, _vp, C.POINTER(_vp), _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "db200_corr_tiled_levels": (C.c_uint, [_i, _i, _i, _i, _i]),
    "db200_corr_volume_pyramid_tiled": (_i, [_vp, _vp, C.POINTER(_vp), _i, _i, _i, _i, _i, _i, _vp, _sz, C.c_uint, _vp]),
    "db200_corr_pyramid_lookup_tiled": (_i, [C.POINTER(_vp), _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, C.c_uint, _vp]),
    "db200_altcorr_forward": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "db200_altcorr_backward": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "db200_altcorr_workspace_bytes": (_sz, [_i] * 7),
    "db200_altcorr_forward_ws": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "db200_altcorr_backward_ws": (_i, [_vp] * 7 + [_i] * 8 + [_vp, _sz, _vp]),
    "db200_altcorr_static_max_bytes": (_sz, []),
    "db200_altcorr_forward_static": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "db200_altcorr_status": (_i, [_vp, _vp]),
    "db200_altcorr_cells": (C.c_longlong, [_vp, _vp]),
    "db200_debug_altcorr_waits": (_i, [_vp, _vp, _vp]),
    "db200_altcorr_pyramid_workspace_bytes": (_sz, [_i] * 4),
    "db200_altcorr_pyramid_forward": (_i, [_vp, _vp, C.POINTER(_vp), C.POINTER(_vp), _vp, _vp, C.POINTER(_vp), C.c_longlong, _i, _i,
                                           _i, _i, _vp, _vp, _vp, C.c_longlong, C.c_longlong, C.c_longlong, _vp, C.c_longlong,
                                           C.c_longlong, C.c_longlong, C.c_longlong, _i, _i, _i, _vp, _sz, _vp]),
    "db200_split_hilo_f16": (_i, [_vp, _vp, _vp, _sz, _sz, _vp, _vp]),
    "db200_frame_distance": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp]),
    "db200_proximity_workspace_bytes": (_sz, [_i, _i, _i]),
    "db200_proximity_max_edges": (_i, [_i, _i, _i, _i, _i]),
    "db200_proximity_edges": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _f, C.c_double, _i, _i, _vp, _vp, _i, _vp, _i,
                                   _vp, _vp, _sz, _vp]),
    "db200_filter_repeated_edges": (_i, [_vp, _vp, _i, _vp, _vp, _i, _vp, _vp]),
    "db200_projmap": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "db200_reproject": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "db200_iproj": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "db200_depth_filter": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "db200_ba_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i, _i, _i]),
    "db200_ba": (_i, [_vp] * 9 + [_i] * 7 + [_f, _f, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "db200_ba_status": (_i, [_vp, _vp]),
    "db200_ba_accumulate": (_i, [_vp] * 9 + [_i] * 11 + [_vp, _sz, _vp]),
    "db200_ba_solve_update": (_i, [_vp] * 7 + [_i] * 6 + [_f, _f, _i, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "db200_ba_system_ptr": (_vp, [_vp, _i, _i, _i, _i, _i, _i, _i, _i]),
    "db200_ba_system_elems": (_sz, [_i, _i]),
}

EXPORTED = tuple(_SIGNATURES)


def lib():
    """Load the shared library (once).  Raises if it has not been built (`__graft_entry__.build()`)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is not built (run `python -c 'import __graft_entry__ as g; "
                "g.build()'` or `make -C droid-slam_reserch_b200/csrc`). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def status_string(code: int) -> str:
    return lib().db200_status_string(int(code)).decode()


def check(code: int, what: str):
    if code < 0:
        extra = f"; last CUDA error: {lib().db200_last_cuda_error().decode()}" if code == -2 else ""
        raise RuntimeError(f"{what}: {status_string(code)} (status {code}){extra}")
    return code
