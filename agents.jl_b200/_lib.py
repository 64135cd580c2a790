"""ctypes binding of libagents_b200.so (the C ABI in include/agents_b200.h).

There is no CPU fallback: if the library is missing this raises, and
`ab_create` itself fails without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("AGENTS_B200_LIB") or os.path.join(_HERE, "libagents_b200.so")   # override: kernel-variant experiments

AB_OK = 0
ERR_INVALID, ERR_CUDA, ERR_OUTSIDE, ERR_SPACING, ERR_OCCUPIED, ERR_NUMERIC, ERR_STATE, ERR_COMM = (
    -1, -2, -3, -4, -5, -6, -7, -8)
SPACE_CONTINUOUS, SPACE_GRID_MULTI, SPACE_GRID_SINGLE = 0, 1, 2
F64, F32, I32, I64, U8 = 0, 1, 2, 3, 4
METRICS = {"chebyshev": 0, "manhattan": 1, "euclidean": 2}
SEARCH = {"approximate": 0, "exact": 1}


class SpaceDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("D", C.c_int32), ("extent", C.c_double * 3), ("spacing", C.c_double),
                ("periodic", C.c_uint8 * 3), ("_pad", C.c_uint8 * 5), ("metric", C.c_int32), ("dtype", C.c_int32),
                ("grid_dims", C.c_int64 * 3)]


class FlockParams(C.Structure):
    _fields_ = [("speed", C.c_double), ("cohere_factor", C.c_double), ("separation", C.c_double),
                ("separate_factor", C.c_double), ("match_factor", C.c_double), ("visual_distance", C.c_double),
                ("search", C.c_int32), ("per_agent_field", C.c_int32), ("kernel", C.c_int32), ("_pad", C.c_int32)]


class PairForceParams(C.Structure):
    _fields_ = [("dt", C.c_double), ("cutoff", C.c_double), ("field_r", C.c_int32), ("field_k", C.c_int32),
                ("field_mass", C.c_int32), ("_pad", C.c_int32)]


class SchellingParams(C.Structure):
    _fields_ = [("min_to_be_happy", C.c_int32), ("relocate", C.c_int32), ("seed", C.c_uint64)]


class ReduceSpec(C.Structure):
    _fields_ = [("field", C.c_int32), ("comp", C.c_int32), ("op", C.c_int32), ("filter_field", C.c_int32),
                ("filter_comp", C.c_int32), ("filter_cmp", C.c_int32), ("filter_value", C.c_double)]


RED = {"sum": 0, "mean": 1, "minimum": 2, "min": 2, "maximum": 3, "max": 3, "count": 4, "length": 4}
CMP = {None: 0, "==": 1, "!=": 2, "<": 3, "<=": 4, ">": 5, ">=": 6}


# every symbol include/agents_b200.h declares: name -> (restype, argtypes)
_H = C.c_void_p
_SIGS = {
    "ab_create": (C.c_int, [C.POINTER(SpaceDesc), C.c_int, C.POINTER(_H)]),
    "ab_destroy": (C.c_int, [_H]),
    "ab_last_error": (C.c_char_p, [_H]),
    "ab_sync": (C.c_int, [_H]),
    "ab_set_async": (C.c_int, [_H, C.c_int]),
    "ab_version": (C.c_int, []),
    "ab_nagents": (C.c_int64, [_H]),
    "ab_alloc_pinned": (C.c_void_p, [C.c_int64]),
    "ab_free_pinned": (None, [C.c_void_p]),
    "ab_register_field": (C.c_int, [_H, C.c_char_p, C.c_int, C.c_int]),
    "ab_field_id": (C.c_int, [_H, C.c_char_p]),
    "ab_resize": (C.c_int, [_H, C.c_int64]),
    "ab_upload": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_int64]),
    "ab_download": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_int64]),
    "ab_build_index": (C.c_int, [_H]),
    "ab_set_refinement": (C.c_int, [_H, C.c_int]),
    "ab_cells": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "ab_nearby_count": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_int, C.c_int, C.c_void_p]),
    "ab_nearby_fill": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "ab_pairs_count": (C.c_int, [_H, C.c_double, C.c_int, C.POINTER(C.c_int64)]),
    "ab_pairs_fill": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64]),
    "ab_nearest_neighbor": (C.c_int, [_H, C.c_double, C.c_void_p, C.c_int64]),
    "ab_nearby_positions_count": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_double, C.c_int, C.c_void_p]),
    "ab_nearby_positions_fill": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_double, C.c_int, C.c_void_p, C.c_void_p]),
    "ab_pairs_nearest": (C.c_int, [_H, C.c_double, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ab_pairs_nearest_fill": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64]),
    "ab_elastic_collisions": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.POINTER(C.c_int64)]),
    "ab_spatial_lookup": (C.c_int, [_H, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "ab_update_vel_field": (C.c_int, [_H, C.c_void_p, C.c_int, C.c_void_p, C.c_double, C.c_void_p, C.c_int64]),
    "ab_move": (C.c_int, [_H, C.c_double, C.c_void_p, C.c_int64]),
    "ab_walk": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]),
    "ab_layout_get": (C.c_int, [_H, C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "ab_layout_set": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_int32, C.c_int64]),
    "ab_randomwalk": (C.c_int, [_H, C.c_double, C.c_uint64, C.c_void_p, C.c_int64, C.c_int, C.c_int]),
    "ab_randomwalk_angles": (C.c_int, [_H, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "ab_step_flock": (C.c_int, [_H, C.POINTER(FlockParams), C.c_int64]),
    "ab_step_flock_active": (C.c_int, [_H, C.POINTER(FlockParams), C.c_int64, C.c_void_p, C.c_int64]),
    "ab_step_pairforce": (C.c_int, [_H, C.POINTER(PairForceParams), C.c_int64]),
    "ab_step_pairforce_active": (C.c_int, [_H, C.POINTER(PairForceParams), C.c_int64, C.c_void_p, C.c_int64]),
    "ab_step_schelling": (C.c_int, [_H, C.POINTER(SchellingParams), C.c_int64, C.POINTER(C.c_int64)]),
    "ab_step_schelling_active": (C.c_int, [_H, C.POINTER(SchellingParams), C.c_int64, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ab_flock_last_counts": (C.c_int, [_H, C.c_void_p, C.c_int64]),
    "ab_reduce": (C.c_int, [_H, C.POINTER(ReduceSpec), C.c_int, C.c_void_p, C.c_void_p]),
    "ab_reduce_static": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                   C.c_double, C.c_void_p, C.c_void_p]),
    "ab_ids_by_property": (C.c_int, [_H, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64]),
    "ab_comm_unique_id": (C.c_int, [C.c_void_p]),
    "ab_comm_init": (C.c_int, [_H, C.c_int, C.c_int, C.c_void_p]),
    "ab_comm_link_local": (C.c_int, [C.POINTER(_H), C.c_int]),
    "ab_step_flock_group": (C.c_int, [C.POINTER(_H), C.c_int, C.POINTER(FlockParams), C.c_int64]),
    "ab_comm_set_capacity": (C.c_int, [_H, C.c_int64]),
    "ab_comm_ipc_export": (C.c_int, [_H, C.c_void_p]),
    "ab_comm_ipc_import": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "ab_index_canonical": (C.c_int, [_H]),
    "ab_comm_direct": (C.c_int, [_H]),
    "ab_comm_align": (C.c_int, [_H]),
    "ab_comm_exchange": (C.c_int, [_H]),
    "ab_comm_wait_stats": (C.c_int, [_H, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_int]),
    "ab_comm_slab": (C.c_int, [_H, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "ab_comm_halo": (C.c_int, [_H]),
    "ab_nlocal": (C.c_int64, [_H]),
    "ab_download_local": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ab_upload_local": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_void_p, C.c_int64]),
    "ab_launch_count": (C.c_int64, [_H]),
    "ab_last_step_ms": (C.c_double, [_H]),
    "ab_fp64_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "ab_set_profiling": (C.c_int, [_H, C.c_int]),
    "ab_kernel_times": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
}

_lib = None


def build(verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    subprocess.check_call(cmd, stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def declared_symbols():
    return sorted(_SIGS)
