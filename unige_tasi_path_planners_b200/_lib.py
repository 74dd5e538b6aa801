"""ctypes binding of the C ABI declared in include/tasi_gpu.h (libtasi_gpu.so, built by build.py).

There is no fallback: if the library is missing, or no CUDA device is present when a planner is created, the
call raises.
"""
import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TASI_GPU_LIB") or os.path.join(PKG, "libtasi_gpu.so")  # (override: development builds)

OK = 0
LOOP_OK = 0
LOOP_FAILURE_NO_GRAPH = -1
LOOP_FAILURE_NO_GOAL = -2
ERR_ARG, ERR_CUDA, ERR_TIMEOUT, ERR_STATE = -3, -4, -5, -6
FD, DFM, MFD = 0, 1, 2

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_u64p = C.POINTER(C.c_uint64)
_i64p = C.POINTER(C.c_int64)
_ip = C.POINTER(C.c_int)


class Stats(C.Structure):
    _fields_ = [("tiles", C.c_uint64), ("tile_visits", C.c_uint64), ("tile_rounds", C.c_uint64), ("elem_evals", C.c_uint64),
                ("elem_lowered", C.c_uint64), ("queue_pushes", C.c_uint64), ("inval_visits", C.c_uint64),
                ("elem_invalidated", C.c_uint64), ("seed_elems", C.c_uint64), ("phases", C.c_uint64),
                ("far_parked", C.c_uint64), ("kernel_ms_solve", C.c_float),
                ("kernel_ms_inval", C.c_float), ("kernel_ms_init", C.c_float), ("step_ms_device", C.c_float),
                ("solver_blocks", C.c_int32),
                ("solver_launches", C.c_int32),
                ("cyc_pop", C.c_uint64), ("cyc_stage", C.c_uint64), ("cyc_relax", C.c_uint64),
                ("cyc_post", C.c_uint64), ("cyc_advance", C.c_uint64),
                ("settled_below", C.c_float), ("focus_hit", C.c_int)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class ShardStats(C.Structure):
    _fields_ = [("n_ranks", C.c_int32), ("bands", C.c_int32), ("bands_per_rank", C.c_int32), ("rows_per_band", C.c_int32),
                ("nccl", C.c_int32), ("fused", C.c_int32), ("exchanges", C.c_uint64), ("inval_exchanges", C.c_uint64),
                ("fused_launches", C.c_uint64), ("remote_activations", C.c_uint64),
                ("wall_ms", C.c_float), ("solver_ms_max_rank", C.c_float), ("solver_ms_min_rank", C.c_float),
                ("busy_frac_max_rank", C.c_float), ("busy_frac_min_rank", C.c_float), ("pad", C.c_float)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "pad"}


# every symbol include/tasi_gpu.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "tasi_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "tasi_destroy": (C.c_int, [C.c_void_p]),
    "tasi_reset": (C.c_int, [C.c_void_p]),
    "tasi_set_occupancy_threshold": (C.c_int, [C.c_void_p, C.c_float]),
    "tasi_set_heuristic_multiplier": (C.c_int, [C.c_void_p, C.c_float]),
    "tasi_set_map": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "tasi_set_map_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64]),
    "tasi_patch_map": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _u64p]),
    "tasi_patch_map_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _u64p]),
    "tasi_set_start": (C.c_int, [C.c_void_p, C.c_float, C.c_float]),
    "tasi_set_goal": (C.c_int, [C.c_void_p, C.c_float, C.c_float]),
    "tasi_set_goals": (C.c_int, [C.c_void_p, _f32p, C.c_int]),
    "tasi_step": (C.c_int, [C.c_void_p, _f32p, _f32p, _u64p, _u64p]),
    "tasi_field_dims": (C.c_int, [C.c_void_p, _ip, _ip]),
    "tasi_read_window": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "tasi_read_field": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tasi_field_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), _i64p]),
    "tasi_cost_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), _i64p]),
    "tasi_synth_map": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int,
                                 C.c_int, C.c_int]),
    "tasi_read_map": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tasi_set_solver_blocks": (C.c_int, [C.c_void_p, C.c_int]),
    "tasi_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "tasi_last_error": (C.c_char_p, [C.c_void_p]),
    "tasi_band_config": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int]),
    "tasi_band_rows": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), _i64p]),
    "tasi_band_import": (C.c_int, [C.c_void_p, C.c_int, _u64p]),
    "tasi_band_begin": (C.c_int, [C.c_void_p]),
    "tasi_band_solve": (C.c_int, [C.c_void_p, _u64p]),
    "tasi_band_solve_until": (C.c_int, [C.c_void_p, C.c_float, _u64p, _f32p]),
    "tasi_band_delta": (C.c_int, [C.c_void_p, _f32p]),
    "tasi_read_parents": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "tasi_set_map_stack": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, _i32p, C.c_int, C.c_int, C.c_int]),
    "tasi_set_focus": (C.c_int, [C.c_void_p, C.c_int, C.c_float]),
    "tasi_create_sharded": (C.c_int, [C.c_int, C.c_int, _i32p, C.c_int, C.POINTER(C.c_void_p)]),
    "tasi_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "tasi_create_sharded_rank": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "tasi_shard_config": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "tasi_shard_set_fused": (C.c_int, [C.c_void_p, C.c_int]),
    "tasi_shard_ipc_size": (C.c_int, []),
    "tasi_shard_ipc_export": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tasi_shard_ipc_import": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tasi_shard_get_stats": (C.c_int, [C.c_void_p, C.POINTER(ShardStats)]),
    "tasi_field_checksum": (C.c_int, [C.c_void_p, _u64p]),
    "tasi_extract_paths": (C.c_int, [C.c_void_p, _f32p, _f32p, _i32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, C.c_int,
                                     _i32p, _f32p, C.c_int, _i32p, _f32p, _i32p, _f32p]),
}

_lib = None


def lib():
    """The loaded CUDA library. Raises if it has not been built (python -m unige_tasi_path_planners_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "CUDA extension missing: %s (build it with `python -m unige_tasi_path_planners_b200.build`); "
                "there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
