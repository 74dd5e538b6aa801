"""TEST INFRASTRUCTURE: the product's own __host__ __device__ headers (csrc/tasi_update.cuh, tasi_extract.cuh,
tasi_parents.cuh) compiled for the host (csrc/tasi_hostops.cpp -> libtasi_hostops.so, g++ -ffp-contract=off), so that
CPU tests can compare them value by value with the oracle / the reference. Nothing in the package imports this: the
product path is the CUDA library and has no CPU route.
"""
import ctypes as C
import os

import numpy as np

from unige_tasi_path_planners_b200 import build

HOSTOPS_PATH = os.path.join(os.path.dirname(os.path.abspath(build.__file__)), "libtasi_hostops.so")
_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)

_hostops = None


def hostops():
    """The host build of the operators, the path walk and the parent field (built on demand)."""
    global _hostops
    if _hostops is None:
        build.build_hostops()
        H = C.CDLL(HOSTOPS_PATH)
        H.tasi_host_fd_tri.restype = C.c_float
        H.tasi_host_fd_tri.argtypes = [C.c_float] * 4
        H.tasi_host_mfd_tri.restype = C.c_float
        H.tasi_host_mfd_tri.argtypes = [C.c_float] * 3
        H.tasi_host_dfm_stencil.restype = C.c_float
        H.tasi_host_dfm_stencil.argtypes = [C.c_float] * 4
        H.tasi_host_update.restype = C.c_float
        H.tasi_host_update.argtypes = [C.c_int, _f32p, _f32p]
        H.tasi_host_fd_tri_n.argtypes = [_f32p] * 5 + [C.c_long]
        H.tasi_host_mfd_tri_n.argtypes = [_f32p] * 4 + [C.c_long]
        H.tasi_host_dfm_stencil_n.argtypes = [_f32p] * 5 + [C.c_long]
        H.tasi_host_extract.restype = C.c_int
        H.tasi_host_extract.argtypes = [C.c_int, _f32p, _u8p, C.c_int, C.c_int, C.c_int] + [C.c_float] * 4 + \
            [C.c_int] * 3 + [_f32p, C.c_int, _f32p, C.c_int, _i32p, _i32p, _f32p, _f32p]
        H.tasi_host_node_triangles.argtypes = [C.c_int, _f32p, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        H.tasi_host_parents.argtypes = [C.c_int, _f32p, _u8p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        _hostops = H
    return _hostops


def host_node_triangles(kind, field, cmap, x, y, thr=255):
    """The 8 triangle values of node (x, y) in the order of Graph::neighbors_8 (tests only)."""
    field = np.ascontiguousarray(field, np.float32)
    cmap = np.ascontiguousarray(cmap, np.uint8)
    out = np.zeros(8, np.float32)
    hostops().tasi_host_node_triangles(int(kind), field.ctypes.data_as(_f32p), cmap.ctypes.data_as(_u8p), cmap.shape[0],
                                       cmap.shape[1], thr, int(x), int(y), out.ctypes.data_as(_f32p))
    return out


def host_parents(kind, field, cmap, thr=255):
    """csrc/tasi_parents.cuh compiled for the host (tests only): (ex, ey, 4) int8 offsets of the two parents."""
    field = np.ascontiguousarray(field, np.float32)
    cmap = np.ascontiguousarray(cmap, np.uint8)
    L, W = cmap.shape
    out = np.zeros(field.shape + (4,), np.int8)
    hostops().tasi_host_parents(int(kind), field.ctypes.data_as(_f32p), cmap.ctypes.data_as(_u8p), L, W, thr,
                                out.ctypes.data_as(C.c_void_p))
    return out


def host_extract(cells, field, cmap, start, goal, thr=255, max_steps=20, lookahead=True, allow_indirect=True,
                 cap=1 << 16):
    """csrc/tasi_extract.cuh compiled for the host, over numpy arrays (tests only). -> (xy, piece costs, total
    cost, total dist, status)."""
    field = np.ascontiguousarray(field, np.float32)
    cmap = np.ascontiguousarray(cmap, np.uint8)
    L, W = cmap.shape
    xy = np.zeros((cap, 2), np.float32)
    costs = np.zeros(cap, np.float32)
    n = np.zeros(2, np.int32)
    tot = np.zeros(2, np.float32)
    rc = hostops().tasi_host_extract(
        int(cells), field.ctypes.data_as(_f32p), cmap.ctypes.data_as(_u8p), L, W, thr, start[0], start[1], goal[0],
        goal[1], max_steps, int(lookahead), int(allow_indirect), xy.ctypes.data_as(_f32p), cap,
        costs.ctypes.data_as(_f32p), cap, n[0:].ctypes.data_as(_i32p), n[1:].ctypes.data_as(_i32p),
        tot[0:].ctypes.data_as(_f32p), tot[1:].ctypes.data_as(_f32p))
    return xy[:n[0]].copy(), costs[:n[1]].copy(), float(tot[0]), float(tot[1]), rc
