"""TEST INFRASTRUCTURE — ctypes loaders for the CPU oracle (oracle/libtasi_oracle.so) and for the
unmodified reference planners behind oracle/ref_harness.cpp (oracle/_ref/libref_{fd,dfm,sg}.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this module. The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
FD, DFM, MFD = 0, 1, 2
KIND_NAMES = {FD: "fd", DFM: "dfm", MFD: "mfd"}
_REF_LIB = {FD: "libref_fd.so", DFM: "libref_dfm.so", MFD: "libref_sg.so"}

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)


def build(ref=True):
    """(Re)build the oracle library and, when the reference sources are present, oracle/_ref."""
    targets = ["oracle"] + (["ref"] if ref else [])
    subprocess.check_call(["make", "-s", "-C", HERE, "-j4"] + targets)


def _p(a, t):
    return a.ctypes.data_as(t)


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "libtasi_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = C.CDLL(path)
        L.tasi_oracle_fd_tri.restype = C.c_float
        L.tasi_oracle_fd_tri.argtypes = [C.c_float] * 4
        L.tasi_oracle_mfd_tri.restype = C.c_float
        L.tasi_oracle_mfd_tri.argtypes = [C.c_float] * 3
        L.tasi_oracle_dfm_stencil.restype = C.c_float
        L.tasi_oracle_dfm_stencil.argtypes = [C.c_float] * 4
        L.tasi_oracle_update.restype = C.c_float
        L.tasi_oracle_update.argtypes = [C.c_int, _u8p, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, C.c_int]
        L.tasi_oracle_solve_sweep.restype = C.c_int
        L.tasi_oracle_solve_sweep.argtypes = [C.c_int, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p,
                                              C.c_int]
        L.tasi_oracle_solve_heap.restype = C.c_uint64
        L.tasi_oracle_solve_heap.argtypes = [C.c_int, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.tasi_oracle_residual.restype = None
        L.tasi_oracle_residual.argtypes = [C.c_int, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p,
                                           C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), _f32p]
        L.tasi_oracle_patch.restype = C.c_uint64
        L.tasi_oracle_patch.argtypes = [_u8p, C.c_int, C.c_int, _u8p, C.c_int, C.c_int, C.c_int, C.c_int]
        _lib = L
    return _lib


def field_shape(kind, L, W):
    return (L, W) if kind == DFM else (L + 1, W + 1)


def goal_elem(gx, gy):
    """Node(Position)/Cell(Position) round with roundf (Node.cpp:14-17, Cell.cpp:20-21)."""
    r = lambda v: int(np.floor(abs(v) + 0.5) * (1 if v >= 0 else -1))
    return r(gx), r(gy)


def solve(kind, cmap, goal, thr=255, method="heap"):
    """Converged field of planner `kind` on uint8 map `cmap` (L x W) with goal position (x, y)."""
    cmap = np.ascontiguousarray(cmap, dtype=np.uint8)
    L, W = cmap.shape
    gx, gy = goal_elem(*goal)
    g = np.empty(field_shape(kind, L, W), dtype=np.float32)
    if method == "heap":
        lib().tasi_oracle_solve_heap(kind, _p(cmap, _u8p), W, L, thr, gx, gy, _p(g, _f32p))
    else:
        lib().tasi_oracle_solve_sweep(kind, _p(cmap, _u8p), W, L, thr, gx, gy, _p(g, _f32p), 1 << 20)
    return g


def residual(kind, cmap, goal, g, thr=255):
    cmap = np.ascontiguousarray(cmap, dtype=np.uint8)
    g = np.ascontiguousarray(g, dtype=np.float32)
    L, W = cmap.shape
    gx, gy = goal_elem(*goal)
    lo, hi, mr = C.c_uint64(0), C.c_uint64(0), C.c_float(0)
    lib().tasi_oracle_residual(kind, _p(cmap, _u8p), W, L, thr, gx, gy, _p(g, _f32p), C.byref(lo), C.byref(hi),
                               C.byref(mr))
    return lo.value, hi.value, mr.value


def patch(cmap, patch_bytes, x, y):
    """In-place Graph::update; returns the number of changed bytes."""
    patch_bytes = np.ascontiguousarray(patch_bytes, dtype=np.uint8)
    h, w = patch_bytes.shape
    L, W = cmap.shape
    return lib().tasi_oracle_patch(_p(cmap, _u8p), W, L, _p(patch_bytes, _u8p), x, y, w, h)


# ------------------------------------------------------------------------------------------------
# the unmodified reference (oracle/_ref)
# ------------------------------------------------------------------------------------------------
_ref_libs = {}


def ref_available():
    return all(os.path.exists(os.path.join(HERE, "_ref", n)) for n in _REF_LIB.values())


def ref_fast_available(kind=DFM):
    """The timing build (-O2 -march=x86-64-v3 -DNDEBUG, oracle/Makefile) exists and this CPU can run it."""
    if not os.path.exists(os.path.join(HERE, "_ref", _REF_LIB[kind].replace(".so", "_fast.so"))):
        return False
    try:
        with open("/proc/cpuinfo") as f:
            flags = next((l for l in f if l.startswith("flags")), "")
    except OSError:
        return False
    fl = set(flags.split())
    return {"avx2", "fma", "bmi2", "movbe"} <= fl


def _ref_lib(kind, fast=False):
    key = (kind, bool(fast))
    if key in _ref_libs:
        return _ref_libs[key]
    name = _REF_LIB[kind].replace(".so", "_fast.so") if fast else _REF_LIB[kind]
    R = C.CDLL(os.path.join(HERE, "_ref", name))
    R.ref_create.restype = C.c_void_p
    R.ref_create.argtypes = [C.c_int]
    R.ref_destroy.argtypes = [C.c_void_p]
    R.ref_set_map.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int, C.c_float]
    R.ref_patch_map.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int, C.c_int, C.c_int]
    R.ref_set_start.argtypes = [C.c_void_p, C.c_float, C.c_float]
    R.ref_set_goal.argtypes = [C.c_void_p, C.c_float, C.c_float]
    R.ref_reset.argtypes = [C.c_void_p]
    R.ref_step.restype = C.c_int
    R.ref_step.argtypes = [C.c_void_p]
    R.ref_times.argtypes = [C.c_void_p, _f32p, _f32p]
    R.ref_counters.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    R.ref_map_size.restype = C.c_uint64
    R.ref_map_size.argtypes = [C.c_void_p]
    R.ref_top_key.restype = C.c_float
    R.ref_top_key.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    R.ref_dump.argtypes = [C.c_void_p, _f32p, _f32p]
    R.ref_dump_info.restype = C.c_int
    R.ref_dump_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
    R.ref_step_costs.restype = C.c_int
    R.ref_step_costs.argtypes = [C.c_void_p, _f32p, C.c_int]
    R.ref_extract.restype = C.c_int
    R.ref_extract.argtypes = [C.c_void_p, C.c_int, _f32p, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, _f32p,
                              _f32p, _f32p]
    R.ref_op_tri.restype = C.c_float
    R.ref_op_tri.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_float] * 2
    R.ref_op_stencil.restype = C.c_float
    R.ref_op_stencil.argtypes = [C.c_void_p] + [C.c_float] * 4
    _ref_libs[key] = R
    return _ref_libs[key]


class RefPlanner:
    """The reference planner `kind` (FD/DFM/MFD) at optimisation `level`, driven through its own public
    API (ReplannerBase.h:39-123)."""

    def __init__(self, kind, level=1, fast=False):
        """fast: the timing build (never for parity: its float contraction is the compiler's choice)."""
        self.kind, self.level = kind, level
        self.R = _ref_lib(kind, fast)
        self.h = C.c_void_p(self.R.ref_create(level))
        if not self.h:
            raise ValueError("bad level %d for kind %d" % (level, kind))
        self.L = self.W = 0

    def close(self):
        if self.h:
            self.R.ref_destroy(self.h)
            self.h = None

    __del__ = close

    def set_map(self, cmap, occ=1.0):
        cmap = np.ascontiguousarray(cmap, dtype=np.uint8)
        self.L, self.W = cmap.shape
        self.R.ref_set_map(self.h, _p(cmap, _u8p), self.W, self.L, occ)

    def patch_map(self, pb, x, y):
        pb = np.ascontiguousarray(pb, dtype=np.uint8)
        h, w = pb.shape
        self.R.ref_patch_map(self.h, _p(pb, _u8p), x, y, w, h)

    def set_start(self, x, y):
        self.R.ref_set_start(self.h, x, y)

    def set_goal(self, x, y):
        self.R.ref_set_goal(self.h, x, y)

    def reset(self):
        self.R.ref_reset(self.h)

    def step(self):
        return self.R.ref_step(self.h)

    def times(self):
        u, p = C.c_float(0), C.c_float(0)
        self.R.ref_times(self.h, C.byref(u), C.byref(p))
        return u.value, p.value

    def counters(self):
        u, e = C.c_uint64(0), C.c_uint64(0)
        self.R.ref_counters(self.h, C.byref(u), C.byref(e))
        return u.value, e.value

    def map_size(self):
        return self.R.ref_map_size(self.h)

    def top_key(self):
        e = C.c_int(0)
        k = self.R.ref_top_key(self.h, C.byref(e))
        return float("inf") if e.value else k

    def dump(self):
        shape = field_shape(self.kind, self.L, self.W)
        g = np.empty(shape, np.float32)
        rhs = np.empty(shape, np.float32)
        self.R.ref_dump(self.h, _p(g, _f32p), _p(rhs, _f32p))
        return g, rhs

    def op_tri(self, n, pa, pb, ga, gb):
        """compute_optimal_cost(n, p_a, p_b, g_a, g_b) of the FD*/MFD* planner on its current map."""
        return self.R.ref_op_tri(self.h, n[0], n[1], pa[0], pa[1], pb[0], pb[1], ga, gb)

    def op_stencil(self, ga, gb, tau, h):
        """Value part of DFMPlanner::compute_optimal_cost."""
        return self.R.ref_op_stencil(self.h, ga, gb, tau, h)

    def closed_mask(self):
        """Elements the reference settled: g == rhs < INF and g <= top key."""
        g, rhs = self.dump()
        return (g == rhs) & np.isfinite(g) & (g <= self.top_key()), g

    def extract(self, field=None, max_steps=20, lookahead=True, allow_indirect=True, cap=1 << 16):
        xy = np.zeros((cap, 2), np.float32)
        cost, dist, ems = C.c_float(0), C.c_float(0), C.c_float(0)
        if field is not None:
            field = np.ascontiguousarray(field, dtype=np.float32)
            fp = _p(field, _f32p)
        else:
            fp = None
        n = self.R.ref_extract(self.h, 1 if field is not None else 0, fp, max_steps, int(lookahead),
                               int(allow_indirect), _p(xy, _f32p), cap, C.byref(cost), C.byref(dist),
                               C.byref(ems))
        return xy[:min(n, cap)].copy(), cost.value, dist.value, ems.value

    def dump_info(self):
        """INFO (level-1 back-pointers) of every expanded element: (ex, ey, 4) int32, INT32_MIN = absent.
        FD*/MFD*: [..., :2] = the node b of the triangle (b, ccw(b)); MS-DFM: the two parent cells."""
        shape = field_shape(self.kind, self.L, self.W)
        out = np.empty(shape + (4,), np.int32)
        have = self.R.ref_dump_info(self.h, out.ctypes.data_as(C.POINTER(C.c_int32)))
        return out if have else None

    def step_costs(self, cap=1 << 16):
        """cost_ of the last extract(): the cost of every piece of the path."""
        out = np.zeros(cap, np.float32)
        n = self.R.ref_step_costs(self.h, _p(out, _f32p), cap)
        return out[:min(n, cap)].copy()
