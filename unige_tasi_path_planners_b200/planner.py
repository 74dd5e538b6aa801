"""Python mirror of the reference's planner interface over the C ABI (include/tasi_gpu.h).

Same method names, argument meaning and return codes as ReplannerBase
(ProjectToolkit/include/ReplannerBase.h:39-123): reset, step, set_occupancy_threshold,
set_heuristic_multiplier, set_map, patch_map(patch, x, y, w, h), set_start, set_goal, and the public fields
u_time, p_time, num_nodes_updated, num_nodes_expanded. The three classes correspond to
FieldDPlanner<O> (FieldDStar/FieldDPlanner.h), DFMPlanner<O> (DynamicFastMarching/DynamicFastMarching.h) and
ShiftedGridPlanner<O> (ShiftedGridFastMarching/ShiftedGridPlanner.h).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import FD, DFM, MFD, LOOP_OK, LOOP_FAILURE_NO_GRAPH, LOOP_FAILURE_NO_GOAL  # noqa: F401


class TasiError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("tasi error %d: %s" % (code, msg))
        self.code = code


class FieldView:
    """What the path extractor reads after a step: the converged value of every element (g == rhs).

    Stands where `const ExpandedMap<E,T>&` stands in the reference (get_expanded_map, ReplannerBase.h:117-119);
    get_g / get_rhs / get_interp_rhs follow ExpandedMap_impl.h:65-97.
    """

    def __init__(self, planner):
        self._p = planner

    def window(self, x0, y0, nx, ny):
        return self._p.read_window(x0, y0, nx, ny)

    def get_g(self, x, y):
        return float(self._p.read_window(int(x), int(y), 1, 1)[0, 0])

    get_rhs = get_g

    def get_interp_rhs(self, x, y):
        """Node value seen by the extractor: the node's own rhs (node planners) or the mean of the 4 cells
        around it (MS-DFM, ExpandedMap_impl.h:87-92)."""
        if self._p.kind != DFM:
            return self.get_g(x, y)
        w = self._p.read_window(int(x) - 1, int(y) - 1, 2, 2)
        return float((w[0, 0] + w[0, 1] + w[1, 0] + w[1, 1]) / np.float32(4.0))


class GpuPlanner:
    kind = FD
    max_level = 1

    def __init__(self, level=1, device=0):
        if not (0 <= level <= self.max_level):
            raise ValueError("optimisation level %d not defined for this planner" % level)
        self._L = _lib.lib()
        h = C.c_void_p()
        rc = self._L.tasi_create(self.kind, level, device, C.byref(h))
        if rc != _lib.OK:
            raise TasiError(rc, "tasi_create failed (no CUDA device? there is no CPU fallback)")
        self._h = h
        self.level = level
        self.device = device
        self.u_time = 0.0
        self.p_time = 0.0
        self.num_nodes_updated = 0
        self.num_nodes_expanded = 0
        self.heuristic_multiplier = 1.0
        self.width = self.length = 0

    # -- lifetime ------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._L.tasi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc < _lib.LOOP_FAILURE_NO_GOAL:
            raise TasiError(rc, self._L.tasi_last_error(self._h).decode())
        return rc

    # -- the reference's public surface -------------------------------------------------------------------------
    def reset(self):
        self._check(self._L.tasi_reset(self._h))

    def set_occupancy_threshold(self, threshold):
        self._check(self._L.tasi_set_occupancy_threshold(self._h, threshold))

    def set_heuristic_multiplier(self, mult):
        self.heuristic_multiplier = mult
        self._check(self._L.tasi_set_heuristic_multiplier(self._h, mult))

    def set_map(self, new_map, w=None, h=None):
        """new_map: uint8 array of h rows by w columns (numpy), or a CUDA torch tensor (stays on the device)."""
        if hasattr(new_map, "is_cuda") and new_map.is_cuda:
            assert new_map.dim() == 2 and new_map.stride(1) == 1 and new_map.element_size() == 1
            self.length, self.width = int(new_map.shape[0]), int(new_map.shape[1])
            rc = self._L.tasi_set_map_device(self._h, C.c_void_p(new_map.data_ptr()), self.width, self.length,
                                             int(new_map.stride(0)))
            return self._check(rc)
        a = np.ascontiguousarray(new_map, dtype=np.uint8)
        if w is not None and h is not None:
            a = a.reshape(h, w)
        self.length, self.width = a.shape
        return self._check(self._L.tasi_set_map(self._h, a.ctypes.data_as(C.c_void_p), self.width, self.length))

    def patch_map(self, patch, x, y, w=None, h=None):
        """Rows x..x+h-1, columns y..y+w-1 take `patch` (h rows of w) — Graph::update argument order."""
        a = np.ascontiguousarray(patch, dtype=np.uint8)
        if w is not None and h is not None:
            a = a.reshape(h, w)
        hh, ww = a.shape
        n = C.c_uint64(0)
        self._check(self._L.tasi_patch_map(self._h, a.ctypes.data_as(C.c_void_p), x, y, ww, hh, C.byref(n)))
        return n.value

    PATCH_STATIC_MAX_CELLS = 1 << 20  # TASI_PATCH_STATIC_MAX_CELLS of include/tasi_gpu.h

    def patch_map_static(self, patch, x, y, w=None, h=None):
        """patch_map through tasi_patch_map_static: the variant that allocates nothing (scratch in a static device
        array shared by every planner on the GPU, at most PATCH_STATIC_MAX_CELLS cells per patch; no reference
        counterpart)."""
        a = np.ascontiguousarray(patch, dtype=np.uint8)
        if w is not None and h is not None:
            a = a.reshape(h, w)
        hh, ww = a.shape
        n = C.c_uint64(0)
        self._check(self._L.tasi_patch_map_static(self._h, a.ctypes.data_as(C.c_void_p), x, y, ww, hh, C.byref(n)))
        return n.value

    def set_start(self, x, y=None):
        if y is None:
            x, y = x
        self._check(self._L.tasi_set_start(self._h, x, y))

    def set_goal(self, x, y=None):
        if y is None:
            x, y = x
        self._check(self._L.tasi_set_goal(self._h, x, y))

    def set_goals(self, goals):
        """Several goals (sequence of (x, y)): value 0 at each, the field is the cost to the nearest one."""
        a = np.ascontiguousarray(goals, dtype=np.float32).reshape(-1, 2)
        self._check(self._L.tasi_set_goals(self._h, a.ctypes.data_as(C.POINTER(C.c_float)), a.shape[0]))

    def step(self):
        u, p = C.c_float(0), C.c_float(0)
        nu, ne = C.c_uint64(0), C.c_uint64(0)
        rc = self._check(self._L.tasi_step(self._h, C.byref(u), C.byref(p), C.byref(nu), C.byref(ne)))
        if rc == LOOP_OK:
            self.u_time, self.p_time = u.value, p.value
            self.num_nodes_updated, self.num_nodes_expanded = nu.value, ne.value
        return rc

    def get_expanded_map(self):
        return FieldView(self)

    # -- field access -----------------------------------------------------------------------------------------
    def field_dims(self):
        ex, ey = C.c_int(0), C.c_int(0)
        self._check(self._L.tasi_field_dims(self._h, C.byref(ex), C.byref(ey)))
        return ex.value, ey.value

    def read_field(self):
        ex, ey = self.field_dims()
        out = np.empty((ex, ey), np.float32)
        self._check(self._L.tasi_read_field(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def read_window(self, x0, y0, nx, ny, out=None):
        if out is None:
            out = np.empty((nx, ny), np.float32)
        self._check(self._L.tasi_read_window(self._h, x0, y0, nx, ny, out.ctypes.data_as(C.c_void_p)))
        return out

    def read_map(self):
        out = np.empty((self.length, self.width), np.uint8)
        self._check(self._L.tasi_read_map(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def synth_map(self, seed, width, length, row0=0, p_obs=10, start=None, goal=None):
        """Generate the synthetic map of synth.py on the device; start/goal are GLOBAL cell coordinates."""
        sx, sy = (int(round(start[0])), int(round(start[1]))) if start is not None else (-1, -1)
        gx, gy = (int(round(goal[0])), int(round(goal[1]))) if goal is not None else (-1, -1)
        self.length, self.width = length, width
        self._check(self._L.tasi_synth_map(self._h, seed, width, length, row0, p_obs, sx, sy, gx, gy))

    def set_solver_blocks(self, max_blocks):
        """Cap the persistent solver CTAs (0 = whole device): lets several planners run side by side."""
        self._check(self._L.tasi_set_solver_blocks(self._h, int(max_blocks)))

    def stats(self):
        s = _lib.Stats()
        self._check(self._L.tasi_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    def field_device(self):
        ptr, pitch = C.c_void_p(), C.c_int64(0)
        self._check(self._L.tasi_field_device(self._h, C.byref(ptr), C.byref(pitch)))
        return ptr.value, pitch.value

    # -- row bands ---------------------------------------------------------------------------------------------
    def band_config(self, global_row0, open_top, open_bottom):
        self._check(self._L.tasi_band_config(self._h, global_row0, int(open_top), int(open_bottom)))

    def band_rows(self, side):
        s, r, n = C.c_void_p(), C.c_void_p(), C.c_int64(0)
        self._check(self._L.tasi_band_rows(self._h, side, C.byref(s), C.byref(r), C.byref(n)))
        return s.value, r.value, n.value

    def band_import(self, side):
        n = C.c_uint64(0)
        self._check(self._L.tasi_band_import(self._h, side, C.byref(n)))
        return n.value

    def band_begin(self):
        return self._check(self._L.tasi_band_begin(self._h))

    def band_solve(self):
        n = C.c_uint64(0)
        self._check(self._L.tasi_band_solve(self._h, C.byref(n)))
        return n.value

    def band_solve_until(self, limit):
        """Solve without letting the schedule's threshold pass `limit`; returns the lowest parked value (inf: none)."""
        n, k = C.c_uint64(0), C.c_float(0)
        self._check(self._L.tasi_band_solve_until(self._h, limit, C.byref(n), C.byref(k)))
        return k.value

    def band_delta(self):
        d = C.c_float(0)
        self._check(self._L.tasi_band_delta(self._h, C.byref(d)))
        return d.value

    def set_map_stack(self, maps, index):
        """A stack of independent maps (equal shapes), block b showing maps[index[b]], an obstacle row between blocks
        (tasi_set_map_stack): each distinct map is uploaded once and replicated on the device."""
        arrs = [np.ascontiguousarray(m, np.uint8) for m in maps]
        L, W = arrs[0].shape
        assert all(a.shape == (L, W) for a in arrs)
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        idx = np.ascontiguousarray(index, np.int32)
        self.length, self.width = len(idx) * (L + 1) - 1, W
        self._check(self._L.tasi_set_map_stack(self._h, ptrs, len(arrs), idx.ctypes.data_as(C.POINTER(C.c_int32)),
                                               len(idx), W, L))

    def read_parents(self, x0=0, y0=0, nx=None, ny=None):
        """Parent (back-pointer) field of the window: (nx, ny, 4) int8 offsets (dx1, dy1, dx2, dy2) of the two parents of
        every element, (0, 0) = absent — what the reference's level-1 planners keep as INFO (tasi_read_parents)."""
        ex, ey = self.field_dims()
        nx = ex - x0 if nx is None else nx
        ny = ey - y0 if ny is None else ny
        out = np.zeros((nx, ny, 4), np.int8)
        self._check(self._L.tasi_read_parents(self._h, x0, y0, nx, ny, out.ctypes.data_as(C.c_void_p)))
        return out

    def field_checksum(self):
        """Order-independent 64-bit checksum of the value field (tasi_field_checksum): equal for two planners iff their
        fields are bit-identical; on a multi-process sharded planner every rank calls it and gets the whole field's."""
        v = C.c_uint64(0)
        self._check(self._L.tasi_field_checksum(self._h, C.byref(v)))
        return v.value

    def set_focus(self, enabled=True, slack=-1.0):
        """Plans stop once the start is settled (the reference's end_condition); later work stays parked and is
        resumed by the next step(). slack < 0 = default."""
        self._check(self._L.tasi_set_focus(self._h, int(bool(enabled)), float(slack)))

    def extract_paths(self, starts=None, goals=None, max_steps=20, lookahead=True, allow_indirect=True,
                      cap_points=None, with_costs=True, row_ranges=None):
        """n walks over the converged field in one kernel launch (tasi_extract_paths). starts / goals: n (x, y)
        pairs, None = the planner's own start and goal (one walk). row_ranges: n (first cell row, cell rows) pairs —
        each walk sees that block of rows as its whole map, positions relative to the block. -> list of dicts {path (k, 2), costs,
        total_cost, total_dist, status}, plus the wall clock of the call in ms."""
        f32p, i32p = C.POINTER(C.c_float), C.POINTER(C.c_int32)
        if starts is None:
            n, sp, gp = 1, None, None
        else:
            st = np.ascontiguousarray(starts, np.float32).reshape(-1, 2)
            gl = np.ascontiguousarray(goals, np.float32).reshape(-1, 2)
            assert st.shape == gl.shape
            n, sp, gp = st.shape[0], st.ctypes.data_as(f32p), gl.ctypes.data_as(f32p)
        rr = None
        if row_ranges is not None:
            rr_a = np.ascontiguousarray(row_ranges, np.int32).reshape(-1, 2)
            assert rr_a.shape[0] == n
            rr = rr_a.ctypes.data_as(i32p)
        cap = int(cap_points) if cap_points else 1 + 3 * max_steps
        capc = 2 * max_steps if not cap_points else cap
        xy = np.empty((n, cap, 2), np.float32)
        sc = np.empty((n, capc), np.float32) if with_costs else None
        npts, ncst, status = (np.zeros(n, np.int32) for _ in range(3))
        totals = np.zeros((n, 2), np.float32)
        ems = C.c_float(0)
        self._check(self._L.tasi_extract_paths(
            self._h, sp, gp, rr, n, max_steps, int(lookahead), int(allow_indirect), xy.ctypes.data_as(f32p), cap,
            npts.ctypes.data_as(i32p), sc.ctypes.data_as(f32p) if with_costs else None, capc,
            ncst.ctypes.data_as(i32p), totals.ctypes.data_as(f32p), status.ctypes.data_as(i32p), C.byref(ems)))
        out = [{"path": xy[i, :npts[i]].copy(), "costs": sc[i, :ncst[i]].copy() if with_costs else None,
                "total_cost": float(totals[i, 0]), "total_dist": float(totals[i, 1]), "status": int(status[i])}
               for i in range(n)]
        return out, ems.value


class PathExtractor:
    """LinearInterpolationPathExtractor (PathExtraction/LinearInterpolationPathExtractor.h:9-22) over a GPU
    planner: same public members, the walk runs on the device field (csrc/tasi_extract.cuh)."""

    def __init__(self, planner):
        self.planner = planner
        self.path_ = np.zeros((0, 2), np.float32)
        self.cost_ = np.zeros(0, np.float32)
        self.total_cost = 0.0
        self.total_dist = 0.0
        self.lookahead = True
        self.max_steps = 20
        self.e_time = 0.0
        self.allow_indirect_traversals = True
        self.status = None

    def extract_path(self):
        res, self.e_time = self.planner.extract_paths(None, None, self.max_steps, self.lookahead,
                                                      self.allow_indirect_traversals)
        r = res[0]
        self.path_, self.cost_ = r["path"], r["costs"]
        self.total_cost, self.total_dist, self.status = r["total_cost"], r["total_dist"], r["status"]


class FieldDPlanner(GpuPlanner):
    """FD* — FieldDPlanner<OPT_LVL> (FieldDStar/FieldDPlanner.h); values on nodes."""
    kind = FD
    max_level = 1


class DFMPlanner(GpuPlanner):
    """MS-DFM — DFMPlanner<OPT_LVL> (DynamicFastMarching/DynamicFastMarching.h); values on cells."""
    kind = DFM
    max_level = 1


class ShiftedGridPlanner(GpuPlanner):
    """MFD* — ShiftedGridPlanner<OPT_LVL> (ShiftedGridFastMarching/ShiftedGridPlanner.h); values on nodes."""
    kind = MFD
    max_level = 2


PLANNERS = {FD: FieldDPlanner, DFM: DFMPlanner, MFD: ShiftedGridPlanner}


class ShardedPlanner(GpuPlanner):
    """One grid over several GPUs behind the same interface (tasi_create_sharded / tasi_create_sharded_rank): the
    planner `kind` with its element rows cut into bands dealt out cyclically over the devices. Same methods as the
    single-device classes; the converged field is bit-identical to theirs.

    devices: the devices this process drives (a device may repeat: several ranks on one GPU), or
    rank / n_ranks / unique_id: this process is rank `rank` of `n_ranks` on `device` (one process per GPU; rank 0 gets
    unique_id from ShardedPlanner.nccl_unique_id() and the caller broadcasts it)."""

    def __init__(self, kind, level=None, devices=(0,), rank=None, n_ranks=None, unique_id=None, device=0,
                 bands_per_device=None, phases_per_exchange=None):
        self.kind = kind
        self.max_level = PLANNERS[kind].max_level
        level = self.max_level if level is None else level
        if not (0 <= level <= self.max_level):
            raise ValueError("optimisation level %d not defined for this planner" % level)
        self._L = _lib.lib()
        h = C.c_void_p()
        if rank is None:
            ids = np.ascontiguousarray(devices, np.int32)
            rc = self._L.tasi_create_sharded(kind, level, ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids), C.byref(h))
            self.device = int(ids[0])
        else:
            buf = C.create_string_buffer(bytes(unique_id), 128) if unique_id is not None else None
            rc = self._L.tasi_create_sharded_rank(kind, level, device, rank, n_ranks, buf, C.byref(h))
            self.device = device
        if rc != _lib.OK:
            raise TasiError(rc, "tasi_create_sharded failed: %s (there is no CPU fallback)"
                            % self._L.tasi_last_error(None).decode())
        self._h = h
        self.level = level
        self.u_time = self.p_time = 0.0
        self.num_nodes_updated = self.num_nodes_expanded = 0
        self.heuristic_multiplier = 1.0
        self.width = self.length = 0
        if bands_per_device or phases_per_exchange:
            self._check(self._L.tasi_shard_config(self._h, int(bands_per_device or 0), int(phases_per_exchange or 0)))

    @staticmethod
    def nccl_unique_id():
        buf = C.create_string_buffer(128)
        rc = _lib.lib().tasi_nccl_unique_id(buf)
        if rc != _lib.OK:
            raise TasiError(rc, "tasi_nccl_unique_id failed (libnccl.so.2 not found?)")
        return buf.raw

    def set_fused(self, enabled=True):
        """Fused mode (default on): boundary rows and activations go straight into the neighbour GPU's memory inside the
        tile visits; off = exchange iterations (NCCL or peer reads)."""
        self._check(self._L.tasi_shard_set_fused(self._h, int(bool(enabled))))

    def connect_peers(self, dist, device):
        """Multi-process form: all-gather the CUDA IPC handles of every rank (torch.distributed, any backend) and map the
        neighbours' memory — after every set_map / synth_map. Without it the planner runs the NCCL iterations."""
        import torch
        n = self._L.tasi_shard_ipc_size()
        mine = C.create_string_buffer(n)
        self._check(self._L.tasi_shard_ipc_export(self._h, mine))
        world = dist.get_world_size()
        t = torch.frombuffer(bytearray(mine.raw), dtype=torch.uint8).to(device)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        blob = b"".join(bytes(o.cpu().numpy().tobytes()) for o in out)
        self._check(self._L.tasi_shard_ipc_import(self._h, C.create_string_buffer(blob, len(blob))))

    def shard_stats(self):
        s = _lib.ShardStats()
        self._check(self._L.tasi_shard_get_stats(self._h, C.byref(s)))
        return s.as_dict()
