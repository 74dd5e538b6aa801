"""Row-band sharding of one large grid, driven from Python over the tasi_band_* primitives.

The PRODUCT path for a sharded grid is the sharded planner inside libtasi_gpu.so (tasi_create_sharded /
tasi_create_sharded_rank, csrc/tasi_shard.cuh; planner.ShardedPlanner; bench record in sharded_bench.py). This module is
the protocol it grew out of, kept because it states the exchange protocol in a form that runs on CPU tensors with gloo
(tests/test_bands.py, world size 2 and 3) and on one GPU over the band primitives of the C ABI.

The reference has no multi-device path; this is the sharded form of the same fixed-point solve.
Band r owns a contiguous range of element rows plus one halo row of values on each open side. One outer
iteration = every band drains its own tile queue (tasi_band_solve), neighbours swap their boundary rows
(point-to-point send/recv: NCCL over NVLink on GPUs), each band merges the received row into its halo
(tasi_band_import, which re-activates the border tiles whose halo got lower), and an all-reduce of the number of
lowered halo values decides global convergence (0 everywhere = every band is at its fixed point with halos that
agree with their owners, i.e. the single-grid fixed point, bit for bit).

`run_bands` is the whole protocol and only needs a band object and a communicator, so the same code runs
  * one process per GPU under torch.distributed (GpuBand + DistComm, backend nccl),
  * all bands in one process on one GPU (GpuBand + LocalComm) — the single-GPU test of the band C ABI,
  * on CPU tensors with the gloo backend (tests/ supplies a CPU band) — the world_size-2 test of the host logic.
"""
import numpy as np

from . import planner as _planner


def split_rows(n_rows, parts):
    """Contiguous, near-equal split of element rows [0, n_rows) into `parts` ranges [(a, b), ...]."""
    base, rem = divmod(n_rows, parts)
    out, a = [], 0
    for r in range(parts):
        b = a + base + (1 if r < rem else 0)
        out.append((a, b))
        a = b
    return out


def band_geometry(kind, length, parts, r):
    """Element rows [a, b) owned by band r of `parts`, and the cell rows [c0, c1) it must store.

    Node planners (FD*, MFD*): the (length+1) node rows are split; updating node row x reads cell rows x-1 and x,
    so an open top adds cell row a-1 and an open bottom keeps cell row b-1 (the cells under the last owned row).
    MS-DFM: cell rows are split and a cell's update reads only its own cost."""
    node = kind != _planner.DFM
    n_elem_rows = length + 1 if node else length
    a, b = split_rows(n_elem_rows, parts)[r]
    open_top, open_bottom = r > 0, r < parts - 1
    if node:
        c0 = a - 1 if open_top else 0
        c1 = b if open_bottom else length
    else:
        c0, c1 = a, b
    return dict(a=a, b=b, c0=c0, c1=c1, open_top=open_top, open_bottom=open_bottom)


class DistComm:
    """torch.distributed point-to-point + all-reduce (nccl on GPUs, gloo on CPU tensors)."""

    def __init__(self, dist, rank, world, device=None):
        self.dist, self.rank, self.world, self.device = dist, rank, world, device

    def exchange(self, band):
        import torch
        ops = []
        for side, peer in ((0, self.rank - 1), (1, self.rank + 1)):
            if 0 <= peer < self.world:
                ops.append(self.dist.P2POp(self.dist.isend, band.send_row(side), peer))
                ops.append(self.dist.P2POp(self.dist.irecv, band.recv_stage(side), peer))
        if ops:
            for w in self.dist.batch_isend_irecv(ops):
                w.wait()
            if self.device is not None:
                torch.cuda.current_stream(self.device).synchronize()

    def total(self, value):
        import torch
        t = torch.tensor([int(value)], dtype=torch.int64, device=self.device if self.device is not None else "cpu")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return int(t.item())

    def lowered_and_min(self, lowered, parked):
        """One all-reduce for both convergence quantities: (any halo value lowered anywhere, lowest parked value)."""
        import torch
        t = torch.tensor([1.0 if lowered else 0.0, -float(parked)], dtype=torch.float64,
                         device=self.device if self.device is not None else "cpu")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        v = t.tolist()
        return v[0] > 0.0, -v[1]

    def reduce_float(self, value, op="min"):
        import torch
        t = torch.tensor([float(value)], dtype=torch.float64, device=self.device if self.device is not None else "cpu")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN if op == "min" else self.dist.ReduceOp.MAX)
        return float(t.item())


def run_bands(band, comm, max_outer=1 << 20):
    """Solve to global convergence. Returns the number of outer iterations (halo exchanges)."""
    band.begin()
    outer = 0
    while True:
        band.solve()
        comm.exchange(band)
        lowered = 0
        if band.open_top:
            lowered += band.import_side(0)
        if band.open_bottom:
            lowered += band.import_side(1)
        outer += 1
        if comm.total(lowered) == 0:
            return outer
        if outer >= max_outer:
            raise RuntimeError("band solve did not converge in %d exchanges" % max_outer)


def run_bands_ordered(band, comm, k_phases=4, max_outer=1 << 20):
    """As run_bands, but all bands advance through the values together: between two halo exchanges no band's
    tile schedule passes a common limit (k_phases threshold widths beyond the lowest pending value anywhere), so
    the bands a wavefront currently spans work at the same time instead of one after the other."""
    band.begin()
    step = k_phases * comm.reduce_float(band.delta(), "max")
    limit = step
    outer = 0
    inf = float("inf")
    while True:
        parked = band.solve_until(limit)
        comm.exchange(band)
        lowered = 0
        if band.open_top:
            lowered += band.import_side(0)
        if band.open_bottom:
            lowered += band.import_side(1)
        outer += 1
        any_lowered, gmin = comm.lowered_and_min(lowered, parked)
        if not any_lowered and gmin == inf:
            return outer
        if gmin < inf:
            limit = max(limit, gmin + step)
        if outer >= max_outer:
            raise RuntimeError("band solve did not converge in %d exchanges" % max_outer)


class BandSet:
    """The bands one rank owns under a CYCLIC distribution: with `world` ranks and `per_rank` bands each, rank r
    owns the global bands r, r + world, r + 2*world, ... of world*per_rank. A wavefront from a corner goal covers
    only some rows at a time; dealing the bands out cyclically gives every GPU a piece of the front for most of
    the sweep instead of one contiguous block that is busy a fraction of it.

    `make_band(g, parts)` builds the band object of global index g (GpuBand, or a CPU stand-in in the tests)."""

    def __init__(self, rank, world, per_rank, make_band):
        self.rank, self.world, self.parts = rank, world, world * per_rank
        self.ids = [rank + world * j for j in range(per_rank)]
        self.bands = {g: make_band(g, self.parts) for g in self.ids}

    def owner(self, g):
        return g % self.world

    def begin(self):
        for b in self.bands.values():
            b.begin()

    def delta(self):
        return max(b.delta() for b in self.bands.values())

    def solve_until(self, limit):
        return min(b.solve_until(limit) for b in self.bands.values())

    def import_all(self):
        lowered = 0
        for b in self.bands.values():
            if b.open_top:
                lowered += b.import_side(0)
            if b.open_bottom:
                lowered += b.import_side(1)
        return lowered

    def exchange(self, comm):
        """Swap boundary rows over every edge (g, g+1) this rank takes part in. Edges are walked in the same global
        order on every rank, so the messages between any two ranks are posted in matching order (NCCL has no tags)."""
        ops, local = [], []
        for g in range(self.parts - 1):
            up, dn = g in self.bands, (g + 1) in self.bands
            if up and dn:  # both ends here (world == 1)
                local.append(g)
            elif up:       # we own the upper band: its last row goes down, the lower band's first row comes up
                peer = self.owner(g + 1)
                ops.append(comm.dist.P2POp(comm.dist.isend, self.bands[g].send_row(1), peer))
                ops.append(comm.dist.P2POp(comm.dist.irecv, self.bands[g].recv_stage(1), peer))
            elif dn:
                peer = self.owner(g)
                ops.append(comm.dist.P2POp(comm.dist.irecv, self.bands[g + 1].recv_stage(0), peer))
                ops.append(comm.dist.P2POp(comm.dist.isend, self.bands[g + 1].send_row(0), peer))
        for g in local:
            self.bands[g + 1].recv_stage(0).copy_(self.bands[g].send_row(1))
            self.bands[g].recv_stage(1).copy_(self.bands[g + 1].send_row(0))
        if ops:
            for w in comm.dist.batch_isend_irecv(ops):
                w.wait()
        if comm.device is not None:
            import torch
            torch.cuda.current_stream(comm.device).synchronize()

    def close(self):
        for b in self.bands.values():
            b.close()


def run_bandset_ordered(bs, comm, k_phases=2, max_outer=1 << 20):
    """run_bands_ordered for a rank that owns several (cyclically dealt) bands."""
    bs.begin()
    step = k_phases * comm.reduce_float(bs.delta(), "max")
    limit = step
    outer = 0
    inf = float("inf")
    while True:
        parked = bs.solve_until(limit)
        bs.exchange(comm)
        lowered = bs.import_all()
        outer += 1
        any_lowered, gmin = comm.lowered_and_min(lowered, parked)
        if not any_lowered and gmin == inf:
            return outer
        if gmin < inf:
            limit = max(limit, gmin + step)
        if outer >= max_outer:
            raise RuntimeError("band solve did not converge in %d exchanges" % max_outer)


def run_bands_local_ordered(bands, k_phases=4, max_outer=1 << 20):
    """run_bands_ordered with all bands in this process."""
    for b in bands:
        b.begin()
    step = k_phases * max(b.delta() for b in bands)
    limit = step
    outer = 0
    inf = float("inf")
    while True:
        parked = [b.solve_until(limit) for b in bands]
        for r in range(len(bands) - 1):
            bands[r + 1].recv_stage(0).copy_(bands[r].send_row(1))
            bands[r].recv_stage(1).copy_(bands[r + 1].send_row(0))
        _sync(bands)
        lowered = 0
        for b in bands:
            if b.open_top:
                lowered += b.import_side(0)
            if b.open_bottom:
                lowered += b.import_side(1)
        outer += 1
        gmin = min(parked)
        if lowered == 0 and gmin == inf:
            return outer
        if gmin < inf:
            limit = max(limit, gmin + step)
        if outer >= max_outer:
            raise RuntimeError("band solve did not converge in %d exchanges" % max_outer)


def run_bands_local(bands, max_outer=1 << 20):
    """All bands in this process (one GPU or CPU): same protocol, rows copied directly between neighbours."""
    for b in bands:
        b.begin()
    outer = 0
    while True:
        for b in bands:
            b.solve()
        for r in range(len(bands) - 1):
            bands[r + 1].recv_stage(0).copy_(bands[r].send_row(1))
            bands[r].recv_stage(1).copy_(bands[r + 1].send_row(0))
        _sync(bands)
        lowered = 0
        for b in bands:
            if b.open_top:
                lowered += b.import_side(0)
            if b.open_bottom:
                lowered += b.import_side(1)
        outer += 1
        if lowered == 0:
            return outer
        if outer >= max_outer:
            raise RuntimeError("band solve did not converge in %d exchanges" % max_outer)


def _sync(bands):
    if bands and getattr(bands[0], "device", None) is not None:
        import torch
        torch.cuda.synchronize(bands[0].device)


class _DevRow:
    """CUDA-array-interface view of `n` float32 at device pointer `ptr` (so torch can wrap it without a copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class GpuBand:
    """One row band on one GPU over the band entry points of the C ABI (include/tasi_gpu.h)."""

    def __init__(self, kind, level, length, width, parts, r, device=0):
        import torch
        self.kind, self.length, self.width, self.parts, self.r, self.device = kind, length, width, parts, r, device
        geo = band_geometry(kind, length, parts, r)
        self.geo = geo
        self.open_top, self.open_bottom = geo["open_top"], geo["open_bottom"]
        self.p = _planner.PLANNERS[kind](level, device=device)
        self.p.band_config(geo["a"], self.open_top, self.open_bottom)
        self._torch = torch
        self._rows = {}

    # -- map ------------------------------------------------------------------------------------------------
    def set_map_from_global(self, cmap, occ=1.0):
        self.p.set_occupancy_threshold(occ)
        self.p.set_map(np.ascontiguousarray(cmap[self.geo["c0"]:self.geo["c1"]]))
        self._bind()

    def synth_map(self, seed, p_obs, start, goal, occ=1.0):
        self.p.set_occupancy_threshold(occ)
        g = self.geo
        self.p.synth_map(seed, self.width, g["c1"] - g["c0"], row0=g["c0"], p_obs=p_obs, start=start, goal=goal)
        self._bind()

    def _bind(self):
        self._rows = {}
        for side in (0, 1):
            s, rcv, n = self.p.band_rows(side)
            dev = "cuda:%d" % self.device
            self._rows[side] = (self._torch.as_tensor(_DevRow(s, n), device=dev),
                                self._torch.as_tensor(_DevRow(rcv, n), device=dev))

    def set_endpoints(self, start, goal):
        self.p.set_start(*start)
        self.p.set_goal(*goal)  # global coordinates; the band pins the goal only if it owns that row

    # -- protocol ---------------------------------------------------------------------------------------------
    def begin(self):
        rc = self.p.band_begin()
        assert rc == 0, rc

    def solve(self):
        return self.p.band_solve()

    def solve_until(self, limit):
        return self.p.band_solve_until(limit)

    def delta(self):
        return self.p.band_delta()

    def send_row(self, side):
        return self._rows[side][0]

    def recv_stage(self, side):
        return self._rows[side][1]

    def import_side(self, side):
        return self.p.band_import(side)

    # -- results ----------------------------------------------------------------------------------------------
    def read_owned(self):
        return self.p.read_field()

    def stats(self):
        return self.p.stats()

    def close(self):
        self.p.close()


def solve_sharded_local(kind, cmap, goal, parts, level=None, device=0, occ=1.0, start=(0.0, 0.0), ordered=0):
    """Plan `cmap` as `parts` row bands on ONE GPU and return the assembled field (the test of the band path)."""
    cls = _planner.PLANNERS[kind]
    level = cls.max_level if level is None else level
    L, W = cmap.shape
    bands = [GpuBand(kind, level, L, W, parts, r, device) for r in range(parts)]
    for b in bands:
        b.set_map_from_global(cmap, occ)
        b.set_endpoints(start, goal)
    outer = run_bands_local_ordered(bands, ordered) if ordered else run_bands_local(bands)
    field = np.concatenate([b.read_owned() for b in bands], axis=0)
    stats = [b.stats() for b in bands]
    for b in bands:
        b.close()
    return field, outer, stats
