"""Worker of tests/test_bands.py: one rank of a world_size-N gloo run of bands.run_bands on CPU tensors.

The band here is a CPU stand-in (oracle element updates swept over the owned rows, test infrastructure); what is
under test is the product's host-side protocol: bands.run_bands, bands.DistComm, bands.band_geometry.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from unige_tasi_path_planners_b200 import bands, synth  # noqa: E402


class CpuBand:
    def __init__(self, kind, cmap, goal, parts, r):
        self.kind, self.cmap = kind, np.ascontiguousarray(cmap)
        L, W = cmap.shape
        geo = bands.band_geometry(kind, L, parts, r)
        self.a, self.b = geo["a"], geo["b"]
        self.open_top, self.open_bottom = geo["open_top"], geo["open_bottom"]
        self.shape = O.field_shape(kind, L, W)
        self.goal = O.goal_elem(*goal)
        self.g = None
        self.stage = [torch.full((self.shape[1],), float("inf")), torch.full((self.shape[1],), float("inf"))]

    def begin(self):
        self.g = np.full(self.shape, np.inf, np.float32)
        gx, gy = self.goal
        if self.a <= gx < self.b:
            self.g[gx, gy] = 0.0
        self.gt = torch.from_numpy(self.g)  # shares memory

    def solve(self):
        L, W = self.cmap.shape
        lib = O.lib()
        mp = self.cmap.ctypes.data_as(O._u8p)
        changed = True
        n = 0
        while changed:
            changed = False
            for x in range(self.a, self.b):
                for y in range(self.shape[1]):
                    if (x, y) == self.goal:
                        continue
                    v = lib.tasi_oracle_update(self.kind, mp, W, L, 255, self.g.ctypes.data_as(O._f32p), x, y)
                    if v < self.g[x, y]:
                        self.g[x, y] = v
                        changed = True
            n += 1
        return n

    def solve_until(self, limit):
        self.solve()  # a band may ignore the limit: it then reports nothing parked
        return float("inf")

    def delta(self):
        return 64.0

    def send_row(self, side):
        return self.gt[self.a] if side == 0 else self.gt[self.b - 1]

    def recv_stage(self, side):
        return self.stage[side]

    def import_side(self, side):
        row = self.a - 1 if side == 0 else self.b
        new = self.stage[side].numpy()
        lower = new < self.g[row]
        self.g[row][lower] = new[lower]
        return int(lower.sum())


def main():
    kind, parts, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    rank = int(os.environ["RANK"])
    dist.init_process_group("gloo", rank=rank, world_size=parts)
    L, W = 22, 17
    goal = (float(L - 4), float(W - 5))
    cmap = synth.synth_map(5, L, W, 12, (2, 2), goal)
    comm = bands.DistComm(dist, rank, parts)
    if len(sys.argv) > 4 and sys.argv[4].startswith("cyclic"):
        per = int(sys.argv[4][6:])
        bs = bands.BandSet(rank, parts, per, lambda g, n: CpuBand(kind, cmap, goal, n, g))
        outer = bands.run_bandset_ordered(bs, comm, 2)
        for g, b in bs.bands.items():
            np.save(out + ".g%d.npy" % g, b.g[b.a:b.b])
        with open(out + ".%d.txt" % rank, "w") as f:
            f.write(str(outer))
        dist.destroy_process_group()
        return
    band = CpuBand(kind, cmap, goal, parts, rank)
    outer = bands.run_bands_ordered(band, comm, 2) if len(sys.argv) > 4 and sys.argv[4] == "ordered" else \
        bands.run_bands(band, comm)
    np.save(out + ".%d.npy" % rank, band.g[band.a:band.b])
    with open(out + ".%d.txt" % rank, "w") as f:
        f.write(str(outer))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
