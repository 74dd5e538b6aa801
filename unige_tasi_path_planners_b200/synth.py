"""Deterministic synthetic traversability grids.

cost(x, y) = 255                      if h mod 1000 < 10 * p_obs
             1 + ((h >> 16) mod 64)   otherwise,        h = splitmix64(seed XOR (x*W + y))

with a 9x9 block of cost 1 cleared around the start and the goal. The same function exists as a CUDA
kernel (tasi_synth_map in csrc/tasi_kernels.cu) for grids too large to build on the host; the two are
checked against each other in tests/.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(z):
    z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def synth_rows(seed, x0, x1, W, p_obs=10):
    """Rows x0..x1-1 of the (unbounded-length) synthetic map of width W."""
    with np.errstate(over="ignore"):
        idx = (np.arange(x0, x1, dtype=np.uint64)[:, None] * np.uint64(W)
               + np.arange(W, dtype=np.uint64)[None, :])
        h = splitmix64(np.uint64(seed) ^ idx)
    obst = (h % np.uint64(1000)) < np.uint64(10 * p_obs)
    cost = (np.uint64(1) + ((h >> np.uint64(16)) % np.uint64(64))).astype(np.uint8)
    cost[obst] = 255
    return cost


def clear_block(cmap, cx, cy, half=4, row0=0):
    """Set the (2*half+1)^2 block of cells centred on (cx, cy) to cost 1 (clipped to the array)."""
    L, W = cmap.shape
    xa, xb = max(cx - half - row0, 0), min(cx + half + 1 - row0, L)
    ya, yb = max(cy - half, 0), min(cy + half + 1, W)
    if xa < xb and ya < yb:
        cmap[xa:xb, ya:yb] = 1


def synth_map(seed, L, W, p_obs=10, start=None, goal=None):
    cmap = np.empty((L, W), np.uint8)
    step = max(1, (1 << 24) // max(W, 1))
    for x0 in range(0, L, step):
        x1 = min(L, x0 + step)
        cmap[x0:x1] = synth_rows(seed, x0, x1, W, p_obs)
    for p in (start, goal):
        if p is not None:
            clear_block(cmap, int(round(p[0])), int(round(p[1])))
    return cmap


def default_endpoints(L, W):
    """start = (5,5), goal = (L-6, W-6): the reference then expands ~100% of the grid."""
    return (5.0, 5.0), (float(L - 6), float(W - 6))
