"""Generate tests/golden/*.npz by running the UNMODIFIED reference planners (oracle/_ref, built from
/root/reference by oracle/Makefile). Run in the build container only:

    python scripts/make_golden.py

/root/reference does not exist on the GPU box, so the vectors are committed. Three families of fixtures:

  ops_{fd,mfd}.npz   one application of the reference's compute_optimal_cost
                     (FieldDStar/FieldDPlanner_impl.h:269-320, ShiftedGridFastMarching/ShiftedGridPlanner_impl.h:399-437)
                     on random fields: per-triangle values and per-node minima over the valid triangles
  ops_dfm.npz        DFMPlanner::compute_optimal_cost (DynamicFastMarching_impl.h:322-342) on random tuples
  field_*.npz        step() on the four maps of Tests/Tests/*.bmp and on seeded synthetic maps: closed set
                     (g == rhs < INF and g <= top key), its values, the reference extractor's path, and a replan
                     (patch_map + set_start + step)
"""
import os
import re
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from unige_tasi_path_planners_b200 import synth  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
INF = np.float32(np.inf)
FD, DFM, MFD = 0, 1, 2
NAMES = {FD: "fd", DFM: "dfm", MFD: "mfd"}
MAXLVL = {FD: 1, DFM: 1, MFD: 2}

# the 8 triangles (p1 = orthogonal neighbour, p2 = diagonal neighbour) around a node, as offsets, and the
# offsets of cell c (the triangle's own cell) and b (the cell across edge n-p1) relative to the node
# (cells touching node (x,y): (x-1,y-1) (x-1,y) (x,y-1) (x,y) — Node.cpp:44-50)
TRIS = [((-1, 0), (-1, -1), (-1, -1), (-1, 0)),   # N with NW: c = top-left cell, b = top-right
        ((-1, 0), (-1, 1), (-1, 0), (-1, -1)),    # N with NE
        ((1, 0), (1, -1), (0, -1), (0, 0)),       # S with SW
        ((1, 0), (1, 1), (0, 0), (0, -1)),        # S with SE
        ((0, -1), (-1, -1), (-1, -1), (0, -1)),   # W with NW
        ((0, -1), (1, -1), (0, -1), (-1, -1)),    # W with SW
        ((0, 1), (-1, 1), (-1, 0), (0, 0)),       # E with NE
        ((0, 1), (1, 1), (0, 0), (-1, 0))]        # E with SE


def cell_cost(cmap, x, y, thr=255):
    L, W = cmap.shape
    if x < 0 or y < 0 or x >= L or y >= W or cmap[x, y] >= thr:
        return INF
    return np.float32(cmap[x, y])


def random_field(rng, shape, scale):
    g = (rng.random(shape) * scale).astype(np.float32)
    # make neighbouring values close to each other often (interesting branches need |g1-g2| <~ c)
    base = np.float32(rng.random() * 500.0)
    g = (base + g).astype(np.float32)
    g[rng.random(shape) < 0.12] = INF
    return g


def make_ops_node(kind):
    rng = np.random.default_rng(100 + kind)
    maps, fields, node_vals, tri_rows = [], [], [], []
    for trial in range(24):
        L, W = 12, 14
        cmap = rng.integers(1, 255, size=(L, W)).astype(np.uint8)
        if trial % 3 == 0:
            cmap = rng.integers(1, 40, size=(L, W)).astype(np.uint8)
        cmap[rng.random((L, W)) < 0.15] = 255
        g = random_field(rng, (L + 1, W + 1), [3.0, 30.0, 300.0][trial % 3])
        r = O.RefPlanner(kind, 1)
        r.set_map(cmap, 1.0)
        out = np.full((L + 1, W + 1), INF, np.float32)
        for x in range(L + 1):
            for y in range(W + 1):
                m = INF
                for (d1, d2, dc, db) in TRIS:
                    p1 = (x + d1[0], y + d1[1])
                    p2 = (x + d2[0], y + d2[1])
                    if not (0 <= p1[0] <= L and 0 <= p1[1] <= W and 0 <= p2[0] <= L and 0 <= p2[1] <= W):
                        continue  # Graph::consecutive_neighbors drops triangles with an off-grid node
                    g1, g2 = g[p1], g[p2]
                    v = np.float32(r.op_tri((x, y), p1, p2, float(g1), float(g2)))
                    c = cell_cost(cmap, x + dc[0], y + dc[1])
                    b = cell_cost(cmap, x + db[0], y + db[1])
                    tri_rows.append((g1, g2, b, c, v))
                    if v < m:
                        m = v
                out[x, y] = m
        r.close()
        maps.append(cmap); fields.append(g); node_vals.append(out)
    tri = np.array(tri_rows, dtype=np.float32)
    # keep the fixture small: every branch is hit thousands of times, 20000 rows are plenty
    sel = np.random.default_rng(7).permutation(len(tri))[:20000]
    np.savez_compressed(os.path.join(OUT, "ops_%s.npz" % NAMES[kind]), maps=np.array(maps), fields=np.array(fields),
                        node_vals=np.array(node_vals), tri=tri[sel])
    print("ops", NAMES[kind], len(tri), "triangles,", np.isfinite(tri[:, 4]).mean(), "finite")


def make_ops_dfm():
    rng = np.random.default_rng(555)
    n = 20000
    ga = (rng.random(n) * 400).astype(np.float32)
    gb = (ga + (rng.random(n) - 0.5).astype(np.float32) * np.float32(120.0)).astype(np.float32)
    gb = np.abs(gb).astype(np.float32)
    tau = rng.integers(1, 255, size=n).astype(np.float32)
    h = np.where(rng.random(n) < 0.5, np.float32(1.0), np.float32(np.sqrt(np.float32(2.0)))).astype(np.float32)
    ga[rng.random(n) < 0.05] = INF
    gb[rng.random(n) < 0.05] = INF
    r = O.RefPlanner(DFM, 1)
    r.set_map(np.ones((4, 4), np.uint8), 1.0)
    # hypotf(1,1) in float is what level 1 passes for the diagonal stencil (DynamicFastMarching_impl.h:270-313)
    out = np.array([r.op_stencil(float(a), float(b), float(t), float(hh)) for a, b, t, hh in zip(ga, gb, tau, h)],
                   dtype=np.float32)
    r.close()
    np.savez_compressed(os.path.join(OUT, "ops_dfm.npz"), ga=ga, gb=gb, tau=tau, h=h, out=out)
    print("ops dfm", n, np.isfinite(out).mean(), "finite")


def bmp_cases():
    import cv2
    d = os.path.join(REF, "Tests", "Tests")
    for f in sorted(os.listdir(d)):
        m = re.match(r"(.+)_(\d+)_(\d+)_(\d+)_(\d+)_\.bmp$", f)
        if not m:
            continue
        img = cv2.imread(os.path.join(d, f), cv2.IMREAD_GRAYSCALE)
        data = (~img).astype(np.uint8)          # Simulator/simulator/run_simulator.py:106-113 (high-res data)
        data = (data + (data == 0)).astype(np.uint8)
        fx, fy, tx, ty = [float(v) for v in m.groups()[1:]]
        yield m.group(1), data, (fx, fy), (tx, ty)


def synth_cases():
    for p_obs in (0, 10, 30):
        L, W = 96, 80
        start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
        yield "synth%d" % p_obs, synth.synth_map(900 + p_obs, L, W, p_obs, start, goal), start, goal


def check_level(kind, level, cmap, start, goal, timeout=20):
    """Lower optimisation levels must close the same values as the top level (one kernel
    serves all levels). Runs in a forked child under a timeout: DFMPlanner<0> does not terminate on some maps
    (an ulp-scale raise/lower cycle between two cells keeps its queue non-empty) — recorded, not a gate."""
    import time
    pid = os.fork()
    if pid == 0:
        res = []
        for lv in (MAXLVL[kind], level):
            r = O.RefPlanner(kind, lv)
            r.set_map(cmap, 1.0); r.set_start(*start); r.set_goal(*goal)
            r.step()
            res.append(r.closed_mask())
        both = res[0][0] & res[1][0]
        same = np.array_equal(res[0][1][both], res[1][1][both])
        print("   level %d vs %d: %d common closed, bit-identical=%s" % (level, MAXLVL[kind], both.sum(), same),
              flush=True)
        os._exit(0)
    t0 = time.time()
    while time.time() - t0 < timeout:
        done, _ = os.waitpid(pid, os.WNOHANG)
        if done:
            return
        time.sleep(0.05)
    os.kill(pid, 9)
    os.waitpid(pid, 0)
    print("   level %d of %s: reference did NOT terminate within %d s on this map" % (level, NAMES[kind], timeout),
          flush=True)


def run_case(name, cmap, start, goal):
    L, W = cmap.shape
    rng = np.random.default_rng(zlib.crc32(name.encode()))
    # the replan patch: a window between start and goal, redrawn (mixed raises and lowers, some obstacles)
    ph, pw = max(4, L // 4), max(4, W // 4)
    px = int(min(max((start[0] + goal[0]) / 2 - ph / 2, 0), L - ph))
    py = int(min(max((start[1] + goal[1]) / 2 - pw / 2, 0), W - pw))
    patch = rng.integers(1, 120, size=(ph, pw)).astype(np.uint8)
    patch[rng.random((ph, pw)) < 0.15] = 255
    for kind in (FD, DFM, MFD):
        out = {"cmap": cmap, "start": np.float32(start), "goal": np.float32(goal), "patch": patch,
               "patch_xy": np.int32([px, py])}
        ref_closed = None
        for level in range(MAXLVL[kind] - 1, -1, -1):
            check_level(kind, level, cmap, start, goal)
        for level in (MAXLVL[kind],):
            r = O.RefPlanner(kind, level)
            r.set_map(cmap, 1.0)
            r.set_start(*start)
            r.set_goal(*goal)
            assert r.step() == 0
            mask, g = r.closed_mask()
            if level == MAXLVL[kind]:
                ref_closed = (mask, g)
                out["mask"] = mask
                out["g"] = np.where(mask, g, INF).astype(np.float32)
                out["top_key"] = np.float32(r.top_key())
                out["expanded"] = np.int64(r.counters()[1])
                allow_indirect = kind != MFD  # FDSTAR/main.cpp:82, DFM/main.cpp:80 true; SGDFM/main.cpp:97 false
                xy, cost, dist, _ = r.extract(None, 20 if kind == DFM else 2000, True, allow_indirect)
                out["path"] = xy
                out["path_cost"] = np.float32(cost)
                out["path_dist"] = np.float32(dist)
                # replan: patch, move the start a little (patches are consumed only on a new start), step
                r.patch_map(patch, px, py)
                s2 = (start[0] + (1.0 if start[0] + 1 < L - 1 else -1.0), start[1])
                r.set_start(*s2)
                assert r.step() == 0
                mask2, g2 = r.closed_mask()
                out["start2"] = np.float32(s2)
                out["mask2"] = mask2
                out["g2"] = np.where(mask2, g2, INF).astype(np.float32)
                out["updated2"] = np.int64(r.counters()[0])
            r.close()
        np.savez_compressed(os.path.join(OUT, "field_%s_%s.npz" % (name, NAMES[kind])), **out)
        print("field", name, NAMES[kind], cmap.shape, "closed", int(out["mask"].sum()), "path", len(out["path"]),
              "cost", float(out["path_cost"]), "replan closed", int(out["mask2"].sum()))


def main():
    if not O.ref_available():
        O.build(ref=True)
    os.makedirs(OUT, exist_ok=True)
    make_ops_node(FD)
    make_ops_node(MFD)
    make_ops_dfm()
    for case in list(bmp_cases()) + list(synth_cases()):
        run_case(*case)


if __name__ == "__main__":
    main()
