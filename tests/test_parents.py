"""Parent (back-pointer) field: csrc/tasi_parents.cuh against the INFO of the reference's level-1
planners (FieldDPlanner_impl.h:91-95, DynamicFastMarching_impl.h:322-342).

The reference's INFO is the node b of the triangle (b, ccw(b)) that gave rhs (FD*/MFD*), or the two cells of the
winning stencil (MS-DFM). Several triangles often tie (walking along an edge through the cheaper of its two cells
is the same cost from either side), and among ties the reference keeps whichever its expansion order met last, so
the bar is: the reported triangle attains the element's value bit for bit, and equals the reference's INFO whenever
the minimum is unique; where they differ the reference's triangle must tie.
"""
import numpy as np
import pytest

import hostops

from unige_tasi_path_planners_b200 import _lib, synth

NBR8 = [(-1, 0), (-1, -1), (0, -1), (1, -1), (1, 0), (1, 1), (0, 1), (-1, 1)]  # Graph::neighbors_8 order


def cases():
    for seed in range(4):
        L, W = 40 + 7 * seed, 50 - 3 * seed
        start, goal = (3.0, 3.0), (float(L - 9), float(W - 8))
        yield seed, synth.synth_map(seed, L, W, [0, 10, 25, 10][seed], start, goal), start, goal


@pytest.mark.parametrize("kind", [0, 2])
def test_node_parents_attain_the_value_and_match_the_reference(oracle, kind):
    have_ref = oracle.ref_available()
    checked = same = ties = 0
    for seed, cmap, start, goal in cases():
        F = oracle.solve(kind, cmap, goal)
        par = hostops.host_parents(kind, F, cmap)
        fin = np.isfinite(F) & (F > 0)
        assert not par[~fin].any()                      # goals and unreached nodes have no parents
        assert np.all(par[fin][:, :2].any(axis=1)) and np.all(par[fin][:, 2:].any(axis=1))
        xs, ys = np.nonzero(fin)
        for x, y in list(zip(xs, ys))[::7]:
            tri = hostops.host_node_triangles(kind, F, cmap, x, y)
            k = NBR8.index((int(par[x, y, 0]), int(par[x, y, 1])))
            assert NBR8[(k + 7) & 7] == (int(par[x, y, 2]), int(par[x, y, 3]))  # second parent = ccw(b)
            assert tri[k] == F[x, y] == tri.min()
            assert not any(tri[j] == tri[k] for j in range(k + 1, 8))          # the last of the minima
        if not have_ref:
            continue
        r = oracle.RefPlanner(kind, 1)
        r.set_map(cmap); r.set_start(*start); r.set_goal(*goal)
        r.step()
        info = r.dump_info()
        closed, _ = r.closed_mask()
        for x, y in zip(*np.nonzero(closed & fin)):
            b = (int(info[x, y, 0]) - x, int(info[x, y, 1]) - y)
            checked += 1
            if b == (int(par[x, y, 0]), int(par[x, y, 1])):
                same += 1
            else:
                tri = hostops.host_node_triangles(kind, F, cmap, x, y)
                assert tri[NBR8.index(b)] == F[x, y], "reference parent of (%d,%d) is not a minimum" % (x, y)
                ties += 1
        r.close()
    if have_ref:
        assert checked > 5000 and same > 0.5 * checked and same + ties == checked


def test_cell_parents_match_the_reference(oracle):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    checked = same = 0
    for seed, cmap, start, goal in cases():
        r = oracle.RefPlanner(1, 1)
        r.set_map(cmap); r.set_start(*start); r.set_goal(*goal)
        r.step()
        info = r.dump_info()
        closed, g = r.closed_mask()
        par = hostops.host_parents(1, np.where(closed, g, np.inf).astype(np.float32), cmap)
        # parents read settled neighbours only: compare where the whole neighbourhood is closed
        inner = closed.copy()
        for dx, dy in NBR8:
            sh = np.zeros_like(closed)
            xs = slice(max(dx, 0), closed.shape[0] + min(dx, 0)); xd = slice(max(-dx, 0), closed.shape[0] + min(-dx, 0))
            ys = slice(max(dy, 0), closed.shape[1] + min(dy, 0)); yd = slice(max(-dy, 0), closed.shape[1] + min(-dy, 0))
            sh[xd, yd] = closed[xs, ys] | ~np.isfinite(g[xs, ys])
            inner &= sh
        for x, y in zip(*np.nonzero(inner & (g > 0))):
            a = (int(info[x, y, 0]) - x, int(info[x, y, 1]) - y)
            b = (0, 0) if (info[x, y, 2], info[x, y, 3]) == (-1, -1) else (int(info[x, y, 2]) - x, int(info[x, y, 3]) - y)
            checked += 1
            same += (a + b) == tuple(int(v) for v in par[x, y])
        r.close()
    assert checked > 3000 and same >= 0.99 * checked


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2])
def test_gpu_parents_equal_host_parents(kind):
    from unige_tasi_path_planners_b200 import PLANNERS
    for seed, cmap, start, goal in cases():
        p = PLANNERS[kind](1)
        p.set_occupancy_threshold(1.0)
        p.set_map(cmap); p.set_start(*start); p.set_goal(*goal)
        with pytest.raises(Exception):
            p.read_parents()  # no plan yet
        assert p.step() == 0
        F = p.read_field()
        got = p.read_parents()
        assert np.array_equal(got, hostops.host_parents(kind, F, cmap))
        w = p.read_parents(5, 7, 11, 13)
        assert np.array_equal(w, got[5:16, 7:20])
        with pytest.raises(Exception):
            p.read_parents(0, 0, F.shape[0] + 1, 3)
        # following the lower parent from the start's element walks downhill to the goal
        x, y = (int(start[0]), int(start[1]))
        for _ in range(F.size):
            if F[x, y] == 0:
                break
            cand = [(int(got[x, y, i]), int(got[x, y, i + 1])) for i in (0, 2) if got[x, y, i:i + 2].any()]
            assert cand
            dx, dy = min(cand, key=lambda d: F[x + d[0], y + d[1]])
            assert F[x + dx, y + dy] < F[x, y]
            x, y = x + dx, y + dy
        assert F[x, y] == 0
        p.close()


def test_parents_against_committed_reference_info(oracle):
    """The same bar against tests/golden/parents.npz (scripts/make_golden_parents.py): the reference's INFO on its
    closed set, committed, so the check also runs where oracle/_ref is absent."""
    import os
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "parents.npz"))
    for i in range(int(d["n"][0])):
        k = "c%d_" % i
        kind, cmap, goal = int(d[k + "kind"][0]), d[k + "map"], tuple(float(v) for v in d[k + "goal"])
        closed, g, info = d[k + "closed"], d[k + "g"], d[k + "info"]
        F = oracle.solve(kind, cmap, goal)
        if kind != 1:
            assert np.array_equal(F[closed], g[closed])  # the oracle's field is the reference's on its closed set
        par = hostops.host_parents(kind, F, cmap)
        checked = same = 0
        for x, y in zip(*np.nonzero(closed & (F > 0))):
            if kind == 1:
                # the reference's pair was taken from ITS neighbour values at expansion time; compare where all 8
                # neighbours are closed too (or unreachable)
                nb = [(x + dx, y + dy) for dx, dy in NBR8]
                if not all(0 <= a < F.shape[0] and 0 <= b < F.shape[1] and (closed[a, b] or not np.isfinite(F[a, b]))
                           for a, b in nb):
                    continue
                a = (int(info[x, y, 0]) - x, int(info[x, y, 1]) - y)
                b = (0, 0) if (info[x, y, 2], info[x, y, 3]) == (-1, -1) else (int(info[x, y, 2]) - x, int(info[x, y, 3]) - y)
                checked += 1
                same += (a + b) == tuple(int(v) for v in par[x, y])
            else:
                b = (int(info[x, y, 0]) - x, int(info[x, y, 1]) - y)
                checked += 1
                if b == (int(par[x, y, 0]), int(par[x, y, 1])):
                    same += 1
                else:  # a tie: the reference's triangle attains the minimum too
                    assert hostops.host_node_triangles(kind, F, cmap, x, y)[NBR8.index(b)] == F[x, y]
        assert checked > 500 and same >= (0.97 if kind == 1 else 0.5) * checked, (kind, checked, same)
