"""Row-band sharding (unige_tasi_path_planners_b200/bands.py): host protocol on CPU with gloo (world_size 2 and 3),
and on the GPU all bands of one grid on one device against the unsharded field (bit-identical)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from unige_tasi_path_planners_b200 import bands, synth

HERE = os.path.dirname(os.path.abspath(__file__))


def test_split_and_geometry():
    assert bands.split_rows(10, 3) == [(0, 4), (4, 7), (7, 10)]
    for kind in (0, 1, 2):
        L = 37
        n = L if kind == 1 else L + 1
        prev_b = 0
        for r in range(4):
            g = bands.band_geometry(kind, L, 4, r)
            assert g["a"] == prev_b and g["b"] > g["a"]
            prev_b = g["b"]
            assert g["open_top"] == (r > 0) and g["open_bottom"] == (r < 3)
            if kind == 1:
                assert (g["c0"], g["c1"]) == (g["a"], g["b"])
            else:
                # every owned node row x needs cell rows x-1 and x (where they exist)
                assert g["c0"] == max(g["a"] - 1, 0) and g["c1"] == min(g["b"], L)
        assert prev_b == n


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("kind,parts,mode", [(0, 2, "plain"), (1, 2, "plain"), (2, 3, "plain"), (0, 2, "ordered")])
def test_gloo_bands_equal_single_domain(oracle, kind, parts, mode, tmp_path):
    out = str(tmp_path / "band")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()), WORLD_SIZE=str(parts),
               OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "_band_worker.py"), str(kind), str(parts), out, mode],
                              env=dict(env, RANK=str(r)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(parts)]
    for p in procs:
        o, _ = p.communicate(timeout=300)
        assert p.returncode == 0, o.decode()[-3000:]
    got = np.concatenate([np.load(out + ".%d.npy" % r) for r in range(parts)], axis=0)
    L, W = 22, 17
    goal = (float(L - 4), float(W - 5))
    cmap = synth.synth_map(5, L, W, 12, (2, 2), goal)
    want = oracle.solve(kind, cmap, goal, method="sweep")
    assert got.shape == want.shape
    if kind == 1:  # MS-DFM's fixed point is order-dependent at the ulp level
        m = np.isfinite(want)
        assert np.array_equal(np.isfinite(got), m)
        assert (np.abs(got[m] - want[m]) / np.maximum(want[m], 1e-30)).max() <= 1e-6
    else:
        assert np.array_equal(got, want)  # same operator, same fixed point: bit-identical, INF included
    outers = [int(open(out + ".%d.txt" % r).read()) for r in range(parts)]
    assert len(set(outers)) == 1 and outers[0] >= 2  # every rank left the loop in the same iteration


@pytest.mark.parametrize("kind,world,per", [(0, 2, 2), (2, 3, 2), (0, 2, 3)])
def test_gloo_cyclic_bandset_equal_single_domain(oracle, kind, world, per, tmp_path):
    """Several cyclically dealt bands per rank: the edge-ordered exchange matches messages between rank pairs."""
    out = str(tmp_path / "band")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()), WORLD_SIZE=str(world),
               OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "_band_worker.py"), str(kind), str(world), out,
                               "cyclic%d" % per], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(world)]
    for p in procs:
        o, _ = p.communicate(timeout=300)
        assert p.returncode == 0, o.decode()[-3000:]
    got = np.concatenate([np.load(out + ".g%d.npy" % g) for g in range(world * per)], axis=0)
    L, W = 22, 17
    goal = (float(L - 4), float(W - 5))
    cmap = synth.synth_map(5, L, W, 12, (2, 2), goal)
    want = oracle.solve(kind, cmap, goal, method="sweep")
    assert got.shape == want.shape
    assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("parts,ordered", [(2, 0), (3, 0), (5, 0), (2, 1), (3, 4), (4, 2)])
def test_gpu_bands_equal_unsharded(kind, parts, ordered):
    from unige_tasi_path_planners_b200 import PLANNERS
    L, W = 230, 190
    start, goal = (5.0, 5.0), (float(L - 60), float(W - 6))
    cmap = synth.synth_map(61, L, W, 10, start, goal)
    p = PLANNERS[kind](PLANNERS[kind].max_level)
    p.set_occupancy_threshold(1.0)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    want = p.read_field()
    p.close()
    got, outer, stats = bands.solve_sharded_local(kind, cmap, goal, parts, start=start, ordered=ordered)
    assert got.shape == want.shape
    assert outer >= 2
    if kind == 1:
        m = np.isfinite(want)
        assert np.array_equal(np.isfinite(got), m)
        assert (np.abs(got[m] - want[m]) / np.maximum(want[m], 1e-30)).max() <= 1e-6
    else:
        assert np.array_equal(got, want)
