"""tasi_patch_map_static — the entry point that allocates nothing (scratch in a file-scope static device array; no
counterpart in the reference) — against tasi_patch_map, through the C ABI:
two calls in a row (the second one smaller than the first, so it runs over the first call's leftovers in the shared
array), then the replan. Same n_changed, same cost grid, same num_nodes_updated, and the replanned fields agree within
the tolerance of the parity tests (FD*/MFD* bit-exact, MS-DFM 1e-4 relative). Collected last on purpose."""
import numpy as np
import pytest

from unige_tasi_path_planners_b200 import PLANNERS, synth
from unige_tasi_path_planners_b200.planner import TasiError

L, W = 300, 260
START, GOAL = (5.0, 5.0), (float(L - 6), float(W - 6))


def _planner(kind, cmap):
    p = PLANNERS[kind](PLANNERS[kind].max_level, device=0)
    p.set_occupancy_threshold(1.0)
    p.set_map(cmap)
    p.set_start(*START)
    p.set_goal(*GOAL)
    assert p.step() == 0
    return p


def _same_field(kind, got, want):
    assert np.array_equal(np.isfinite(got), np.isfinite(want)), "reachability differs"
    m = np.isfinite(want)
    if kind == 1:
        assert (np.abs(got[m] - want[m]) / np.maximum(want[m], 1e-30)).max(initial=0.0) <= 1e-4
    else:
        assert np.array_equal(got[m], want[m])


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2])
def test_static_patch_equals_patch_map_on_two_calls_in_a_row(kind, gpu_lib):
    cmap = synth.synth_map(2024, L, W, 10, START, GOAL)
    a, b = _planner(kind, cmap), _planner(kind, cmap)          # a: tasi_patch_map, b: tasi_patch_map_static
    patches = [(synth.synth_rows(71, 0, 120, 96, 30), 90, 70),   # 120 x 96: raises and lowers costs
               (synth.synth_rows(72, 0, 40, 33, 0), 20, 150)]    # 40 x 33 of free cells: smaller than the first
    try:
        for i, (patch, x, y) in enumerate(patches):
            na = a.patch_map(patch, x, y)
            nb = b.patch_map_static(patch, x, y)
            assert na == nb and na > 0
            assert np.array_equal(a.read_map(), b.read_map())
            pos = (8.0 + i, 9.0)
            a.set_start(*pos)
            b.set_start(*pos)
            assert a.step() == 0 and b.step() == 0
            assert a.num_nodes_updated == b.num_nodes_updated and a.num_nodes_updated > 0
            _same_field(kind, b.read_field(), a.read_field())
        # ... and the other way round on the same handles: the two entry points can be mixed
        patch, x, y = synth.synth_rows(73, 0, 64, 64, 20), 200, 10
        assert a.patch_map_static(patch, x, y) == b.patch_map(patch, x, y)
        a.set_start(12.0, 12.0)
        b.set_start(12.0, 12.0)
        assert a.step() == 0 and b.step() == 0
        _same_field(kind, a.read_field(), b.read_field())
    finally:
        a.close()
        b.close()


@pytest.mark.gpu
def test_static_patch_rejects_what_does_not_fit(gpu_lib):
    side = 1100
    start, goal = synth.default_endpoints(side, side)
    p = PLANNERS[0](1, device=0)
    try:
        p.set_occupancy_threshold(1.0)
        p.synth_map(5, side, side, 0, 10, start, goal)
        before = p.read_map()
        big = np.ones((1025, 1024), np.uint8)                   # 1 024 cells over the limit
        assert big.size > p.PATCH_STATIC_MAX_CELLS
        with pytest.raises(TasiError) as e:
            p.patch_map_static(big, 0, 0)
        assert "maximum" in str(e.value) and str(p.PATCH_STATIC_MAX_CELLS) in str(e.value)
        assert np.array_equal(p.read_map(), before)             # nothing was applied
        with pytest.raises(TasiError):
            p.patch_map_static(np.ones((8, 8), np.uint8), side - 4, 0)   # outside the map, as tasi_patch_map
        exact = np.full((1024, 1024), 7, np.uint8)              # exactly the limit is accepted
        assert p.patch_map_static(exact, 10, 10) > 0
        assert np.array_equal(p.read_map()[10:1034, 10:1034], exact)
    finally:
        p.close()
