"""The sharded planner behind the C ABI (tasi_create_sharded, csrc/tasi_shard.cuh): one grid cut into cyclically dealt row
bands, stacked per rank into one tile queue, exchanged and steered on the device streams.

The reference has no multi-device path; the contract is bit-identity with the single-device field, which the
other tests pin to the reference. With one GPU the ranks are given as a repeated device id ([0, 0, 0]): every rank then
has its own stack, stream and buffers and the whole protocol runs (pack, event-ordered peer reads, merge, reduction, limit);
NCCL needs distinct devices and is exercised by `bench.py --gpus N` (bands record, verified by checksum).
"""
import numpy as np
import pytest

from unige_tasi_path_planners_b200 import synth

pytestmark = pytest.mark.gpu
KINDS = [0, 1, 2]


def single(kind, cmap, start, goal):
    from unige_tasi_path_planners_b200 import PLANNERS
    p = PLANNERS[kind](PLANNERS[kind].max_level)
    p.set_occupancy_threshold(1.0)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    return p


def sharded(kind, ranks, per_rank, phases=2):
    from unige_tasi_path_planners_b200 import ShardedPlanner
    p = ShardedPlanner(kind, devices=[0] * ranks, bands_per_device=per_rank, phases_per_exchange=phases)
    p.set_occupancy_threshold(1.0)
    return p


def same(kind, got, want):
    assert got.shape == want.shape
    assert np.array_equal(np.isfinite(got), np.isfinite(want))
    m = np.isfinite(want)
    if kind == 1:  # MS-DFM: the fixed point is reached to an ulp or so whatever the order (north_star: 1e-4)
        assert (np.abs(got[m] - want[m]) / np.maximum(want[m], 1e-30)).max(initial=0.0) <= 1e-4
    else:
        bad = got[m] != want[m]
        assert not bad.any(), "%d elements differ" % bad.sum()


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("ranks,per_rank", [(1, 1), (1, 3), (2, 1), (2, 2), (3, 2)])
def test_sharded_plan_equals_single_device(kind, ranks, per_rank):
    L, W = 400, 300
    start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
    cmap = synth.synth_map(51 + kind, L, W, 10, start, goal)
    ref = single(kind, cmap, start, goal)
    want = ref.read_field()
    p = sharded(kind, ranks, per_rank)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    got = p.read_field()
    same(kind, got, want)
    ss = p.shard_stats()
    assert ss["n_ranks"] == ranks and ss["bands"] <= ranks * per_rank
    # one rank: the fused mode (border rows and activations written straight across the band edges, global phases);
    # several ranks on ONE device: the exchange iterations (persistent kernels of two ranks must not share a GPU)
    assert ss["fused"] == (1 if ranks == 1 else 0)
    assert (ss["fused_launches"] >= 1) if ss["fused"] else (ss["exchanges"] >= 1)
    if ss["fused"] and ss["bands"] > 1:
        assert ss["remote_activations"] > 0
    if kind != 1:
        assert p.field_checksum() == ref.field_checksum()
    assert np.array_equal(p.read_map(), cmap)
    w = p.read_window(100, 50, 40, 60)
    assert np.array_equal(w, got[100:140, 50:110])
    st = p.stats()
    assert st["tile_visits"] > 0
    p.close()
    ref.close()


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("mode", ["mixed", "raise", "lower"])
@pytest.mark.parametrize("ranks,per_rank", [(2, 2), (1, 4)])
def test_sharded_replan_equals_from_scratch(kind, mode, ranks, per_rank):
    """patch_map + step on the sharded grid (invalidate across bands, then re-propagate) == a single-device plan of the
    patched map; the patch straddles band boundaries."""
    L, W = 420, 260
    start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
    cmap = synth.synth_map(61, L, W, 10, start, goal)
    p = sharded(kind, ranks, per_rank)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    rng = np.random.default_rng(7)
    x, y, h, w = 90, 60, 180, 120   # rows 90..269: crosses several of the 128-row bands
    old = cmap[x:x + h, y:y + w]
    if mode == "mixed":
        patch = synth.synth_rows(99, 0, h, w, 15)
    elif mode == "raise":
        patch = np.where(rng.random(old.shape) < 0.1, 255, np.minimum(old.astype(np.int32) + 9, 255)).astype(np.uint8)
    else:
        patch = np.maximum(old.astype(np.int32) - 9, 1).astype(np.uint8)
    changed = p.patch_map(patch, x, y)
    assert changed == int((patch != old).sum())
    patched = cmap.copy()
    patched[x:x + h, y:y + w] = patch
    p.set_start(8.0, 8.0)
    assert p.step() == 0
    ref = single(kind, patched, (8.0, 8.0), goal)
    same(kind, p.read_field(), ref.read_field())
    assert np.array_equal(p.read_map(), patched)
    if mode != "lower":
        ss = p.shard_stats()
        assert ss["inval_exchanges"] >= 1 or ss["fused_launches"] >= 2
    # a second patch right away, and a step without a new start in between changes nothing (ReplannerBase.h:56-59)
    patch2 = np.full((30, 40), 3, np.uint8)
    p.patch_map(patch2, 120, 100)
    assert p.step() == 0
    same(kind, p.read_field(), ref.read_field())
    p.set_start(9.0, 9.0)
    assert p.step() == 0
    patched[120:150, 100:140] = patch2
    ref2 = single(kind, patched, (9.0, 9.0), goal)
    same(kind, p.read_field(), ref2.read_field())
    p.close(); ref.close(); ref2.close()


@pytest.mark.parametrize("kind", [0, 2])
def test_sharded_walk_equals_single_device_walk(kind):
    L, W = 380, 300
    start, goal = (200.0, 40.0), (float(L - 6), float(W - 6))   # the start sits next to a band boundary (row 192)
    cmap = synth.synth_map(71, L, W, 10, start, goal)
    ref = single(kind, cmap, start, goal)
    want, _ = ref.extract_paths(None, None, 20)
    p = sharded(kind, 2, 3)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    got, _ = p.extract_paths(None, None, 20)
    assert np.array_equal(got[0]["path"], want[0]["path"])
    assert got[0]["total_cost"] == want[0]["total_cost"] and got[0]["status"] == want[0]["status"]
    p.close(); ref.close()


@pytest.mark.parametrize("kind", KINDS)
def test_fused_and_iteration_modes_agree(kind):
    L, W = 500, 333
    start, goal = (250.0, 160.0), (10.0, 300.0)   # a goal in the first band: fronts run up AND down the stack
    cmap = synth.synth_map(81, L, W, 12, start, goal)
    ref = single(kind, cmap, start, goal)
    want = ref.read_field()
    for fused in (True, False):
        p = sharded(kind, 1, 5)
        p.set_fused(fused)
        p.set_map(cmap)
        p.set_start(*start)
        p.set_goal(*goal)
        assert p.step() == 0
        assert p.shard_stats()["fused"] == int(fused)
        same(kind, p.read_field(), want)
        p.close()
    ref.close()


def test_sharded_flags_and_errors():
    from unige_tasi_path_planners_b200 import ShardedPlanner, TasiError
    p = ShardedPlanner(0, devices=[0, 0])
    assert p.step() == -1
    p.set_map(np.ones((100, 80), np.uint8))
    assert p.step() == -2
    p.set_goal(90.0, 70.0)
    p.set_start(2.0, 2.0)
    assert p.step() == 0
    assert p.field_dims() == (101, 81)
    with pytest.raises(TasiError):
        p.set_focus(True)
    with pytest.raises(TasiError):
        p.band_begin()
    # same goal element: no re-initialisation; a new goal re-plans (ReplannerBase.h:99-108)
    f0 = p.read_field()
    p.set_goal(90.2, 70.1)
    assert p.step() == 0 and p.num_nodes_expanded == 0
    p.set_goal(10.0, 10.0)
    assert p.step() == 0 and p.num_nodes_expanded > 0
    assert p.read_field()[10, 10] == 0.0 and f0[90, 70] == 0.0
    p.close()
    with pytest.raises(TasiError):
        ShardedPlanner(0, devices=[0, 0]).set_map(np.ones((3, 3), np.uint8))  # too small for two bands of rows... one tile
