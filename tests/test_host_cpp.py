"""The C++ host side (unige_tasi_path_planners_b200/host/GpuPlanners.h): the planner classes with the
reference's ReplannerBase surface, over the C ABI.

CPU part: the classes compile stand-alone and fail loudly without a device. GPU part: host/plan_demo (plan +
patch + replan through the classes) gives the oracle's field; oracle/_ref/ref_dropin_* (built in the build
container from the reference's sources, travels as a binary) runs the UNMODIFIED reference extractor and demo
loop over the GPU planner classes and over the reference's CPU planner side by side.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import hostops

from unige_tasi_path_planners_b200 import build, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KINDS = {"fd": 0, "dfm": 1, "mfd": 2}
DROPIN = {"fd": "ref_dropin_fd", "dfm": "ref_dropin_dfm", "mfd": "ref_dropin_sg"}


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="module")
def demo():
    build.build()
    return build.build_host_demo()


def test_host_classes_compile_and_link(demo):
    assert os.access(demo, os.X_OK)
    r = subprocess.run([demo], capture_output=True, text=True)
    assert r.returncode == 64 and "usage" in r.stderr


def test_host_classes_have_no_cpu_fallback(demo):
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([demo, "fd", "1", "32", "32", "1"], capture_output=True, text=True)
    assert r.returncode == 70
    assert "no CPU fallback" in r.stderr


def test_header_mirrors_the_reference_surface():
    """Every public member of ReplannerBase (ReplannerBase.h:39-123, 144-156) the demos touch exists by name."""
    src = open(os.path.join(ROOT, "unige_tasi_path_planners_b200", "host", "GpuPlanners.h")).read()
    for name in ["void reset()", "int step()", "void set_occupancy_threshold(float", "void set_heuristic_multiplier(float",
                 "void set_map(const std::shared_ptr<uint8_t> &", "void patch_map(const std::shared_ptr<uint8_t> &",
                 "void set_start(const Position &", "void set_goal(const Position &", "get_expanded_map()", "get_grid()",
                 "u_time", "p_time", "num_nodes_updated", "num_nodes_expanded", "heuristic_multiplier",
                 "initialize_graph", "initialize_search", "goal_set", "new_goal", "new_start",
                 "class GpuFieldDPlanner", "class GpuDFMPlanner", "class GpuShiftedGridPlanner"]:
        assert name in src, name


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["fd", "dfm", "mfd"])
def test_plan_demo_matches_oracle(demo, oracle, kind, tmp_path):
    L, W, seed = 150, 210, 17
    px, py, ph, pw, pseed = 40, 50, 30, 44, 99
    dump = str(tmp_path / "f")
    r = subprocess.run([demo, kind, "1", str(L), str(W), str(seed), "--patch", str(px), str(py), str(ph), str(pw),
                        str(pseed), "--dump", dump], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = [json.loads(l) for l in r.stdout.splitlines()]
    assert [l["step"] for l in lines] == ["plan", "replan"]
    k = KINDS[kind]
    start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
    cmap = synth.synth_map(seed, L, W, 10, start, goal)
    shape = oracle.field_shape(k, L, W)
    for what, cm in (("plan", cmap), ("replan", None)):
        if cm is None:
            cm = cmap.copy()
            oracle.patch(cm, synth.synth_rows(pseed, 0, ph, pw, 15), px, py)
        got = np.fromfile(dump + "." + what, np.float32).reshape(shape)
        want = oracle.solve(k, cm, goal)
        assert np.array_equal(np.isfinite(got), np.isfinite(want))
        m = np.isfinite(want)
        if k == 1:
            assert (np.abs(got[m] - want[m]) / np.maximum(want[m], 1e-30)).max() <= 1e-4
        else:
            assert np.array_equal(got[m], want[m])
        # GpuPathExtractor (device walk, reference defaults) == the same walk on the host over the dumped field
        from unige_tasi_path_planners_b200 import _lib
        path = np.fromfile(dump + "." + what + ".path", np.float32).reshape(-1, 2)
        hw = hostops.host_extract(k == 1, got, cm, start if what == "plan" else (8.0, 8.0), goal, 255, 20)
        assert np.array_equal(path, hw[0])
        if oracle.ref_available() and k != 1:  # the reference's own extractor over the dumped field (oracle/_ref)
            rp = oracle.RefPlanner(k, 1)
            rp.set_map(cm)
            rp.set_start(*(start if what == "plan" else (8.0, 8.0)))
            rp.set_goal(*goal)
            xy, _, _, _ = rp.extract(field=got, max_steps=20)
            rp.close()
            assert np.array_equal(xy, path)
        line = lines[0 if what == "plan" else 1]
        assert line["path_points"] == len(path) and line["path_cost"] == pytest.approx(hw[2], rel=1e-6)
    assert lines[1]["num_nodes_updated"] > 0 and lines[0]["tile_visits"] > 0
    # FieldMap hand-off: get_rhs(start element) is the field value there
    ex = (5, 5)
    assert lines[0]["start_value"] == pytest.approx(float(np.fromfile(dump + ".plan", np.float32).reshape(shape)[ex]))


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["fd", "mfd"])
@pytest.mark.parametrize("devices", ["0,0", "0,0,0"])
def test_plan_demo_over_several_ranks_equals_one_device(demo, kind, devices, tmp_path):
    """GpuFieldDPlanner<1>(devices) — the sharded planner behind the same class — plans and replans the same field, walks
    the same path and reports the same counters of the reference's interface as the one-device object."""
    L, W, seed = 300, 210, 23
    args = [kind, "1", str(L), str(W), str(seed), "--patch", "110", "50", "60", "44", "99"]
    outs = []
    for extra, tag in ((["--device", "0"], "one"), (["--devices", devices], "many")):
        dump = str(tmp_path / tag)
        r = subprocess.run([demo] + args + ["--dump", dump] + extra, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        outs.append(([json.loads(l) for l in r.stdout.splitlines()], dump))
    (one, d1), (many, d2) = outs
    for what, a, b in (("plan", one[0], many[0]), ("replan", one[1], many[1])):
        assert np.array_equal(np.fromfile(d1 + "." + what, np.float32), np.fromfile(d2 + "." + what, np.float32))
        assert np.array_equal(np.fromfile(d1 + "." + what + ".path", np.float32),
                              np.fromfile(d2 + "." + what + ".path", np.float32))
        for key in ("finite", "sum", "start_value", "start_interp_rhs", "path_points", "path_cost", "path_dist",
                    "num_nodes_updated"):
            assert a[key] == b[key], (what, key)


@pytest.mark.gpu
@pytest.mark.parametrize("kind,level", [("fd", 1), ("fd", 0), ("mfd", 1), ("dfm", 1)])
def test_reference_extractor_and_demo_loop_over_gpu_planner(kind, level):
    """The reference's own extractor + demo loop: CPU planner vs GPU planner class, same patches, step by step."""
    exe = os.path.join(ROOT, "oracle", "_ref", DROPIN[kind])
    if not os.access(exe, os.X_OK):
        pytest.skip("oracle/_ref drop-in binaries not built (the reference sources are absent)")
    r = subprocess.run([exe, str(level), "96", "120", "5", "40", "6"], capture_output=True, text=True, timeout=600)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, r.stderr
    steps = [l for l in lines if "iter" in l]
    assert len(steps) >= 5
    # third arm: the device walk (GpuPathExtractor) equals the reference extractor's walk over the same GPU field
    assert all(s["device_walk_identical"] for s in steps)
    if kind == "dfm":
        # MS-DFM fields agree to ~1e-7, not bit for bit, and the extractor's edge interpolation amplifies an ulp
        # into ~1e-3 of a cell: same number of points, each within 0.05 cell, path cost within 1e-3 (north_star)
        for s in steps:
            assert s["ref_points"] == s["gpu_points"]
            assert s["max_point_diff"] <= 0.05
            assert abs(s["ref_cost"] - s["gpu_cost"]) <= 1e-3 * abs(s["ref_cost"])
    else:
        assert r.returncode == 0, r.stdout[-2000:]
        assert all(s["identical"] for s in steps)
        assert lines[-1]["mismatching_steps"] == 0
