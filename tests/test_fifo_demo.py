"""host/fifo_demo: the demos' planner process over the GPU planners, driven through the reference's FIFO wire
protocol (Tests/Planners/FDSTAR/main.cpp:49-175; the simulator side of Simulator/simulator/run_simulator.py:38-113
is played here). CPU part: builds, prints usage, exits 70 without a device. GPU part: a whole drive to the goal;
every step's path must equal the walk over the oracle's converged field of the map known at that step.
"""
import os
import struct
import subprocess
import threading

import numpy as np
import pytest

import hostops

from unige_tasi_path_planners_b200 import _lib, build, synth

KINDS = {"fd": 0, "dfm": 1, "mfd": 2}


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="module")
def exe():
    build.build()
    return build.build_fifo_demo()


def test_fifo_demo_builds_and_prints_usage(exe):
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 64 and "usage" in r.stderr


def test_fifo_demo_has_no_cpu_fallback(exe):
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([exe, "fd", "1", "/nonexistent/in", "/nonexistent/out", "1", "1", "5", "5"],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 70 and "no CPU fallback" in r.stderr


class Simulator:
    """The simulator's half of the protocol: reveals a square of the true world around the reported position."""

    def __init__(self, to_planner, from_planner, world, known, radius, endpoints=None, tof=False):
        self.w = open(to_planner, "wb", buffering=0)     # the planner opens its input first (main.cpp:52-53)
        self.r = open(from_planner, "rb", buffering=0)
        self.world, self.known, self.radius = world, known, radius
        self.endpoints, self.tof = endpoints, tof
        self.steps = []      # (position, known map at that step, path, piece costs, totals, times, tof records)

    def get(self, n):
        b = b""
        while len(b) < n:
            c = self.r.read(n - len(b))
            assert c, "planner closed the pipe"
            b += c
        return b

    def run(self):
        L, W = self.known.shape
        assert self.get(1) == b"\x00"
        self.w.write(b"\x00" + struct.pack("<ii", W, L) + self.known.tobytes() + struct.pack("<i", 1))
        if self.endpoints is not None:  # DFM/main.cpp:62-66
            self.w.write(struct.pack("<ffff", *self.endpoints) + struct.pack("<b", int(self.tof)))
        while True:
            code = self.get(1)
            if code == b"\x02":
                self.w.write(b"\x02")
                return
            assert code == b"\x01"
            x, y, step_cost = struct.unpack("<fff", self.get(12))
            cx, cy = int(np.floor(x + 0.5)), int(np.floor(y + 0.5))
            x0, y0 = max(cx - self.radius, 0), max(cy - self.radius, 0)
            x1, y1 = min(cx + self.radius + 1, L), min(cy + self.radius + 1, W)
            patch = np.ascontiguousarray(self.world[x0:x1, y0:y1])
            self.known[x0:x1, y0:y1] = patch
            self.w.write(b"\x01" + struct.pack("<iiii", x0, y0, x1 - x0, y1 - y0) + patch.tobytes() +
                         struct.pack("<i", 1))
            assert self.get(1) == b"\x03"
            n = struct.unpack("<i", self.get(4))[0]
            path = np.frombuffer(self.get(8 * n), np.float32).reshape(n, 2).copy()
            costs = np.frombuffer(self.get(4 * max(n - 1, 0)), np.float32).copy()  # receive_path reads n - 1
            tail = self.get(20)
            total_dist, total_cost, u, p, e = struct.unpack("<fffff", tail)
            recs = None
            if self.tof:
                assert self.get(1) == b"\x04"
                cnt = struct.unpack("<q", self.get(8))[0]
                recs = np.frombuffer(self.get(16 * cnt), np.dtype([("x", "<i4"), ("y", "<i4"), ("g", "<f4"),
                                                                    ("rhs", "<f4")])).copy()
            self.steps.append(((x, y), self.known.copy(), path, costs, (total_cost, total_dist), (u, p, e), recs,
                               step_cost))


@pytest.mark.gpu
@pytest.mark.parametrize("kind,focus", [("fd", False), ("mfd", False), ("dfm", False), ("fd", True)])
def test_fifo_drive_to_the_goal(exe, oracle, kind, focus, tmp_path):
    L, W = 90, 110
    start, goal = (5.0, 6.0), (80.0, 100.0)
    world = synth.synth_map(51, L, W, 12, start, goal)
    if kind == "dfm":
        # the extractor reads a cell field through 4-cell node means (ExpandedMap_impl.h:87-97): next to obstacles
        # those are INF and its walk (the reference's, bit for bit) circles in place; a smooth world lets it drive
        xx, yy = np.meshgrid(np.arange(L), np.arange(W), indexing="ij")
        world = (25 + 20 * np.sin(xx / 9.0) * np.cos(yy / 7.0)).astype(np.uint8)
    known = np.full((L, W), 20, np.uint8)
    fin, fout = str(tmp_path / "to_planner"), str(tmp_path / "from_planner")
    os.mkfifo(fin)
    os.mkfifo(fout)
    tof = kind == "fd" and not focus
    if kind == "dfm":
        cmd = [exe, "dfm", "1", fin, fout, "--max-iter", "40"]
    else:
        cmd = [exe, kind, "1", fin, fout, str(start[0]), str(start[1]), str(goal[0]), str(goal[1]), "1" if tof else "0"]
    if focus:
        cmd.append("--focus")
    proc = subprocess.Popen(cmd, stderr=subprocess.PIPE, text=True)
    sim = {}

    def play():
        s = Simulator(fin, fout, world, known, 6, endpoints=(start + goal) if kind == "dfm" else None, tof=tof)
        sim["s"] = s
        s.run()

    t = threading.Thread(target=play, daemon=True)
    t.start()
    t.join(timeout=120)
    assert not t.is_alive(), "protocol stalled"
    assert proc.wait(timeout=30) == 0, proc.stderr.read()
    steps = sim["s"].steps
    assert len(steps) >= 10
    k = KINDS[kind]
    off = 0.5 if kind == "dfm" else 0.0   # DFM/main.cpp:92 reports the cell centre
    indirect = kind != "mfd"              # SGDFM/main.cpp:97
    for i, (pos, kmap, path, costs, totals, times, recs, _) in enumerate(steps):
        p0 = (pos[0] - off, pos[1] - off)
        assert np.allclose(path[0], p0, atol=1e-5, rtol=0)  # (the float32 "+ 0.5" of the DFM report is not exactly undone)
        p0 = (float(path[0][0]), float(path[0][1]))
        want = oracle.solve(k, kmap, goal)
        hw = hostops.host_extract(k == 1, want, kmap, p0, goal, 255, 20, True, indirect)
        if k != 1:
            assert np.array_equal(path, hw[0]), "step %d" % i
            assert np.array_equal(costs[:len(hw[1])], hw[1]) and not costs[len(hw[1]):].any()
        assert totals[0] == pytest.approx(hw[2], rel=1e-3)
        assert len(costs) == len(path) - 1 and all(v >= 0 for v in times)
        if recs is not None:  # the tof record: every reached element with g == rhs == its converged value
            assert len(recs) == int(np.isfinite(want).sum())
            assert np.array_equal(recs["g"], want[recs["x"], recs["y"]]) and np.array_equal(recs["g"], recs["rhs"])
    # the rover advanced at most 5 cells per step and ended on the goal
    last = steps[-1][2]
    if kind == "dfm":  # the walk circles the goal cell (node values are 4-cell means): close is what it gets
        assert np.hypot(last[-1][0] - goal[0], last[-1][1] - goal[1]) <= 3.0
    else:
        assert any(tuple(q) == goal for q in last)
    for (a, *_), (b, *_) in zip(steps, steps[1:]):
        assert np.hypot(round(b[0] - off) - round(a[0] - off), round(b[1] - off) - round(a[1] - off)) <= 7.1
