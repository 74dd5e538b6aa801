"""TEST INFRASTRUCTURE — BASELINE.json configs[0]: "Field D* Level 1 initial plan + replan on a Tests/Tests map via the
Tests/Planners demo (run_test.py, CPU)", headless.

The simulator's half of the drive, restated from the reference's Python (no GUI, no perf, no plots):
  * world and first belief:  simulation_data(img, filter_radius=3, low_res_penalty=10)   Tests/run_test.py:101,
                             Simulator/simulator/run_simulator.py:106-113
  * c-space of a node planner: cv2.dilate with the ellipse kernel (cspace_diam + 2)      Tests/run_test.py:94-97,102
  * every step: round_patch_update(belief, world, rover, radius=5), dilate, send the patch rectangle
                                                                                        Tests/run_test.py:134-146,
                                                                                        run_simulator.py:9-28
  * wire protocol of the FD* demo process                                               Tests/Planners/FDSTAR/main.cpp:49-175
                                                                                        run_simulator.py:38-103,153-154
  * a rover that reports the same position twice is stuck and the drive ends            Tests/run_test.py:124-131

`drive(cmd, ...)` plays that against ANY planner process speaking the protocol: the reference's own
field_d_planner_1_no_heur (oracle/_ref, compiled from the unmodified main.cpp) or host/fifo_demo (the GPU planners), and
returns the step records; `compare(ref_steps, gpu_steps)` checks the second drive against the first step by step.
The four maps travel as gray images in tests/golden/cfg1_maps.npz (scripts/make_golden_cfg1.py reads the BMPs).
"""
import os
import struct
import subprocess
import threading
import zlib

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
MAPS_NPZ = os.path.join(HERE, "golden", "cfg1_maps.npz")
REF_DEMO = os.path.join(ROOT, "oracle", "_ref", "field_d_planner_1_no_heur")
CSPACE_DIAM = 1      # Tests/run_test.py:21
SENSOR_RADIUS = 5    # Tests/run_test.py:134 (radius = 5 * upscale, upscale 1)
MAX_STEPS = 400

REC = np.dtype([("x", "<i4"), ("y", "<i4"), ("g", "<f4"), ("rhs", "<f4")])


def load_maps():
    """{name: (gray image uint8, (from_x, from_y), (to_x, to_y))} — the four Tests/Tests/*.bmp fixtures."""
    z = np.load(MAPS_NPZ)
    out = {}
    for name in z["names"]:
        name = str(name)
        ends = z["ends_" + name]
        out[name] = (z["img_" + name], (float(ends[0]), float(ends[1])), (float(ends[2]), float(ends[3])))
    return out


def simulation_data(img_h, filter_radius=3, low_res_penalty=10):
    """run_simulator.py:106-113: cost = ~pixel with 0 -> 1; the belief is the blurred image plus a penalty (saturating)."""
    img_l = cv2.GaussianBlur(img_h, (filter_radius, filter_radius), 0)
    h = ~img_h
    h = h + (h == 0).astype(np.uint8)
    lo = ~img_l
    lo = lo + (lo == 0).astype(np.uint8)
    lo = cv2.add(lo, low_res_penalty)
    return lo, h


def round_patch_update(data_l, data_h, center, radius):
    """run_simulator.py:9-28: the disc of `radius` around `center` (column, row) of the belief takes the world's values;
    returns the new belief, the (top, left) of the square around the disc clipped at the top/left border, and its slices."""
    off_x = off_y = 0
    top, bottom = center[1] - radius, center[1] + radius + 1
    left, right = center[0] - radius, center[0] + radius + 1
    if top < 0:
        off_y, top = top, 0
    if left < 0:
        off_x, left = left, 0
    data_m = data_h.copy()
    c = (left + radius + off_x, top + radius + off_y)
    data_l = data_l.copy()
    cv2.circle(data_l, c, radius, 0, cv2.FILLED)
    cv2.circle(data_m, c, radius, 0, cv2.FILLED)
    data_l = (data_h - data_m) + data_l
    return data_l, (top, left), (slice(top, bottom), slice(left, right))


class Drive:
    """One drive of one planner process. steps[i] = dict(pos, step_cost, belief (the c-space map the planner knows when
    it plans step i), path, costs, dist, cost, times (u, p, e in ms as the process reports them), recs (tof) )."""

    def __init__(self, img, to_planner, from_planner, tof=True):
        self.w = open(to_planner, "wb", buffering=0)   # the planner opens its input first (main.cpp:52-53)
        self.r = open(from_planner, "rb", buffering=0)
        self.tof = tof
        self.data_l, self.data_h = simulation_data(img)
        k = CSPACE_DIAM + 2                             # 'n'-type planners: one more in radius (run_test.py:94-97)
        self.kernel = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
        self.steps = []
        self.stuck = False
        self.ended = False

    def get(self, n):
        b = b""
        while len(b) < n:
            c = self.r.read(n - len(b))
            if not c:
                raise EOFError("planner closed the pipe")
            b += c
        return b

    def run(self):
        assert self.get(1) == b"\x00"
        cspace = cv2.dilate(self.data_l, self.kernel)
        known = cspace.copy()                           # what the planner holds: the first map + the patch rectangles
        min_cost = int(cv2.minMaxLoc(cspace)[0])
        self.w.write(b"\x00" + struct.pack("<ii", cspace.shape[1], cspace.shape[0]) + cspace.tobytes() +
                     struct.pack("<i", min_cost))
        prev = None
        while True:
            code = self.get(1)
            if code == b"\x02":
                self.w.write(b"\x02")
                self.ended = True
                return
            assert code == b"\x01", code
            x, y, step_cost = struct.unpack("<fff", self.get(12))
            if prev == (x, y) or len(self.steps) >= MAX_STEPS:
                self.stuck = prev == (x, y)
                return                                  # run_test.py:127-131 kills the process here; the caller does
            prev = (x, y)
            center = (int(round(y)), int(round(x)))
            self.data_l, pos, rng = round_patch_update(self.data_l, self.data_h, center, SENSOR_RADIUS)
            cspace = cv2.dilate(self.data_l, self.kernel)
            patch = np.ascontiguousarray(cspace[rng[0], rng[1]])
            known[rng[0], rng[1]] = patch
            min_cost = int(cv2.minMaxLoc(cspace)[0])
            self.w.write(b"\x01" + struct.pack("<iiii", pos[0], pos[1], patch.shape[0], patch.shape[1]) + patch.tobytes() +
                         struct.pack("<i", min_cost))
            assert self.get(1) == b"\x03"
            n = struct.unpack("<i", self.get(4))[0]
            path = np.frombuffer(self.get(8 * n), np.float32).reshape(n, 2).copy()
            costs = np.frombuffer(self.get(4 * max(n - 1, 0)), np.float32).copy()
            dist, cost, u, p, e = struct.unpack("<fffff", self.get(20))
            recs = None
            if self.tof:
                assert self.get(1) == b"\x04"
                cnt = struct.unpack("<q", self.get(8))[0]
                recs = np.frombuffer(self.get(16 * cnt), REC).copy()
            self.steps.append({"pos": (x, y), "step_cost": step_cost, "belief": known.copy(), "path": path,
                               "costs": costs, "dist": dist, "cost": cost, "times": (u, p, e), "recs": recs})


def replay_beliefs(img, positions):
    """The c-space map a planner holds at each step of a drive that reported `positions` (the simulator is deterministic
    given the positions): the first map with every patch rectangle applied, as Drive.run builds it."""
    data_l, data_h = simulation_data(img)
    k = CSPACE_DIAM + 2
    kernel = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
    known = cv2.dilate(data_l, kernel).copy()
    out = []
    for (x, y) in positions:
        data_l, _, rng = round_patch_update(data_l, data_h, (int(round(float(y))), int(round(float(x)))), SENSOR_RADIUS)
        known[rng[0], rng[1]] = cv2.dilate(data_l, kernel)[rng[0], rng[1]]
        out.append(known.copy())
    return out


def load_trace():
    """{name: [digest dicts]} from tests/golden/cfg1_trace.npz (the reference process's drive, scripts/make_golden_cfg1.py)."""
    z = np.load(os.path.join(HERE, "golden", "cfg1_trace.npz"))
    out = {}
    for key in z.files:
        if not key.startswith("n_"):
            continue
        name = key[2:]
        steps = []
        for i in range(int(z[key][0])):
            k = "%s_%03d_" % (name, i)
            steps.append({"pos": (float(z[k + "pos"][0]), float(z[k + "pos"][1])), "path": z[k + "path"],
                          "costs": z[k + "costs"], "cost": float(z[k + "tot"][0]), "dist": float(z[k + "tot"][1]),
                          "closed_xy": z[k + "xy"].astype(np.int64), "closed_crc": int(z[k + "crc"][0])})
        out[name] = steps
    return out


def compare_with_trace(trace_steps, steps, field_at):
    """A drive (or an oracle replay) against the committed digest of the reference's drive. field_at(i) -> dense node field
    of step i. Same bars as compare()."""
    assert len(trace_steps) == len(steps), "drives differ in length: %d vs %d" % (len(trace_steps), len(steps))
    n = 0
    for i, (a, b) in enumerate(zip(trace_steps, steps)):
        if b is not None:
            assert a["pos"] == b["pos"], "step %d: rover at %r vs %r" % (i, a["pos"], b["pos"])
            assert np.array_equal(a["path"], b["path"]), "step %d: paths differ" % i
            assert np.array_equal(a["costs"], b["costs"]), "step %d: piece costs differ" % i
            assert np.float32(a["cost"]) == np.float32(b["cost"]) and np.float32(a["dist"]) == np.float32(b["dist"]), i
        xy = a["closed_xy"]
        f = field_at(i)
        got = np.ascontiguousarray(f[xy[:, 0], xy[:, 1]], dtype=np.float32)
        assert zlib.crc32(got.tobytes()) == a["closed_crc"], "step %d: values of the reference's closed set differ" % i
        n += len(xy)
    return {"steps": len(steps), "closed_values_compared": n}


def drive(cmd_of, img, workdir, tag, tof=True, timeout=300):
    """cmd_of(fifo_in, fifo_out) -> argv of the planner process. Returns the finished Drive."""
    fin, fout = os.path.join(workdir, "to_%s" % tag), os.path.join(workdir, "from_%s" % tag)
    for f in (fin, fout):
        if os.path.exists(f):
            os.remove(f)
        os.mkfifo(f)
    proc = subprocess.Popen(cmd_of(fin, fout), stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    box = {}

    def play():
        try:
            d = Drive(img, fin, fout, tof)
            box["d"] = d
            d.run()
        except Exception as ex:  # noqa: BLE001 — reported by the caller
            box["err"] = ex

    t = threading.Thread(target=play, daemon=True)
    t.start()
    t.join(timeout=timeout)
    alive = t.is_alive()
    d = box.get("d")
    if alive or (d is not None and not d.ended):
        proc.kill()
    rc = proc.wait(timeout=30)
    err = proc.stderr.read().decode(errors="replace")
    if alive:
        raise RuntimeError("%s: protocol stalled\n%s" % (tag, err))
    if "err" in box:
        raise RuntimeError("%s: %r (rc %s)\n%s" % (tag, box["err"], rc, err))
    d.returncode = rc
    d.stderr = err
    return d


def ref_cmd(name, start, goal, tof=True):
    """argv of the reference's process: <map> <from_x> <from_y> <to_x> <to_y> <cspace> <fifo_in> <fifo_out> <gui> <tof> <out>
    (FDSTAR/main.cpp:16-31)."""
    return lambda fin, fout: [REF_DEMO, name, str(start[0]), str(start[1]), str(goal[0]), str(goal[1]), str(CSPACE_DIAM),
                              fin, fout, "0", "1" if tof else "0", "/tmp"]


def gpu_cmd(exe, start, goal, tof=True, extra=()):
    return lambda fin, fout: [exe, "fd", "1", fin, fout, str(start[0]), str(start[1]), str(goal[0]), str(goal[1]),
                              "1" if tof else "0"] + list(extra)


def closed_set(step, start_cell_of=None):
    """The part of the reference's tof record that is final at the moment plan() returned: consistent nodes whose value
    does not exceed the highest key of the start cell's corner nodes — end_condition() returns once that key is <= the top
    key of the queue (FieldDPlanner_impl.h:225-256, NO_HEURISTIC key = min(g, rhs)), and no entry of the queue can lower a
    value at or below its own key. Returns (x, y, g) arrays."""
    r = step["recs"]
    cx, cy = int(np.floor(step["pos"][0] + 0.5)), int(np.floor(step["pos"][1] + 0.5))
    corner = ((r["x"] == cx) | (r["x"] == cx + 1)) & ((r["y"] == cy) | (r["y"] == cy + 1))
    k = np.minimum(r["g"][corner], r["rhs"][corner])
    k = k[np.isfinite(k) & np.isfinite(r["rhs"][corner])]
    top = float(k.max()) if len(k) else -1.0
    m = (r["g"] == r["rhs"]) & np.isfinite(r["g"]) & (r["g"] <= top)
    return r["x"][m], r["y"][m], r["g"][m]


def field_of(step, shape):
    """Dense (L+1) x (W+1) node field from a GPU-side tof record (every reached node, g == rhs)."""
    f = np.full(shape, np.inf, np.float32)
    r = step["recs"]
    f[r["x"], r["y"]] = r["g"]
    return f


def digest(step):
    """What a committed trace keeps of one reference step (small): the closed set as (x, y) + a CRC of its values."""
    x, y, g = closed_set(step)
    o = np.lexsort((y, x))
    return {"pos": step["pos"], "path": step["path"], "costs": step["costs"], "cost": step["cost"], "dist": step["dist"],
            "closed_xy": np.stack([x[o], y[o]], 1).astype(np.int16), "closed_crc": zlib.crc32(g[o].tobytes())}


def compare(ref_steps, gpu_steps, shape):
    """The GPU drive against the reference drive, step by step. Returns a report dict; raises AssertionError on the first
    difference that is not allowed. Bars (BASELINE.json north star): reported positions, path points, piece costs, total
    cost and length bit-identical (FD* is bit-exact against the reference operator; the walk is the reference's, restated);
    every value of the reference's closed set bit-identical in the GPU field."""
    assert len(ref_steps) == len(gpu_steps), "drives differ in length: %d vs %d" % (len(ref_steps), len(gpu_steps))
    n_closed = 0
    for i, (a, b) in enumerate(zip(ref_steps, gpu_steps)):
        assert a["pos"] == b["pos"], "step %d: rover at %r vs %r" % (i, a["pos"], b["pos"])
        assert np.array_equal(a["belief"], b["belief"]), "step %d: the two drives saw different maps" % i
        assert np.array_equal(a["path"], b["path"]), "step %d: paths differ" % i
        assert np.array_equal(a["costs"], b["costs"]), "step %d: piece costs differ" % i
        assert np.float32(a["cost"]) == np.float32(b["cost"]) and np.float32(a["dist"]) == np.float32(b["dist"]), i
        if a["recs"] is not None and b["recs"] is not None:
            x, y, g = closed_set(a)
            f = field_of(b, shape)
            assert np.array_equal(f[x, y], g), "step %d: %d of %d closed values differ" % (
                i, int((f[x, y] != g).sum()), len(g))
            n_closed += len(g)
    return {"steps": len(ref_steps), "closed_values_compared": n_closed}
