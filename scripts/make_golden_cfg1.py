"""Generates tests/golden/cfg1_maps.npz (the four Tests/Tests/*.bmp fixtures as gray images + the endpoints their file
names carry) and tests/golden/cfg1_trace.npz (what the reference's own demo process, oracle/_ref/field_d_planner_1_no_heur,
reported at every step of the headless drive of tests/cfg1_drive.py). Run in the build container (needs /root/reference
and `make -C oracle ref`).

    python scripts/make_golden_cfg1.py
"""
import os
import re
import sys
import tempfile

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = os.environ.get("TASI_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "tests", "golden")


def main():
    d = os.path.join(REF, "Tests", "Tests")
    maps = {"names": []}
    for f in sorted(os.listdir(d)):
        m = re.match(r"(.+)_(\d+)_(\d+)_(\d+)_(\d+)_\.bmp$", f)
        if not m:
            continue
        name = m.group(1)
        maps["names"].append(name)
        maps["img_" + name] = cv2.imread(os.path.join(d, f), cv2.IMREAD_GRAYSCALE)
        maps["ends_" + name] = np.array([float(v) for v in m.groups()[1:]], np.float32)
    maps["names"] = np.array(maps["names"])
    np.savez_compressed(os.path.join(OUT, "cfg1_maps.npz"), **maps)
    import cfg1_drive as D
    trace = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, (img, start, goal) in D.load_maps().items():
            dr = D.drive(D.ref_cmd(name, start, goal), img, tmp, "ref")
            print(name, img.shape, start, goal, "steps", len(dr.steps), "ended", dr.ended, "stuck", dr.stuck, "rc", dr.returncode)
            trace["n_" + name] = np.array([len(dr.steps), int(dr.ended), int(dr.stuck)])
            for i, s in enumerate(dr.steps):
                g = D.digest(s)
                k = "%s_%03d_" % (name, i)
                trace[k + "pos"] = np.array(g["pos"], np.float32)
                trace[k + "path"] = g["path"]
                trace[k + "costs"] = g["costs"]
                trace[k + "tot"] = np.array([g["cost"], g["dist"]], np.float32)
                trace[k + "xy"] = g["closed_xy"]
                trace[k + "crc"] = np.array([g["closed_crc"]], np.uint32)
    np.savez_compressed(os.path.join(OUT, "cfg1_trace.npz"), **trace)
    print("sizes", {f: os.path.getsize(os.path.join(OUT, f)) for f in ("cfg1_maps.npz", "cfg1_trace.npz")})


if __name__ == "__main__":
    main()
