"""cfg1 (BASELINE.json configs[0]) on the GPU box: the headless drive of tests/cfg1_drive.py over the four Tests/Tests maps,
once with the reference's own process (oracle/_ref/field_d_planner_1_no_heur, CPU) and once with host/fifo_demo (the GPU
planners), compared step by step (positions, paths, piece costs, totals, the reference's closed set: bit-identical), with
the per-step update / planning / extraction times both processes report (ms) written to <out>/r02_cfg1_<map>.jsonl.

    python scripts/cfg1_profile.py [out_dir]      (default gpurun_out)
"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import cfg1_drive as D
    from unige_tasi_path_planners_b200 import build
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    exe = build.build_fifo_demo()
    have_ref = os.path.exists(D.REF_DEMO)
    summary = []
    with tempfile.TemporaryDirectory() as tmp:
        for name, (img, start, goal) in D.load_maps().items():
            shape = (img.shape[0] + 1, img.shape[1] + 1)
            g = D.drive(D.gpu_cmd(exe, start, goal), img, tmp, "gpu")
            g2 = D.drive(D.gpu_cmd(exe, start, goal, extra=["--focus"]), img, tmp, "gpuf")
            rep = D.compare_with_trace(D.load_trace()[name], g.steps, lambda i: D.field_of(g.steps[i], shape))
            r = None
            if have_ref:
                r = D.drive(D.ref_cmd(name, start, goal), img, tmp, "ref")
                rep = D.compare(r.steps, g.steps, shape)
            same_focus = len(g2.steps) == len(g.steps) and all(
                a["pos"] == b["pos"] and (a["path"].shape == b["path"].shape and (a["path"] == b["path"]).all())
                for a, b in zip(g.steps, g2.steps))
            with open(os.path.join(out, "r02_cfg1_%s.jsonl" % name), "w") as f:
                for i, s in enumerate(g.steps):
                    rec = {"map": name, "step": i, "pos": s["pos"], "path_points": int(len(s["path"])), "path_cost": s["cost"],
                           "gpu_ms": dict(zip("upe", s["times"])), "gpu_focus_ms": dict(zip("upe", g2.steps[i]["times"]))
                           if i < len(g2.steps) else None,
                           "ref_ms": dict(zip("upe", r.steps[i]["times"])) if r else None,
                           "reached_nodes": int(len(s["recs"])),
                           "ref_expanded_map": int(len(r.steps[i]["recs"])) if r else None}
                    f.write(json.dumps(rec) + "\n")
            tot = lambda steps, k: sum(s["times"][k] for s in steps)
            summary.append({"map": name, "grid": "%dx%d" % img.shape, "steps": rep["steps"],
                            "closed_values_compared": rep["closed_values_compared"], "identical": True,
                            "against": "live reference process" if r else "committed trace of the reference process",
                            "focus_drive_identical": bool(same_focus),
                            "first_plan_ms": {"gpu": g.steps[0]["times"][1], "gpu_focus": g2.steps[0]["times"][1],
                                              "ref": r.steps[0]["times"][1] if r else None},
                            "replan_ms_mean": {"gpu": (tot(g.steps[1:], 0) + tot(g.steps[1:], 1)) / max(len(g.steps) - 1, 1),
                                               "gpu_focus": (tot(g2.steps[1:], 0) + tot(g2.steps[1:], 1)) / max(len(g2.steps) - 1, 1),
                                               "ref": ((tot(r.steps[1:], 0) + tot(r.steps[1:], 1)) / max(len(r.steps) - 1, 1)) if r else None},
                            "extract_ms_mean": {"gpu": tot(g.steps, 2) / len(g.steps), "ref": tot(r.steps, 2) / len(r.steps) if r else None}})
            print(json.dumps(summary[-1]), flush=True)
    with open(os.path.join(out, "r02_cfg1_summary.json"), "w") as f:
        json.dump({"workload": "cfg1: FD* level 1, Tests/Tests maps, headless demo drive (sensor radius 5, blur 3, penalty 10)",
                   "maps": summary}, f, indent=1)


if __name__ == "__main__":
    main()
