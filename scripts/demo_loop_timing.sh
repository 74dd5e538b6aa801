#!/bin/bash
# The reference's demo loop (patch -> step -> extract -> advance) over the CPU planner and the GPU planner side by
# side (oracle/_ref/ref_dropin_*), with and without focused planning: per-step planner times and walk times.
mkdir -p gpurun_out
for fam in fd dfm sg; do
  for focus in 0 1; do
    timeout 600 oracle/_ref/ref_dropin_$fam 1 ${SIDE:-1024} ${SIDE:-1024} 5 ${ITERS:-14} ${RADIUS:-40} $focus \
      > gpurun_out/demo_${fam}_focus$focus.jsonl 2> gpurun_out/demo_${fam}_focus$focus.err
    echo "$fam focus=$focus rc=$?"
  done
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/demo_*_focus*.jsonl")):
    rows = [json.loads(l) for l in open(f) if l.startswith("{") and '"iter"' in l]
    if not rows: print(f, "no rows"); continue
    r0, rest = rows[0], rows[1:]
    avg = lambda k, i: sum(r[k][i] for r in rest) / max(len(rest), 1)
    print(f.split("/")[-1], "steps", len(rows), "identical", sum(r["identical"] for r in rows), "device_walk_identical",
          sum(r["device_walk_identical"] for r in rows),
          "| first plan ms ref %.1f gpu %.2f | replans: ref update+plan %.3f+%.3f gpu %.3f+%.3f | walk ref %.3f gpu %.3f" % (
              r0["ref_ms"][1], r0["gpu_ms"][1], avg("ref_ms", 0), avg("ref_ms", 1), avg("gpu_ms", 0), avg("gpu_ms", 1),
              sum(r["ref_walk_ms"] for r in rest) / max(len(rest), 1), sum(r["device_walk_ms"] for r in rest) / max(len(rest), 1)))
PY
