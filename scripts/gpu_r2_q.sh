#!/bin/bash
# round 2, visit Q (1 GPU): key-ordered release of the parked tiles (TASI_REL_PASSES 1 | 2 | 4 | 8)
mkdir -p gpurun_out
for rp in 1 4 8 2; do
TASI_REL_PASSES=$rp timeout 600 python scripts/quick_bench.py 1024 4096 16384 > gpurun_out/quick_rel$rp.jsonl 2>&1; echo "quick rel=$rp rc=$?"
done
python - <<'PY'
import json
for c in (1, 2, 4, 8):
    for l in open("gpurun_out/quick_rel%d.jsonl" % c):
        try: d = json.loads(l)
        except Exception: print(l.strip()[:200]); continue
        print("rel", c, "kind", d["kind"], d["N"], "ms %.2f" % d["kernel_ms_solve"], "visits/tile %.2f" % d["visits_per_tile"], "ev/el %.1f" % d["evals_per_elem"],
              "cyc/visit %.0f" % ((d["cyc_pop"] + d["cyc_stage"] + d["cyc_relax"] + d["cyc_post"]) / max(d["tile_visits"], 1)), "pop %.2f" % (d["cyc_pop"] / (d["cyc_pop"] + d["cyc_stage"] + d["cyc_relax"] + d["cyc_post"])))
PY
TASI_REL_PASSES=4 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_focus.py -m gpu -x -q 2>&1 | tail -2
