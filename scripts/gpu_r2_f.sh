#!/bin/bash
# round 2, visit F (1 GPU): compacted relaxation (TASI_COMPACT=1) against the strip relaxation: parity tests, then timings.
mkdir -p gpurun_out
TASI_COMPACT=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_config_sizes.py -m gpu -x -q > gpurun_out/pytest_compact.log 2>&1; echo "pytest compact rc=$?"
tail -n 4 gpurun_out/pytest_compact.log
for c in 0 1; do
  TASI_COMPACT=$c timeout 600 python scripts/quick_bench.py 1024 4096 16384 > gpurun_out/quick_compact$c.jsonl 2>&1; echo "quick compact=$c rc=$?"
done
python - <<'PY'
import json
for c in (0, 1):
    for l in open("gpurun_out/quick_compact%d.jsonl" % c):
        try: d = json.loads(l)
        except Exception: print(l.strip()[:200]); continue
        print(c, d["kind"], d["N"], "ms %.2f" % d["kernel_ms_solve"], "visits/tile %.2f" % d["visits_per_tile"], "rounds/visit %.1f" % d["rounds_per_visit"], "evals/elem %.1f" % d["evals_per_elem"],
              "cyc/round %.0f" % (d["cyc_relax"] / max(d["tile_rounds"], 1)), "pop %.2f stage %.2f relax %.2f post %.2f" % tuple(d[k] / (d["cyc_pop"] + d["cyc_stage"] + d["cyc_relax"] + d["cyc_post"]) for k in ("cyc_pop", "cyc_stage", "cyc_relax", "cyc_post")))
PY
