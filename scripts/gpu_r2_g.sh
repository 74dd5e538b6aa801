#!/bin/bash
# round 2, visit G (1 GPU): warp-local compaction (TASI_COMPACT=2): parity tests, then timings against strips (0) and CTA list (1).
mkdir -p gpurun_out
TASI_COMPACT=2 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_config_sizes.py -m gpu -x -q > gpurun_out/pytest_compact2.log 2>&1; echo "pytest compact2 rc=$?"
tail -n 4 gpurun_out/pytest_compact2.log
TASI_COMPACT=2 timeout 600 python scripts/quick_bench.py 1024 4096 16384 > gpurun_out/quick_compact2.jsonl 2>&1; echo "quick compact=2 rc=$?"
python - <<'PY'
import json
for c in (2,):
    for l in open("gpurun_out/quick_compact%d.jsonl" % c):
        try: d = json.loads(l)
        except Exception: print(l.strip()[:200]); continue
        print(c, d["kind"], d["N"], "ms %.2f" % d["kernel_ms_solve"], "visits/tile %.2f" % d["visits_per_tile"], "rounds/visit %.1f" % d["rounds_per_visit"], "evals/elem %.1f" % d["evals_per_elem"],
              "cyc/round %.0f" % (d["cyc_relax"] / max(d["tile_rounds"], 1)), "pop %.2f stage %.2f relax %.2f post %.2f" % tuple(d[k] / (d["cyc_pop"] + d["cyc_stage"] + d["cyc_relax"] + d["cyc_post"]) for k in ("cyc_pop", "cyc_stage", "cyc_relax", "cyc_post")))
PY
