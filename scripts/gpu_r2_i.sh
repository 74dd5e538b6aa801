#!/bin/bash
# round 2, visit I (1 GPU): cfg1 drive (GPU process vs the reference process, live) + per-step profile; TMA staging: parity, then timings against vector-load staging
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cfg1.py -x -q > gpurun_out/pytest_cfg1.log 2>&1; echo "cfg1 pytest rc=$?"; tail -n 3 gpurun_out/pytest_cfg1.log
timeout 600 python scripts/cfg1_profile.py gpurun_out > gpurun_out/cfg1_profile.log 2>&1; echo "cfg1 profile rc=$?"; tail -n 5 gpurun_out/cfg1_profile.log | cut -c1-700
TASI_TMA=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/pytest_tma.log 2>&1; echo "pytest tma rc=$?"
tail -n 4 gpurun_out/pytest_tma.log
TASI_COMPACT=0 TASI_TMA=1 timeout 600 python scripts/quick_bench.py 1024 4096 16384 > gpurun_out/quick_tma1.jsonl 2>&1; echo "quick tma=1 rc=$?"
TASI_COMPACT=0 TASI_TMA=0 timeout 600 python scripts/quick_bench.py 1024 4096 16384 > gpurun_out/quick_tma0.jsonl 2>&1; echo "quick tma=0 rc=$?"
python - <<'PY'
import json
for c in (0, 1):
    for l in open("gpurun_out/quick_tma%d.jsonl" % c):
        try: d = json.loads(l)
        except Exception: print(l.strip()[:200]); continue
        print("tma", c, "kind", d["kind"], d["N"], "ms %.2f" % d["kernel_ms_solve"], "visits/tile %.2f" % d["visits_per_tile"],
              "cyc_stage/visit %.0f" % (d["cyc_stage"] / max(d["tile_visits"], 1)), "cyc/visit %.0f" % ((d["cyc_pop"] + d["cyc_stage"] + d["cyc_relax"] + d["cyc_post"]) / max(d["tile_visits"], 1)))
PY
