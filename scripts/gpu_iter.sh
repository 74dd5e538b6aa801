#!/bin/bash
# Development iteration on the GPU box: parity suite, then the per-planner timing probe.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python scripts/quick_bench.py ${SIZES:-4096 16384} > gpurun_out/quick.log 2>&1; echo "quick rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/quick.log'):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(d['kind'], d['N'], 'ms=%.2f'%d['kernel_ms_solve'], 'Mc/s=%.0f'%d['mcells_s'], 'vis/tile=%.2f'%d['visits_per_tile'], 'rnd/vis=%.2f'%d['rounds_per_visit'], 'ev/el=%.1f'%d['evals_per_elem'], 'low/el=%.1f'%d['lowered_per_elem'], 'phases', d['phases'], 'cyc/visit pop %.0f stage %.0f relax %.0f post %.0f' % tuple(d[k]/max(d['tile_visits'],1) for k in ('cyc_pop','cyc_stage','cyc_relax','cyc_post')), 'cyc/round %.0f' % (d['cyc_relax']/max(d['tile_rounds'],1)), 'adv_ms %.2f' % (d['cyc_advance']/1.965e6), 'blk', d['solver_blocks'], 'rc', d['rc'])
PY
