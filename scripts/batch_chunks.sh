#!/bin/bash
# development probe: cfg5 sample with different stack sizes (queries per multi-goal plan)
for c in 64 128 32; do echo "chunk $c"; TASI_BATCH_CHUNK=$c python bench.py --workload cfg5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print({k:(round(v['queries_per_s']),round(v['seconds'],3)) for k,v in d['per_planner'].items()})"; done
