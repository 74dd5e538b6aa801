#!/bin/bash
# round 2, visit X: cfg3 repair latency against band factor / rings of the repair launches
run() { echo "== $*"; env "$@" python bench.py --workload cfg3 --steps 5 --warmup 1 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('plan %.1f' % d['initial_plan']['ms'], {k:(round(v['latency_ms'],2),round(v['device_ms'],2),v['equals_from_scratch_plan']) for k,v in d['replan'].items()})"; }
run TASI_REPAIR_BUCKETS=0
for f in 3 4 6 8 12; do run TASI_REPAIR_FACTOR=$f; done
