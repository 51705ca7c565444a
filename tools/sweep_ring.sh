#!/bin/bash
run() {
  env "$@" python bench.py --steps 60 --warmup 6 --no-cpu-baseline --stage-steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$*', 'ms/step', round(d['ms_per_step'],4), 'feature', d['stages']['feature']['ms'], 'G edges/s', round(d['value']/1e9,3))"
}
for st in 2 3 4; do for c in 1 2; do run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=$c XG_GATHER_RING_STAGES=$st; done; done
run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=3 XG_GATHER_RING_STAGES=2
run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=4 XG_GATHER_RING_STAGES=2
