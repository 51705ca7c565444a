#!/bin/bash
# 1-GPU sweep of the gather variants x one/two streams; prints ms/step and the feature stage time
run() {
  env "$@" python bench.py --steps 60 --warmup 6 --no-cpu-baseline --stage-steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$*', 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'feature', d['stages']['feature']['ms'], 'G edges/s', round(d['value']/1e9,3))"
}
run XG_SINGLE_STREAM=1 XG_GATHER_VARIANT=0
run XG_SINGLE_STREAM=0 XG_GATHER_VARIANT=0
run XG_SINGLE_STREAM=0 XG_GATHER_VARIANT=1
run XG_SINGLE_STREAM=1 XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=1
run XG_SINGLE_STREAM=1 XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=2
run XG_SINGLE_STREAM=0 XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=1
run XG_SINGLE_STREAM=0 XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=2
