#!/bin/bash
run() {
  env "$@" python bench.py --steps 60 --warmup 6 --no-cpu-baseline --stage-steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$*', 'ms/step', round(d['ms_per_step'],4), {k:v['ms'] for k,v in d['stages'].items()})"
}
run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=2 XG_GATHER_RING_STAGES=3
run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=2 XG_GATHER_RING_STAGES=3 XG_SINGLE_STREAM=1
run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=2 XG_GATHER_RING_STAGES=3 XG_L2_FETCH=32
run XG_GATHER_VARIANT=2 XG_GATHER_RING_CTAS=2 XG_GATHER_RING_STAGES=3 XG_L2_FETCH=128
