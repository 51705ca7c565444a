#!/bin/bash
run() {
  env "$@" python bench.py --steps 80 --warmup 8 --no-cpu-baseline --stage-steps 4 $EXTRA 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$* $EXTRA', 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'G edges/s', round(d['value']/1e9,3))"
}
EXTRA="--in-flight 2" run XG_SAMPLE_LANES=1
EXTRA="--in-flight 3" run XG_SAMPLE_LANES=1
EXTRA="--in-flight 2" run XG_SAMPLE_LANES=2
EXTRA="--in-flight 3" run XG_SAMPLE_LANES=2
EXTRA="--in-flight 4" run XG_SAMPLE_LANES=2
EXTRA="--in-flight 3" run XG_SAMPLE_LANES=2 XG_GATHER_RING_CTAS=1 XG_GATHER_RING_STAGES=4
