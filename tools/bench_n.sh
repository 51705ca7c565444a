#!/bin/bash
# tools/bench_n.sh N [extra bench args]: run bench.py on N GPUs the way the driver does, print a short summary
N=$1; shift
if [ "$N" = 1 ]; then CMD="python bench.py --gpus 1"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N"; fi
$CMD "$@" 2>gpurun_out/bench_n.err | tail -1 > gpurun_out/bench_n_last.json
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_n_last.json').read())
print('N', d['n_gpus'], 'rep', d['config'].get('replicate'), 'G edges/s', round(d['value']/1e9,3), 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']/1e9,3),
      'roofline', d['roofline']['bound'], d['roofline']['frac'], 'feature ms', d['stages']['feature']['ms'], 'sample ms', round(sum(v['ms'] for k,v in d['stages'].items() if k!='feature'),3))
print(d['roofline']['note'])
P
