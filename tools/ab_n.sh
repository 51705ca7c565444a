#!/bin/bash
# A/B harness at N GPUs: tools/ab_n.sh N "" "XG_GATHER_RING_STAGES=6" ...
N=$1; shift
for cfg in "$@"; do
  out=$(env $cfg python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 60 --warmup 10 --no-secondary --stage-steps 10 2>/dev/null | tail -1)
  python - "$cfg" <<PY
import json,sys
d=json.loads('''$out''')
st={k:v["ms"] for k,v in d["stages"].items()}
r=d["roofline"]
print(f"{sys.argv[1] or 'default':50s} G/s {d['value']/1e9:.3f} ms/step {d['ms_per_step']:.4f} gather {st['feature']:.4f} frac {r['frac']} sample_sum {sum(v for k,v in st.items() if k not in ('feature','label')):.4f}")
PY
done
