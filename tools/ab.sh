#!/bin/bash
# A/B harness: runs bench.py once per environment assignment given as arguments ("" = default) and prints ms/step + stages.
# usage (on the GPU box): tools/ab.sh "" "XG_TABLE_PERSIST_MB=64" ...   (XG_AB_ARGS = extra bench.py arguments)
for cfg in "$@"; do
  out=$(env $cfg python bench.py --steps 60 --warmup 10 --no-cpu-baseline --no-secondary --stage-steps 10 $XG_AB_ARGS 2>/dev/null | tail -1)
  python - "$cfg" <<PY
import json,sys
d=json.loads('''$out''')
st={k:v["ms"] for k,v in d["stages"].items() if not k.startswith("remap")}
print(f"{sys.argv[1] or 'default':40s} ms/step {d['ms_per_step']:.4f}  e2e {d['e2e']['ms_per_step']:.4f}  sample_sum {sum(v for k,v in st.items() if k not in ('feature','label')):.4f}  {st}")
PY
done
