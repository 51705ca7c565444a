#!/bin/bash
# ncu of the routed gather with 4 of 5 rows in pinned host memory (config 5 shape, 1 GPU): PCIe / DRAM byte counters
CMD="python bench.py --shape mag240M-homo --cache-pct 0.2 --steps 4 --warmup 3 --stage-steps 1 --in-flight 1 --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r02q_host_tier_short.json 2> gpurun_out/r02q_host_tier_short.err || { tail -5 gpurun_out/r02q_host_tier_short.err; exit 1; }
ncu --metrics gpu__time_duration.sum,pcie__read_bytes.sum,pcie__write_bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_aperture_sysmem.sum,lts__t_sectors_srcunit_tex_aperture_sysmem_op_read.sum \
  --clock-control none -k regex:gather_rows -s 3 -c 2 --csv --log-file gpurun_out/r02q_host_tier_ncu.csv $CMD > /dev/null 2>&1
grep -v "^==" gpurun_out/r02q_host_tier_ncu.csv | cut -d, -f5,13,14,15 | head -20
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02q_host_tier_short.json').read().strip().splitlines()[-1])
print(d["roofline"]["note"]); print(d["stages"]["feature"], d["roofline"]["frac"])
PY
