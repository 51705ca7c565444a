#!/bin/bash
# minimal re-capture of the gather's DRAM traffic after a source change (plain run first, then one ncu launch)
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-secondary --stage-steps 1"
timeout 60 $CMD > gpurun_out/r02u_short.json 2>/dev/null || exit 1
timeout 80 ncu --set full --clock-control none --import-source on -k regex:gather_rows_ring -s 3 -c 1 -f -o gpurun_out/r02u_gather $CMD > gpurun_out/r02u_ncu.log 2>&1
tail -2 gpurun_out/r02u_ncu.log
