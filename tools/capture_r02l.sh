#!/bin/bash
# default bench (both arms) + one ncu --set full capture of the gather on the same code
set -x
python bench.py > gpurun_out/r02l_bench_1gpu.json 2> gpurun_out/r02l_bench_1gpu.err || exit 1
python bench.py --impl reference > gpurun_out/r02l_bench_reference.json 2> gpurun_out/r02l_bench_reference.err || exit 1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-secondary --stage-steps 1"
$CMD > gpurun_out/r02l_short.json 2>/dev/null || exit 1
ncu --set full --clock-control none --import-source on -k regex:gather_rows_ring -s 3 -c 1 -f -o gpurun_out/r02l_gather $CMD > gpurun_out/r02l_ncu.log 2>&1
tail -3 gpurun_out/r02l_ncu.log
tail -c 600 gpurun_out/r02l_bench_1gpu.json
