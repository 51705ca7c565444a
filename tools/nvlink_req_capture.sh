#!/bin/bash
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/nvlink_req_probe tools/nvlink_req_probe.cu || exit 1
timeout 120 /tmp/nvlink_req_probe > gpurun_out/r02n_nvlink_req_probe.txt 2>&1 || { cat gpurun_out/r02n_nvlink_req_probe.txt; exit 1; }
timeout 120 /tmp/nvlink_req_probe --local >> gpurun_out/r02n_nvlink_req_probe.txt 2>&1
cat gpurun_out/r02n_nvlink_req_probe.txt
timeout 300 ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
  --log-file gpurun_out/r02n_nvlink_req_ncu.csv /tmp/nvlink_req_probe --one-way > gpurun_out/r02n_ncu_stdout.txt 2>&1
