#!/bin/bash
set -x
python tools/peer_gather_ncu.py > gpurun_out/r02m_peer_gather.json 2> gpurun_out/r02m_peer_gather.err || { tail -20 gpurun_out/r02m_peer_gather.err; exit 1; }
cat gpurun_out/r02m_peer_gather.json
ncu --set full --clock-control none --import-source on --metrics nvlrx__bytes.sum,nvltx__bytes.sum,lts__t_sectors_srcunit_ltcfabric.sum,lts__t_sectors_srcunit_tex.sum \
  -k regex:gather_rows_ring -s 2 -c 1 -f -o gpurun_out/r02m_peer_gather python tools/peer_gather_ncu.py --iters 4 > gpurun_out/r02m_ncu.log 2>&1
tail -4 gpurun_out/r02m_ncu.log
