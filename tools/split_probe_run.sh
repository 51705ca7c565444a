set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR tools/gather_split_probe.py --replicate 0.4 2>&1 | grep -v Warning | tail -3 > gpurun_out/split_A.log
XG_GATHER_RING_CTAS=1 XG_GATHER_RING_STAGES=6 $TR tools/gather_split_probe.py --replicate 0.4 2>&1 | tail -3 > gpurun_out/split_B.log
XG_GATHER_RING_CTAS=2 XG_GATHER_RING_STAGES=3 $TR tools/gather_split_probe.py --replicate 0.4 2>&1 | tail -3 > gpurun_out/split_C.log
cat gpurun_out/split_A.log gpurun_out/split_B.log gpurun_out/split_C.log
