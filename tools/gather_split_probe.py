#!/usr/bin/env python
"""Probe (torchrun, N >= 2): does the routed gather get faster when every 32-row group is homogeneous (all local or
all peer rows) instead of mixed?  Same rows, same kernel; only the ORDER of the node list changes.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/gather_split_probe.py --replicate 0.4"""
import argparse, os, sys, json
from pathlib import Path
import numpy as np
import torch
import torch.distributed as dist
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench
from xgnn_b200 import dataset as D, ops

ap = argparse.ArgumentParser()
ap.add_argument("--replicate", type=float, default=0.6)
a = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
st = D.build_store("papers100M", local, rank, world, replicate_pct=a.replicate, replicate_topology=True)
ptrs = bench.exchange_parts(st, dist, torch)
D.attach(st, None, None, ptrs["feat_part"])
sm = ops.MinibatchSampler(st.graph, bench.FANOUT, 8000, seed=1, num_slots=2, device=local)
seeds = bench.seed_batches(st.train, rank, world, 8000, 3)
res = {}
for b in range(3):
    mb = sm.sample(seeds[b], batch_key=b)
    nodes = mb.input_nodes.clone()
    mb.release()
    table = st.local["cache_table"]
    r = (table[nodes.long()].to(torch.int64) & 0xFFFFFFFF)
    n_rep = st.local.get("num_replicated", 0)
    peer = (r >= n_rep) & (((r - n_rep) % world) != rank)
    order = torch.argsort(peer.to(torch.int8), stable=True)
    variants = {"mixed": nodes, "split": nodes[order].contiguous()}
    out = torch.empty((nodes.numel(), st.dim), dtype=st.dtype, device="cuda")
    # "concurrent": the two classes as two launches on two streams AT THE SAME TIME (rank + destination lists per class)
    pos = torch.arange(nodes.numel(), device="cuda", dtype=torch.int32)
    lists = [(r[m].to(torch.int32).contiguous(), pos[m].contiguous()) for m in (~peer, peer)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    ts = []
    for it in range(6):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for sx, (src, dst) in zip(streams, lists):
            sx.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(sx):
                ops.CombineCacheData(out, st.store, src, dst)
        for sx in streams:
            torch.cuda.current_stream().wait_stream(sx)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res.setdefault("concurrent", []).append(min(ts[1:]))
    want = torch.empty_like(out)
    ops.ExtractGGMS(st.store, nodes, st.dim, st.dtype, out=want)
    assert torch.equal(out.view(torch.int32), want.view(torch.int32)), "concurrent two-class gather differs from the mixed one"
    for name, nd in variants.items():
        ts = []
        for it in range(6):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.ExtractGGMS(st.store, nd, st.dim, st.dtype, out=out); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        res.setdefault(name, []).append(min(ts[1:]))
    res.setdefault("peer_frac", []).append(float(peer.float().mean()))
t = torch.tensor([np.mean(res["mixed"]), np.mean(res["split"]), np.mean(res["concurrent"])], device="cuda", dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"world": world, "replicate": a.replicate, "peer_frac": np.mean(res["peer_frac"]), "mixed_ms": t[0].item(), "split_ms": t[1].item(), "concurrent_ms": t[2].item(),
                      "ring": [os.environ.get("XG_GATHER_RING_CTAS", "3"), os.environ.get("XG_GATHER_RING_STAGES", "4")]}))
dist.destroy_process_group()
