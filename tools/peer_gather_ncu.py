#!/usr/bin/env python
"""ONE process, two devices: the routed gather on cuda:0 with a fifth of its rows in cuda:1's HBM (read through P2P
loads inside the kernel), sized like the headline minibatch (4.9 M rows x 512 B out of a 111 M-row store).  Made to run
under ncu (one process, one profiled GPU) for the NVLink-side counters of the mixed gather - the multi-rank bench
cannot be profiled:
  python tools/peer_gather_ncu.py            # prints CUDA-event times
  ncu --set full --clock-control none -k regex:gather_rows_ring -s 2 -c 1 -o X python tools/peer_gather_ncu.py
Placement = what bench.py --gpus N builds: hottest `--replicate` of the ranks in local HBM (laid out by node id,
resolved by rank/select words), the rest sharded rank % 2 over {cuda:0, cuda:1}.  Node ids are uniform, so the peer
share is (1 - replicate) / 2.  Row CONTENTS are uninitialised memory: this is a traffic probe, parity is tested elsewhere.
"""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from xgnn_b200 import ops  # noqa: E402
from xgnn_b200._clib import check, lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nodes", type=int, default=111_059_956)
ap.add_argument("--rows", type=int, default=4_918_000)
ap.add_argument("--dim", type=int, default=128)
ap.add_argument("--replicate", type=float, default=0.6)
ap.add_argument("--iters", type=int, default=6)
a = ap.parse_args()

check(lib.xg_enable_peer_access(0, 1))
check(lib.xg_enable_peer_access(1, 0))
torch.cuda.set_device(0)
V, dim = a.nodes, a.dim
n_rep = int(V * a.replicate)
rest = V - n_rep
rows = [(rest - p + 1) // 2 for p in range(2)]
replica = torch.empty((n_rep, dim), dtype=torch.float32, device="cuda:0")
parts = [torch.empty((max(rows[0], 1), dim), dtype=torch.float32, device="cuda:0"),
         torch.empty((max(rows[1], 1), dim), dtype=torch.float32, device="cuda:1")]
table = torch.arange(V, dtype=torch.int32, device="cuda:0")  # rank == node id: both tiers are laid out by node id
words = ops.build_replica_words(table, n_rep)
store = ops.make_feature_store(dim * 4, None, table, parts, local_parts=(0,), replica=replica, num_replicated=n_rep,
                               replica_words=words)
g = torch.Generator(device="cuda:0").manual_seed(7)
nodes = torch.randperm(V, generator=g, device="cuda:0", dtype=torch.int64)[: a.rows].to(torch.int32).contiguous()
out = torch.empty((a.rows, dim), dtype=torch.float32, device="cuda:0")
stats = torch.zeros(4, dtype=torch.int64, device="cuda:0")
ops.ExtractGGMS(store, nodes, dim, torch.float32, stats=stats, out=out)  # statistics variant (register-staged kernel)
torch.cuda.synchronize()
c = stats.tolist()
ts = []
for _ in range(a.iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.ExtractGGMS(store, nodes, dim, torch.float32, out=out)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
rb = dim * 4
local_b, peer_b = c[0] * rb, c[1] * rb
hbm_ms = (local_b + a.rows * rb + a.rows * 4) / 6544.7e6
nvl_ms = peer_b / 770e6
print(json.dumps({"rows": a.rows, "row_bytes": rb, "local": c[0], "peer": c[1], "host": c[2], "ms": [round(t, 4) for t in ts],
                  "best_ms": min(ts[1:]), "roofline_ms": {"hbm": round(hbm_ms, 4), "nvlink_in_770": round(nvl_ms, 4)},
                  "peer_GBps": round(peer_b / min(ts[1:]) / 1e6, 1), "cover_all": bool(getattr(words, "cover_all", False))}))
