#!/usr/bin/env python
"""Probe 3: the bench's own 2-GPU store; time the feature gather of a real minibatch's input nodes
standalone (one rank / both ranks), and with the same nodes permuted / replaced by uniform random ones."""
import ctypes
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from xgnn_b200 import dataset as D  # noqa: E402
from xgnn_b200 import ops  # noqa: E402
from xgnn_b200._clib import check, lib  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    scale = int(os.environ.get("PROBE_SCALE", "1"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    st = D.build_store("papers100M", local, rank, world, scale=scale)
    ptrs = bench.exchange_parts(st, dist, torch)
    D.attach(st, ptrs["indptr"], ptrs["indices"], ptrs["feat_part"])
    sm = ops.MinibatchSampler(st.graph, bench.FANOUT, 8000, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                              label=st.label, seed=2024, num_slots=2, device=local)
    seeds = bench.seed_batches(st.train, rank, world, 8000, 2)[1]
    mb = sm.wait(sm.submit(torch.from_numpy(seeds.view(np.int32)).cuda(), batch_key=1))
    nodes = mb.input_nodes.clone()
    n = nodes.numel()
    mb.release()
    out_buf = torch.empty((n, st.dim), dtype=st.dtype, device="cuda")
    perm = nodes[torch.randperm(n, device="cuda")]
    rnd = torch.randint(0, st.num_node, (n,), device="cuda", dtype=torch.int64).to(torch.int32)
    srt = torch.sort(nodes.to(torch.int64) & 0xFFFFFFFF)[0].to(torch.int32)
    stream = torch.cuda.current_stream().cuda_stream

    def gather(idx):
        check(lib.xg_extract_ggms(ctypes.byref(st.store), ctypes.c_void_p(idx.data_ptr()), n,
                                  ctypes.c_void_p(out_buf.data_ptr()), None, ctypes.c_void_p(stream)))

    def timed(idx, active, reps=4):
        best = None
        for _ in range(reps + 1):
            dist.barrier()
            torch.cuda.synchronize()
            if rank in active:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                gather(idx)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1)
                best = ms if best is None else min(best, ms)
        return best or 0.0

    stats = torch.zeros(3, dtype=torch.int64, device="cuda")
    check(lib.xg_extract_ggms(ctypes.byref(st.store), ctypes.c_void_p(nodes.data_ptr()), n,
                              ctypes.c_void_p(out_buf.data_ptr()), ctypes.c_void_p(stats.data_ptr()),
                              ctypes.c_void_p(stream)))
    torch.cuda.synchronize()
    if rank == 0:
        print(json.dumps({"rows": n, "local/peer/host": stats.tolist(), "scale": scale}), flush=True)
    for name, idx in (("minibatch input nodes", nodes), ("same nodes, permuted", perm), ("same nodes, sorted", srt),
                      ("uniform random nodes", rnd)):
        for active in ((0,), (0, 1)):
            ms = timed(idx, active)
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rank == 0:
                print(json.dumps({"case": name, "active_ranks": len(active), "ms_max": round(float(t.item()), 3),
                                  "rows_GBps": round(n * st.row_bytes / float(t.item()) / 1e6, 1)}), flush=True)
    dist.barrier()
    del sm
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
