#!/usr/bin/env python
"""NVLink peer-gather probe (2 GPUs, one box).  Answers: what limits in-kernel P2P row gathers?

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/p2p_probe.py

rank 1 owns a feature table (cudaMalloc, exported over CUDA IPC); rank 0 maps it and times
  (a) a bulk peer copy (cudaMemcpy of the mapped range),
  (b) xg_extract (gather_rows_kernel) over the mapped table with sequential and random row ids,
      for several table sizes (TLB reach) and row sizes,
  (c) the same gathers from a same-process peer allocation (cudaDeviceEnablePeerAccess) for comparison.
Prints one JSON line per measurement on rank 0.
"""
import ctypes
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from xgnn_b200 import ops  # noqa: E402
from xgnn_b200._clib import check, lib  # noqa: E402


def view_u8(ptr, nbytes):
    return torch.as_tensor(ops._DevMem(ptr, (nbytes,), "|u1", None), device="cuda")


def time_ms(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def gather_cases(tag, src_ptr, table_bytes, out):
    n = 2_000_000
    for row_bytes in (512, 4096):
        rows_total = table_bytes // row_bytes
        dst = torch.empty(n * row_bytes, dtype=torch.uint8, device="cuda")
        for pattern in ("seq", "rand"):
            if pattern == "seq":
                idx = (torch.arange(n, device="cuda", dtype=torch.int64) % rows_total).to(torch.int32)
            else:
                idx = torch.randint(0, rows_total, (n,), device="cuda", dtype=torch.int64).to(torch.int32)
            st = torch.cuda.current_stream().cuda_stream

            def run():
                check(lib.xg_extract(ctypes.c_void_p(dst.data_ptr()), ctypes.c_void_p(src_ptr),
                                     ctypes.c_void_p(idx.data_ptr()), n, row_bytes, 3, ctypes.c_void_p(st)))  # 3 = XG_U8
            ms = time_ms(run)
            out({"case": tag, "table_GB": round(table_bytes / 1e9, 2), "row_bytes": row_bytes, "pattern": pattern,
                 "ms": round(ms, 3), "GBps": round(n * row_bytes / ms / 1e6, 1)})
        del dst


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def out(d):
        print(json.dumps(d), flush=True)

    for table_bytes in (1 << 30, 8 << 30, 32 << 30):
        p = ctypes.c_void_p()
        check(lib.xg_device_malloc(ctypes.byref(p), table_bytes))
        view_u8(p.value, table_bytes)[:: 4096].fill_(1)
        torch.cuda.synchronize()
        hnd = (ctypes.c_ubyte * 64)()
        check(lib.xg_ipc_export(p, hnd))
        blob = torch.tensor(list(bytes(hnd)), dtype=torch.uint8, device="cuda")
        allb = [torch.empty_like(blob) for _ in range(world)]
        dist.all_gather(allb, blob)
        if rank == 0:
            peer = ctypes.c_void_p()
            hb = bytes(allb[1].cpu().numpy().tolist())
            check(lib.xg_ipc_import((ctypes.c_ubyte * 64).from_buffer_copy(hb), ctypes.byref(peer)))
            # (a) bulk copy
            n_copy = min(table_bytes, 4 << 30)
            local_buf = torch.empty(n_copy, dtype=torch.uint8, device="cuda")
            src = view_u8(peer.value, n_copy)
            ms = time_ms(lambda: local_buf.copy_(src))
            out({"case": "ipc bulk copy", "table_GB": round(table_bytes / 1e9, 2), "ms": round(ms, 3),
                 "GBps": round(n_copy / ms / 1e6, 1)})
            del local_buf, src
            # local reference
            gather_cases("local", p.value, table_bytes, out)
            gather_cases("ipc peer", peer.value, table_bytes, out)
            check(lib.xg_ipc_close(peer))
        dist.barrier()
        check(lib.xg_device_free(p))
        dist.barrier()

    if rank == 0:  # (c) same-process peer allocation on device 1
        check(lib.xg_enable_peer_access(0, 1))
        for table_bytes in (1 << 30, 32 << 30):
            with torch.cuda.device(1):
                q = ctypes.c_void_p()
                check(lib.xg_device_malloc(ctypes.byref(q), table_bytes))
            gather_cases("same-process peer", q.value, table_bytes, out)
            with torch.cuda.device(1):
                check(lib.xg_device_free(q))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
