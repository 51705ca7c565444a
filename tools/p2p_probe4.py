#!/usr/bin/env python
"""Probe 4: bisect why the bench's IPC-shared shards gather slowly while a plain allocation does not.
Each trial: rank 1 (and rank 0) allocate a list of blocks, export, import; rank 0 times a remote-only
random 512-B row gather (xg_extract) out of every imported block."""
import ctypes
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from xgnn_b200 import ops  # noqa: E402
from xgnn_b200._clib import check, lib  # noqa: E402

N = 1_000_000
ROW = 512


def view_u8(ptr, nbytes):
    return torch.as_tensor(ops._DevMem(ptr, (nbytes,), "|u1", None), device="cuda")


def gather_ms(src_ptr, nbytes, dst):
    rows = nbytes // ROW
    idx = torch.randint(0, rows, (N,), device="cuda", dtype=torch.int64).to(torch.int32)
    st = torch.cuda.current_stream().cuda_stream
    best = 1e9
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.xg_extract(ctypes.c_void_p(dst.data_ptr()), ctypes.c_void_p(src_ptr), ctypes.c_void_p(idx.data_ptr()),
                             N, ROW, 3, ctypes.c_void_p(st)))
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def trial(name, sizes, rank, world, dst, via_torch=False, fill=False):
    mine, raw = [], []
    for nbytes in sizes:
        t = None
        if via_torch:  # what bench.exchange_parts does: a torch tensor first, then a cudaMalloc copy of it
            t = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
            t[:: 4096] = 1
        p = ctypes.c_void_p()
        check(lib.xg_device_malloc(ctypes.byref(p), nbytes))
        if t is not None:
            view_u8(p.value, nbytes).copy_(t)
            torch.cuda.synchronize()
            del t
        elif fill:
            view_u8(p.value, nbytes)[:: 512] = 1
            torch.cuda.synchronize()
        hnd = (ctypes.c_ubyte * 64)()
        check(lib.xg_ipc_export(p, hnd))
        mine.append(bytes(hnd))
        raw.append(p.value)
    if via_torch:
        torch.cuda.empty_cache()
    blob = torch.tensor(list(b"".join(mine)), dtype=torch.uint8, device="cuda")
    allb = [torch.empty_like(blob) for _ in range(world)]
    dist.all_gather(allb, blob)
    hb = bytes(allb[1 - rank].cpu().numpy().tolist())
    peers = []
    for k in range(len(sizes)):
        out = ctypes.c_void_p()
        check(lib.xg_ipc_import((ctypes.c_ubyte * 64).from_buffer_copy(hb[64 * k:64 * k + 64]), ctypes.byref(out)))
        peers.append(out.value)
    dist.barrier()
    if rank == 0:
        for k, nbytes in enumerate(sizes):
            ms_l = gather_ms(raw[k], nbytes, dst)
            ms_r = gather_ms(peers[k], nbytes, dst)
            print(json.dumps({"trial": name, "block": k, "bytes": nbytes, "local_va": hex(raw[k]), "peer_va": hex(peers[k]),
                              "local_ms": round(ms_l, 3), "peer_ms": round(ms_r, 3),
                              "peer_GBps": round(N * ROW / ms_r / 1e6, 1)}), flush=True)
    dist.barrier()
    for q in peers:
        check(lib.xg_ipc_close(ctypes.c_void_p(q)))
    dist.barrier()
    for q in raw:
        check(lib.xg_device_free(ctypes.c_void_p(q)))
    dist.barrier()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dst = torch.empty(N * ROW, dtype=torch.uint8, device="cuda")
    A, B, C = 27_765_000, 403_586_076, 3_553_918_464  # the bench's shard sizes at --scale 8 (indptr, indices, feat)
    trial("one block 4 GiB", [4 << 30], rank, world, dst)
    trial("one block, bench feat size", [C], rank, world, dst)
    trial("three blocks A,B,C", [A, B, C], rank, world, dst)
    trial("three blocks C,B,A", [C, B, A], rank, world, dst)
    trial("three blocks A,B,C via torch copy", [A, B, C], rank, world, dst, via_torch=True)
    trial("three blocks, 2 MiB-rounded sizes", [(x + (2 << 20) - 1) // (2 << 20) * (2 << 20) for x in (A, B, C)], rank, world,
          dst)
    trial("full-scale sizes A,B,C", [222_119_916, 3_228_688_610 // 2 * 2, 28_431_348_736], rank, world, dst)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
