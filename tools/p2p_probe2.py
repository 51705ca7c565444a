#!/usr/bin/env python
"""NVLink peer-gather probe, part 2: what happens when BOTH GPUs gather at once and when local and
peer rows are mixed in one launch (the GGMS feature gather at P = 2).

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/p2p_probe2.py
"""
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

ROW = 512
N = 2_000_000


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    part_bytes = int(os.environ.get("PROBE_PART_GB", "16")) << 30
    rows_part = part_bytes // ROW
    p = ctypes.c_void_p()
    check(lib.xg_device_malloc(ctypes.byref(p), part_bytes))
    fill = os.environ.get("PROBE_FILL", "0")
    if fill != "0":  # real data instead of never-written pages
        v = torch.as_tensor(ops._DevMem(p.value, (part_bytes // 8,), "<i8", None), device="cuda")
        step = 1 << 27
        for o in range(0, v.numel(), step):
            if fill == "1":
                v[o:o + step].random_()
            else:
                v[o:o + step].fill_(0x0102030405060708)
        torch.cuda.synchronize()
    hnd = (ctypes.c_ubyte * 64)()
    check(lib.xg_ipc_export(p, hnd))
    blob = torch.tensor(list(bytes(hnd)), dtype=torch.uint8, device="cuda")
    allb = [torch.empty_like(blob) for _ in range(world)]
    dist.all_gather(allb, blob)
    peer = ctypes.c_void_p()
    hb = bytes(allb[1 - rank].cpu().numpy().tolist())
    check(lib.xg_ipc_import((ctypes.c_ubyte * 64).from_buffer_copy(hb), ctypes.byref(peer)))
    parts = [None, None]
    parts[rank], parts[1 - rank] = p.value, peer.value

    V = 2 * rows_part
    table = torch.arange(V, dtype=torch.int64, device="cuda").to(torch.int32)  # rank == node id
    store = ops.make_feature_store(ROW, None, table, parts=parts, local_parts=(rank,))
    dst = torch.empty(N * ROW, dtype=torch.uint8, device="cuda")
    g = torch.Generator(device="cuda")
    g.manual_seed(7 + rank)
    r = torch.randint(0, rows_part, (N,), device="cuda", dtype=torch.int64, generator=g)
    idx_remote = (2 * r + (1 - rank)).to(torch.int32)   # every row in the peer part
    idx_local = (2 * r + rank).to(torch.int32)
    idx_mixed = torch.randint(0, V, (N,), device="cuda", dtype=torch.int64, generator=g).to(torch.int32)
    # mixed but warp-segregated: 32-row groups are all-local or all-remote
    grp = (torch.arange(N, device="cuda") // 32) % 2
    idx_seg = torch.where(grp == 0, idx_local, idx_remote)
    # 20 % of the rows on the peer (the two-tier store's regime): mixed at random, grouped per warp, fully sorted
    pick = torch.rand(N, device="cuda", generator=g) < 0.2
    idx_mix20 = torch.where(pick, idx_remote, idx_local)
    grp5 = (torch.arange(N, device="cuda") // 32) % 5 == 0
    idx_seg20 = torch.where(grp5, idx_remote, idx_local)
    idx_sort20 = torch.cat([idx_local[: int(N * 0.8)], idx_remote[: N - int(N * 0.8)]])
    st = torch.cuda.current_stream().cuda_stream

    def gather(idx):
        check(lib.xg_extract_ggms(ctypes.byref(store), ctypes.c_void_p(idx.data_ptr()), N,
                                  ctypes.c_void_p(dst.data_ptr()), None, ctypes.c_void_p(st)))

    def timed(idx, active, reps=5):
        """both ranks enter; only `active` ranks launch; returns best ms of this rank (or None)"""
        best = None
        for _ in range(reps + 1):
            dist.barrier()
            torch.cuda.synchronize()
            if rank in active:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(4):
                    gather(idx)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 4
                best = ms if best is None else min(best, ms)
        return best

    cases = [("remote only, rank0 alone", idx_remote, (0,)),
             ("remote only, both ranks", idx_remote, (0, 1)),
             ("local only, both ranks", idx_local, (0, 1)),
             ("mixed 50/50, rank0 alone", idx_mixed, (0,)),
             ("mixed 50/50, both ranks", idx_mixed, (0, 1)),
             ("warp-segregated 50/50, rank0 alone", idx_seg, (0,)),
             ("warp-segregated 50/50, both ranks", idx_seg, (0, 1)),
             ("20 % peer, mixed, both ranks", idx_mix20, (0, 1)),
             ("20 % peer, warp-segregated, both ranks", idx_seg20, (0, 1)),
             ("20 % peer, local rows first then peer rows, both ranks", idx_sort20, (0, 1))]
    for name, idx, active in cases:
        ms = timed(idx, active)
        t = torch.tensor([ms if ms is not None else 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"case": name, "part_GB": part_bytes >> 30, "fill": os.environ.get("PROBE_FILL", "0"), "ms_max": round(float(t.item()), 3),
                              "rows_GBps": round(N * ROW / float(t.item()) / 1e6, 1)}), flush=True)
    # rank 0 local-only while rank 1 reads rank 0's memory
    dist.barrier()
    ms = timed(idx_local if rank == 0 else idx_remote, (0, 1))
    allms = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
    dist.all_gather(allms, torch.tensor([ms], dtype=torch.float64, device="cuda"))
    if rank == 0:
        print(json.dumps({"case": "rank0 local-only + rank1 remote-only (reads rank0)",
                          "ms_rank0": round(float(allms[0].item()), 3), "ms_rank1": round(float(allms[1].item()), 3)}),
              flush=True)
    dist.barrier()
    check(lib.xg_ipc_close(peer))
    dist.barrier()
    check(lib.xg_device_free(p))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
