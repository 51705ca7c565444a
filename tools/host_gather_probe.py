#!/usr/bin/env python
"""PCIe probe for host-resident rows (GPUExtractMissData path): in-kernel UVA gather of random rows out of a
pinned host table vs a plain pinned H2D copy of the same number of bytes.  One GPU."""
import ctypes
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from xgnn_b200._clib import check, lib  # noqa: E402

torch.cuda.set_device(0)
TABLE_GB = 8
host = torch.empty(TABLE_GB << 30, dtype=torch.uint8).pin_memory()
host[:: 4096] = 1
st = torch.cuda.current_stream().cuda_stream


def time_ms(fn, reps=4):
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


n_copy = 2 << 30
dst = torch.empty(n_copy, dtype=torch.uint8, device="cuda")
ms = time_ms(lambda: dst.copy_(host[:n_copy], non_blocking=True))
print(json.dumps({"case": "pinned H2D cudaMemcpy", "GB": n_copy / 1e9, "ms": round(ms, 2), "GBps": round(n_copy / ms / 1e6, 1)}))
for row_bytes in (512, 1536, 4096):
    rows_total = (TABLE_GB << 30) // row_bytes
    n = (1 << 30) // row_bytes
    out = torch.empty(n * row_bytes, dtype=torch.uint8, device="cuda")
    for pattern in ("seq", "rand"):
        idx = (torch.arange(n, device="cuda") if pattern == "seq" else torch.randint(0, rows_total, (n,), device="cuda")).to(torch.int32)
        ms = time_ms(lambda: check(lib.xg_extract(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(host.data_ptr()),
                                                 ctypes.c_void_p(idx.data_ptr()), n, row_bytes, 3, ctypes.c_void_p(st))))
        print(json.dumps({"case": "UVA gather from pinned host", "row_bytes": row_bytes, "pattern": pattern,
                          "ms": round(ms, 2), "GBps": round(n * row_bytes / ms / 1e6, 1)}))
