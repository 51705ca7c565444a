#!/usr/bin/env python
"""Does cudaLimitMaxL2FetchGranularity change what a random 4-byte gather costs on B200?"""
import ctypes
import json
import sys

import torch

rt = ctypes.CDLL("libcudart.so.12")
gran = int(sys.argv[1]) if len(sys.argv) > 1 else 0
torch.cuda.set_device(0)
torch.zeros(1, device="cuda")
if gran:
    rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran))  # cudaLimitMaxL2FetchGranularity = 0x05
else:
    rc = 0
v = ctypes.c_size_t(0)
rt.cudaDeviceGetLimit(ctypes.byref(v), 5)
table = torch.arange(2 << 30, device="cuda", dtype=torch.int32)  # 8 GiB
idx = torch.randint(0, table.numel(), (64 << 20,), device="cuda")
out = torch.empty(idx.numel(), dtype=torch.int32, device="cuda")
best = 1e9
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    torch.index_select(table, 0, idx, out=out)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(json.dumps({"requested": gran, "set_rc": rc, "limit_now": v.value, "ms": round(best, 3),
                  "G_lookups_per_s": round(idx.numel() / best / 1e6, 2)}))
