"""GPU probe: cost per key of one fill (insert + number) at a FIXED load factor as a function of the hash-table size:
how much of the 126 MB L2 a randomly probed table can really use."""
import json, sys, os
sys.path.insert(0, ".")
import torch
from xgnn_b200 import ops

torch.cuda.set_device(0)
V = 111_059_956
g = torch.Generator(device="cuda"); g.manual_seed(1)
for table_mb in [4, 8, 16, 24, 32, 48, 64, 84, 128, 256]:
    slots = table_mb * 1_000_000 // 8
    max_items = int(slots / 1.25)
    n = int(slots * 0.4)
    keys = torch.randint(0, V, (n,), device="cuda", dtype=torch.int32, generator=g)
    t = ops.OrderedHashTable(max_items)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    times = []
    for it in range(6):
        t.Reset()
        torch.cuda.synchronize()
        ev[0].record()
        t.FillWithDupRevised(keys)
        ev[1].record()
        torch.cuda.synchronize()
        times.append(ev[0].elapsed_time(ev[1]))
    print(json.dumps({"table_MB": table_mb, "keys": n, "fill_ms": round(min(times[1:]), 4),
                      "ns_per_key": round(min(times[1:]) * 1e6 / n, 4)}), flush=True)
    del t
