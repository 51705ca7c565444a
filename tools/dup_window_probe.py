#!/usr/bin/env python
"""How many of a layer's sampled neighbours are duplicates of an EARLIER neighbour (a) anywhere in the fill, (b) inside
the same window of W consecutive edges - what a shared-memory pre-dedup of a CTA tile could take off the global table.
Measured on real minibatches of the synthetic shapes (1 GPU):   python tools/dup_window_probe.py [--shape twitter]"""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from xgnn_b200 import dataset as D  # noqa: E402
from xgnn_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shape", default="papers100M")
ap.add_argument("--scale", type=int, default=1)
ap.add_argument("--batches", type=int, default=3)
a = ap.parse_args()
torch.cuda.set_device(0)
st = D.attach(D.build_store(a.shape, 0, 0, 1, scale=a.scale, with_features=False))
sm = ops.MinibatchSampler(st.graph, bench.FANOUT, 8000, seed=3, num_slots=2, device=0)
seeds = bench.seed_batches(st.train, 0, 1, 8000, a.batches)
out = {"shape": a.shape, "layers": {}}
for b in range(a.batches):
    mb = sm.sample(seeds[b], batch_key=b)
    for l in range(mb.num_layer):
        row = mb.row(l).to(torch.int64) & 0xFFFFFFFF  # local id of every edge's neighbour, in edge (= insert) order
        n = row.numel()
        rec = out["layers"].setdefault(f"L{l}", {"edges": 0, "dup_anywhere": 0, "dup_in_1024": 0, "dup_in_4096": 0, "dup_in_warp32": 0})
        rec["edges"] += n
        rec["dup_anywhere"] += n - int(torch.unique(row).numel())
        for w, key in ((1024, "dup_in_1024"), (4096, "dup_in_4096"), (32, "dup_in_warp32")):
            pad = (-n) % w
            r = torch.cat([row, torch.arange(pad, device=row.device) + (1 << 40)]).view(-1, w)
            s, _ = torch.sort(r, dim=1)
            rec[key] += int((s[:, 1:] == s[:, :-1]).sum())
    mb.release()
for rec in out["layers"].values():
    for k in list(rec):
        if k != "edges":
            rec[k + "_frac"] = round(rec[k] / max(rec["edges"], 1), 5)
print(json.dumps(out))
