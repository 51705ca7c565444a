#!/usr/bin/env python
"""Print the device timeline of the two phases of consecutive minibatches (XG_TIMELINE=1), 1 GPU.
  XG_TIMELINE=1 python tools/timeline.py [--scale N] [--in-flight K]"""
import argparse
import os
import sys
from pathlib import Path

import numpy as np
import torch

os.environ["XG_TIMELINE"] = "1"
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from xgnn_b200 import dataset as D  # noqa: E402
from xgnn_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=int, default=1)
ap.add_argument("--in-flight", type=int, default=2)
ap.add_argument("--steps", type=int, default=24)
ap.add_argument("--no-extract", action="store_true", help="sample phase alone (no feature store, no labels): what the captured graph costs without the gather beside it")
a = ap.parse_args()
torch.cuda.set_device(0)
st = D.attach(D.build_store("papers100M", 0, 0, 1, scale=a.scale))
stream = torch.cuda.Stream(priority=-1)
sm = ops.MinibatchSampler(st.graph, bench.FANOUT, 8000, store=None if a.no_extract else st.store, feat_dim=st.dim,
                          feat_dtype=st.dtype, label=None if a.no_extract else st.label, seed=2024, num_slots=a.in_flight + 1,
                          device=0, stream=stream)
batches = bench.seed_batches(st.train, 0, 1, 8000, a.steps)
d = [torch.from_numpy(b.view(np.int32)).cuda() for b in batches]
pend, rows = [], []
for i in range(a.steps):
    pend.append(sm.submit(d[i], batch_key=i))
    if len(pend) == a.in_flight:
        mb = sm.wait(pend.pop(0))
        rows.append(mb.timeline())
        mb.release()
for h in pend:
    mb = sm.wait(h)
    rows.append(mb.timeline())
    mb.release()
t0 = rows[8][0]
print("batch  sample[begin end]  extract[begin end]   (ms, relative)   sample_ms extract_ms step_ms")
for i in range(8, len(rows)):
    r = rows[i]
    step = rows[i][3] - rows[i - 1][3]
    print(f"{i:3d}   {r[0]-t0:8.3f} {r[1]-t0:8.3f}   {r[2]-t0:8.3f} {r[3]-t0:8.3f}     {r[1]-r[0]:6.3f} {r[3]-r[2]:6.3f} {step:6.3f}")
