#!/usr/bin/env python
"""Per-stage device time of the sample phase WITH the previous batch's gather running beside it (XG_STAGE_OVERLAP=1:
stage marks on both streams, kernels launched one by one instead of as a graph), next to the back-to-back stage times
of bench.py's stage pass.  1 GPU.
  python tools/stage_overlap.py [--scale N] [--steps K]"""
import argparse
import ctypes
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from xgnn_b200 import dataset as D  # noqa: E402
from xgnn_b200 import ops  # noqa: E402
from xgnn_b200._clib import check, lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=int, default=1)
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--in-flight", type=int, default=2)
a = ap.parse_args()
torch.cuda.set_device(0)
st = D.attach(D.build_store("papers100M", 0, 0, 1, scale=a.scale))
batches = bench.seed_batches(st.train, 0, 1, 8000, a.steps)
d = [torch.from_numpy(b.view(np.int32)).cuda() for b in batches]
L = len(bench.FANOUT)


def stages_of(hnd):
    out, ms = {}, ctypes.c_float(0)
    for name, code in bench.STAGES.items():
        for l in (range(L) if code in (1, 2, 3) else [0]):
            check(lib.xg_batch_stage_ms(hnd, code, l, ctypes.byref(ms)))
            out[f"{name}.L{l}" if code in (1, 2, 3) else name] = ms.value
    return out


def run(overlap):
    os.environ["XG_STAGE_OVERLAP"] = "1" if overlap else "0"
    stream = torch.cuda.Stream(priority=-1)
    sm = ops.MinibatchSampler(st.graph, bench.FANOUT, 8000, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                              label=st.label, seed=2024, num_slots=a.in_flight + 1, device=0, stream=stream,
                              use_graph=False, stage_timing=True)
    acc, n, pend = {}, 0, []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def finish(i, hnd):
        nonlocal n
        mb = sm.wait(hnd)
        if i >= 8:
            for k, v in stages_of(hnd).items():
                acc[k] = acc.get(k, 0.0) + v
            n += 1
        mb.release()

    for i in range(a.steps):
        if i == 8:
            torch.cuda.synchronize()
            e0.record(stream)
        pend.append((i, sm.submit(d[i], batch_key=i)))
        if len(pend) == (a.in_flight if overlap else 1):
            finish(*pend.pop(0))
    for p in pend:
        finish(*p)
    e1.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    del sm
    return {k: v / n for k, v in acc.items()}


alone = run(False)
beside = run(True)
print(f"{'stage':16s} {'back to back':>12s} {'beside gather':>14s}  ratio")
tot_a = tot_b = 0.0
for k in alone:
    print(f"{k:16s} {alone[k]:12.4f} {beside[k]:14.4f}  {beside[k] / max(alone[k], 1e-9):5.2f}")
    if k not in ("feature", "label"):
        tot_a += alone[k]
        tot_b += beside[k]
print(f"{'sample phase':16s} {tot_a:12.4f} {tot_b:14.4f}  {tot_b / tot_a:5.2f}")
