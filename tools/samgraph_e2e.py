#!/usr/bin/env python
"""End to end through the REFERENCE-FACING API (xgnn_b200.samgraph.torch = samgraph_* C-ABI), one GPU.

  python tools/samgraph_e2e.py [--shape papers100M --scale 16 --epochs 3 --dir /dev/shm/xg_ds]

1. dumps a synthetic dataset of the named shape in the reference's on-disk format (meta.txt + *.bin);
2. runs the loop of example/samgraph/sgnn/train_graphsage.py:180-260 without the model: config -> init ->
   for epoch / step: sample_once(); key = get_next_batch(); blocks, feat, label = get_dgl_blocks(key, L);
3. reports sampled edges/s the way the reference's scripts do (kLogEpochNumSample / epoch wall time) next to what
   the operator-level engine (xg_sampler, what bench.py drives) does on the same dataset and batches.
Prints one JSON line.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def dump_dataset(path, shape, scale):
    from xgnn_b200 import dataset as D
    V, E, dim, dtype, ncls, frac = D.SHAPES[shape]
    V, E = V // scale, E // scale
    dev = torch.device("cuda", 0)
    os.makedirs(path, exist_ok=True)
    indptr, indices, ne = D.gen_csr(0, 1, V, V, D.degree_constant(V, E), D.SYNTH_SEED, dev)
    indptr.cpu().numpy().view(np.uint32).tofile(os.path.join(path, "indptr.bin"))
    indices[:ne].cpu().numpy().view(np.uint32).tofile(os.path.join(path, "indices.bin"))
    del indices
    with open(os.path.join(path, "feat.bin"), "wb") as f:
        step = 1 << 20
        for o in range(0, V, step):
            n = min(step, V - o)
            f.write(D.gen_features(n, dim, dtype, dev, begin=o).cpu().numpy().tobytes())
    (np.arange(V, dtype=np.uint64) % ncls).tofile(os.path.join(path, "label.bin"))
    train = D.train_set(V, max(1, int(V * frac)))
    train.tofile(os.path.join(path, "train_set.bin"))
    np.zeros(0, np.uint32).tofile(os.path.join(path, "valid_set.bin"))
    np.zeros(0, np.uint32).tofile(os.path.join(path, "test_set.bin"))
    np.arange(V, dtype=np.uint32).tofile(os.path.join(path, "cache_by_degree.bin"))  # degree is non-increasing in the id
    with open(os.path.join(path, "meta.txt"), "w") as f:
        f.write(f"NUM_NODE {V}\nNUM_EDGE {ne}\nFEAT_DIM {dim}\nNUM_CLASS {ncls}\nNUM_TRAIN_SET {len(train)}\n"
                f"NUM_VALID_SET 0\nNUM_TEST_SET 0\n")
        if dtype == torch.float16:
            f.write("FEAT_DATA_TYPE F16\n")
    torch.cuda.empty_cache()
    return V, ne, len(train)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="papers100M")
    ap.add_argument("--scale", type=int, default=16)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--batch", type=int, default=8000)
    ap.add_argument("--dir", default="/dev/shm/xg_ds")
    ap.add_argument("--background", action="store_true", help="sam.start(): a producer thread samples ahead of the consumer")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    t0 = time.perf_counter()
    V, E, T = dump_dataset(a.dir, a.shape, a.scale)
    t_dump = time.perf_counter() - t0

    import xgnn_b200.samgraph.torch as sam
    fanout = [5, 10, 15]
    sam.config({"dataset_path": a.dir, "_arch": sam.kArch3, "_sample_type": sam.kKHop3, "batch_size": a.batch,
                "num_epoch": a.epochs, "_cache_policy": sam.kCacheByDegree, "cache_percentage": 1.0,
                "max_sampling_jobs": 10, "max_copying_jobs": 2, "omp_thread_num": 8, "num_layer": 3, "num_hidden": 256,
                "lr": 0.003, "dropout": 0.5, "fanout": fanout, "num_fanout": 3, "sampler_ctx": "cuda:0",
                "trainer_ctx": "cuda:0"})
    t0 = time.perf_counter()
    sam.init()
    t_init = time.perf_counter() - t0
    steps = sam.steps_per_epoch()
    per_epoch = []
    if a.background:
        sam.start()
    for ep in range(a.epochs):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        edges = rows = 0
        for st in range(steps):
            if not a.background:
                sam.sample_once()
            key = sam.get_next_batch()
            blocks, feat, label = sam.get_dgl_blocks(key, 3)
            edges += sum(sam.get_graph_num_edge(key, l) for l in range(3))
            rows += feat.shape[0]
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        per_epoch.append({"s": round(dt, 4), "edges": edges, "rows": rows,
                          "log_num_sample": sam.get_log_epoch_value(ep, sam.kLogEpochNumSample)})
    sam.shutdown()
    use = per_epoch[1:] or per_epoch  # first epoch discarded as warm-up, like the reference (common_config.py:174-175)
    eps = sum(e["edges"] for e in use) / sum(e["s"] for e in use)
    print(json.dumps({"api": "samgraph.torch (samgraph_* C-ABI)", "background_producer": a.background, "shape": a.shape, "scale": a.scale, "num_node": V,
                      "num_edge": E, "train_set": T, "batch": a.batch, "steps_per_epoch": steps,
                      "sampled_edges_per_s": eps, "ms_per_step": 1e3 * sum(e["s"] for e in use) / (len(use) * steps),
                      "epochs": per_epoch, "dump_s": round(t_dump, 1), "init_s": round(t_init, 1),
                      "note": "one batch at a time (sample_once -> get_next_batch -> get_dgl_blocks): no look-ahead, "
                              "so the two phases of consecutive batches do not overlap as they do under bench.py"}))


if __name__ == "__main__":
    main()
