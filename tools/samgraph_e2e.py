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
    ap.add_argument("--arch5", action="store_true", help="decoupled: sampler process(es) on the first GPUs, trainer process(es) on the next")
    ap.add_argument("--samplers", type=int, default=1)
    ap.add_argument("--trainers", type=int, default=1)
    ap.add_argument("--dump-only", action="store_true")
    a = ap.parse_args()
    if a.arch5:
        return main_arch5(a)
    torch.cuda.set_device(0)
    t0 = time.perf_counter()
    V, E, T = dump_dataset(a.dir, a.shape, a.scale)
    t_dump = time.perf_counter() - t0
    if a.dump_only:
        print(json.dumps({"num_node": V, "num_edge": E, "train_set": T}))
        return

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


def main_arch5(a):
    """S sampler processes (cuda:0..S-1) -> task queue -> T trainer processes (cuda:S..), as multi_gpu/train_graphsage.py
    runs arch5; no model.  The parent must not touch CUDA before the fork, so the dataset is dumped by a child."""
    import subprocess
    r = subprocess.run([sys.executable, __file__, "--dump-only", "--shape", a.shape, "--scale", str(a.scale), "--dir", a.dir],
                       capture_output=True, text=True, check=True)
    meta = json.loads(r.stdout.strip().splitlines()[-1])
    import xgnn_b200.samgraph.torch as sam
    ngpu = int(os.environ.get("XG_NUM_GPU", "0")) or int(subprocess.run(
        ["nvidia-smi", "-L"], capture_output=True, text=True).stdout.count("GPU "))
    S, T = a.samplers, a.trainers
    sam.config({"dataset_path": a.dir, "_arch": sam.kArch5, "_sample_type": sam.kKHop3, "batch_size": a.batch,
                "num_epoch": a.epochs, "_cache_policy": sam.kCacheByDegree, "cache_percentage": 1.0,
                "max_sampling_jobs": 10, "max_copying_jobs": 2, "omp_thread_num": 8, "num_layer": 3, "num_hidden": 256,
                "lr": 0.003, "dropout": 0.5, "fanout": [5, 10, 15], "num_fanout": 3, "num_sample_worker": S,
                "num_train_worker": T})
    sam.data_init()
    pids = []

    def fork(fn):
        pid = os.fork()
        if pid == 0:
            try:
                fn()
                os._exit(0)
            except BaseException:  # noqa: BLE001
                import traceback
                traceback.print_exc()
                os._exit(1)
        pids.append(pid)

    def sampler(w):
        sam.sample_init(w, f"cuda:{w % ngpu}")
        for _ in range(a.epochs * sam.num_local_step()):
            sam.sample_once()
        sam.shutdown()

    def trainer(t):
        dev = (S + t) % ngpu
        sam.train_init(t, f"cuda:{dev}")
        torch.cuda.set_device(dev)
        steps = sam.steps_per_epoch()
        total = a.epochs * steps
        mine = total // T + (1 if t < total % T else 0)
        warm = steps // T  # about one epoch of this trainer's share
        sam.extract_start(mine)
        edges = rows = nbytes = 0
        recv = copy = 0.0
        t0 = None
        for i in range(mine):
            if i == warm:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
            key = sam.get_next_batch()
            blocks, feat, label = sam.get_dgl_blocks(key, 3)
            if i >= warm:
                e = sum(sam.get_graph_num_edge(key, l) for l in range(3))
                edges += e
                rows += feat.shape[0]
                nbytes += 4 * (feat.shape[0] + label.shape[0] + 2 * e + sum(b[3] + 1 for b in blocks))
                recv += sam.get_log_step_value_by_key(key, sam.kLogL1RecvTime)
                copy += sam.get_log_step_value_by_key(key, sam.kLogL1CopyTime)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        n = mine - warm
        with open(os.path.join(a.dir, f"arch5_{t}.json"), "w") as f:
            json.dump({"trainer": t, "device": dev, "batches": n, "s": dt, "edges": edges, "rows": rows,
                       "task_MB": nbytes / n / 1e6, "recv_ms": 1e3 * recv / n, "extract_ms": 1e3 * copy / n}, f)
        sam.shutdown()

    for w in range(S):
        fork(lambda w=w: sampler(w))
    for t in range(T):
        fork(lambda t=t: trainer(t))
    bad = sum(sam.wait_one_child() for _ in pids)
    assert bad == 0
    per = [json.load(open(os.path.join(a.dir, f"arch5_{t}.json"))) for t in range(T)]
    print(json.dumps({"api": "samgraph.torch arch5 (decoupled sampler / trainer processes)", "shape": a.shape, "scale": a.scale,
                      **meta, "samplers": S, "trainers": T, "gpus": ngpu, "batch": a.batch,
                      "sampled_edges_per_s": sum(p["edges"] for p in per) / max(p["s"] for p in per),
                      "trainers_detail": per,
                      "note": "recv_ms = dequeue + ONE peer copy of the packed Task out of the sampler's HBM (incl. waiting "
                              "for the sampler); extract_ms = feature + label gather on the trainer GPU"}))


if __name__ == "__main__":
    main()
