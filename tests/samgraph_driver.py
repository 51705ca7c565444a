"""Runs the reference-facing API (xgnn_b200.samgraph.torch) the way the reference's training scripts do
(example/samgraph/sgnn/train_graphsage.py:180-260, 394-412) and checks every batch against the CPU oracle.

Own process per scenario: the engine is a process-wide singleton configured once (operation.cc:52,66).

  python tests/samgraph_driver.py single  <dataset_dir> <sample_type>
  python tests/samgraph_driver.py arch6   <dataset_dir> <sample_type> <num_worker>
"""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import oracle_lib as O  # noqa: E402
import synth  # noqa: E402

ENGINE_SEED = 0x58474E4E  # engine.cc create_sampler
BATCH, EPOCHS = 96, 2


def load(d):
    meta = dict(l.split() for l in open(os.path.join(d, "meta.txt")))
    V, dim = int(meta["NUM_NODE"]), int(meta["FEAT_DIM"])
    g = dict(V=V, dim=dim, indptr=np.fromfile(os.path.join(d, "indptr.bin"), np.uint32),
             indices=np.fromfile(os.path.join(d, "indices.bin"), np.uint32),
             feat=np.fromfile(os.path.join(d, "feat.bin"), np.float32).reshape(V, dim),
             label=np.fromfile(os.path.join(d, "label.bin"), np.uint64).astype(np.int64),
             train=np.fromfile(os.path.join(d, "train_set.bin"), np.uint32), prob=None, alias=None, prefix=None)
    k = int(os.environ.get("SAMGRAPH_EMPTY_FEAT", "0"))
    if k:  # mock features: node v reads row v & (2^k - 1) of the engine's pattern table (engine.cc load_dataset)
        rb = dim * 4
        r = np.arange(1 << k, dtype=np.int64)[:, None]
        table = ((31 * r + np.arange(rb, dtype=np.int64)[None, :]) & 0x3F).astype(np.uint8).view(np.float32)
        g["feat"] = table[np.arange(V) & ((1 << k) - 1)]
    if os.path.exists(os.path.join(d, "prob_table.bin")):
        g["prob"] = np.fromfile(os.path.join(d, "prob_table.bin"), np.float32)
        g["alias"] = np.fromfile(os.path.join(d, "alias_table.bin"), np.uint32)
        g["prefix"] = np.fromfile(os.path.join(d, "prob_prefix_table.bin"), np.float32)
    return g


def run_config(sam, d, sample_type, g, arch, extra):
    maxdeg = int(np.diff(g["indptr"].astype(np.int64)).max())
    cfg = {"dataset_path": d, "_arch": arch, "_sample_type": sample_type, "batch_size": BATCH, "num_epoch": EPOCHS,
           "_cache_policy": sam.kCacheByDegree, "cache_percentage": 0.5, "max_sampling_jobs": 10, "max_copying_jobs": 2,
           "omp_thread_num": 4, "num_layer": 2, "num_hidden": 16, "lr": 0.003, "dropout": 0.5}
    if sample_type == sam.kRandomWalk:
        cfg.update(random_walk_length=3, random_walk_restart_prob=0.5, num_random_walk=4, num_neighbor=5)
        fan = [5, 5]
    else:
        fan = [maxdeg, maxdeg] if sample_type in (sam.kKHop0, sam.kKHop3, sam.kKHop2) else [4, 6]
        cfg.update(fanout=fan, num_fanout=2)
    cfg.update(extra)
    if os.environ.get("XG_TEST_PRESAMPLE"):  # cache ranking from presampling instead of cache_by_degree.bin
        cfg.update(_cache_policy=sam.kCacheByPreSample, presample_epoch=int(os.environ["XG_TEST_PRESAMPLE"]))
    sam.config(cfg)
    return fan


def expected_batches(g, worker, num_worker):
    """DistAlignedShuffler: padded train set, one more Fisher-Yates pass per epoch, contiguous slice per worker"""
    train = g["train"]
    pad = (-len(train)) % num_worker
    data = np.concatenate([train, train[:pad]]) if pad else train.copy()
    n_local = len(data) // num_worker
    out = []
    for ep in range(EPOCHS):
        data = synth.minstd_shuffle_epoch(data, ep)
        mine = data[worker * n_local:(worker + 1) * n_local]
        out.append([mine[o:o + BATCH] for o in range(0, n_local, BATCH)])
    return out


def check_batch(sam, g, key, seeds, fan, sample_type):
    import torch
    u32 = lambda t: t.cpu().numpy().view(np.uint32)  # noqa: E731
    if sample_type in (sam.kKHop0, sam.kKHop2, sam.kKHop3):
        ref = O.minibatch(g["indptr"], g["indices"], seeds, fan, seed=1)
    else:
        ref = O.minibatch_ext(g, seeds, fan, sample_type, ENGINE_SEED, key, rw=(3, 0.5, 4))
    weighted = sample_type == sam.kRandomWalk
    blocks, feat, label = (sam.get_dgl_blocks_with_weights if weighted else sam.get_dgl_blocks)(key, len(fan))
    assert np.array_equal(u32(sam.get_graph_input_nodes(key)), ref["input_nodes"]), "input_nodes"
    assert np.array_equal(u32(sam.get_graph_output_nodes(key)), seeds), "output_nodes"
    for l, blk in enumerate(blocks):
        row, col, nsrc, ndst = blk[:4]
        assert np.array_equal(u32(row), ref["row"][l]) and np.array_equal(u32(col), ref["col"][l]), f"layer {l} edges"
        assert nsrc == ref["num_src"][l] and ndst == ref["num_dst"][l], f"layer {l} sizes"
        assert sam.get_graph_num_edge(key, l) == len(ref["row"][l])
        if weighted:
            assert np.array_equal(u32(blk[4]), ref["data"][l]), f"layer {l} data"
        if sample_type in (sam.kKHop0, sam.kKHop2, sam.kKHop3):
            ip, c = sam.get_graph_csc_indptr(key, l), ref["col"][l]
            if len(c) and not np.all(c[1:] >= c[:-1]):
                assert ip is None, "padded (duplicate) seeds have no CSC form"
            else:
                assert np.array_equal(u32(ip), O.coo_to_csc_indptr(c, ref["num_dst"][l])), f"layer {l} csc"
    if sample_type in (sam.kKHop0, sam.kKHop2, sam.kKHop3):
        cblocks, _, _ = sam.get_dgl_blocks_csc(key, len(fan), with_feat=False)
        assert len(cblocks) == len(fan)
    assert feat.dtype == torch.float32 and feat.cpu().numpy().tobytes() == g["feat"][ref["input_nodes"]].tobytes(), "feat"
    assert np.array_equal(label.cpu().numpy(), g["label"][seeds]), "label"
    return feat  # the caller may hold it across batches (ownership test)


def consume(sam, g, fan, sample_type, worker, num_worker, background):
    batches = expected_batches(g, worker, num_worker)
    num_step, local = sam.steps_per_epoch(), sam.num_local_step()
    assert local == len(batches[0]) and num_step == local * num_worker, (local, num_step)
    assert sam.num_epoch() == EPOCHS and sam.feat_dim() == g["dim"] and sam.num_class() == 47
    if background:
        (sam.extract_start if num_worker > 1 else (lambda n: sam.start()))(EPOCHS * local)
    held = None
    for ep in range(EPOCHS):
        for st in range(local):
            if not background:
                sam.sample_once()
            key = sam.get_next_batch()
            assert key == ep * num_step + worker * local + st, (key, ep, st)
            feat = check_batch(sam, g, key, batches[ep][st], fan, sample_type)
            if held is None:
                held, held_copy = feat, feat.clone()
            assert sam.get_log_step_value_by_key(key, sam.kLogL1NumNode) == feat.shape[0]
            assert sam.get_log_step_value_by_key(key, sam.kLogL1SampleTime) > 0
        assert sam.get_log_epoch_value(ep, sam.kLogEpochNumSample) > 0
        assert sam.get_log_epoch_value(ep, sam.kLogEpochSampleTime) > 0
    # a tensor handed out for the very first batch is still intact: its slot was not recycled under it
    import torch
    assert torch.equal(held, held_copy), "a held batch tensor was overwritten"
    sam.report_epoch_average(EPOCHS - 1)
    sam.shutdown()


def run_arch5(sam, g, d, sample_type, num_sampler, num_trainer):
    """arch5 (example/samgraph/multi_gpu/train_graphsage.py): S sampler processes and T trainer processes; a Task is
    sampled in one process / GPU and its features are gathered in another.  Trainers take Tasks off one shared queue in
    whatever order they arrive, so every batch is checked by its KEY: key -> (epoch, sampler, local step) -> the seeds
    DistAlignedShuffler gives that sampler for that step."""
    import torch
    ngpu = int(os.environ.get("XG_TEST_NUM_GPU", "1"))
    fan = run_config(sam, d, sample_type, g, sam.kArch5, {"num_sample_worker": num_sampler, "num_train_worker": num_trainer})
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
        local = sam.num_local_step()
        assert sam.steps_per_epoch() == local * num_sampler
        for _ in range(EPOCHS * local):
            sam.sample_once()
        assert sam.get_log_epoch_value(0, sam.kLogEpochSampleTime) > 0
        sam.shutdown()  # returns once every Task this sampler sent has been pulled

    def trainer(t):
        sam.train_init(t, f"cuda:{(num_sampler + t) % ngpu}")  # a different GPU than the samplers when the box has one
        num_step, local = sam.steps_per_epoch(), sam.num_local_step()
        total = EPOCHS * num_step
        mine = total // num_trainer + (1 if t < total % num_trainer else 0)
        expected = [expected_batches(g, w, num_sampler) for w in range(num_sampler)]
        background = t % 2 == 1
        if background:
            sam.extract_start(mine)
        seen, held = [], None
        for _ in range(mine):
            if not background:
                sam.sample_once()
            key = sam.get_next_batch()
            ep, step = divmod(key, num_step)
            w, st = divmod(step, local)
            feat = check_batch(sam, g, key, expected[w][ep][st], fan, sample_type)
            if held is None:
                held, held_copy = feat, feat.clone()
            assert sam.get_log_step_value_by_key(key, sam.kLogL1NumNode) == feat.shape[0]
            seen.append(key)
        assert len(set(seen)) == len(seen), "a Task was delivered twice"
        assert torch.equal(held, held_copy), "a held batch tensor was overwritten"
        with open(os.path.join(d, f"arch5_keys_{os.getppid()}_{t}.txt"), "w") as f:
            f.write(" ".join(map(str, seen)))
        sam.shutdown()

    for w in range(num_sampler):
        fork(lambda w=w: sampler(w))
    for t in range(num_trainer):
        fork(lambda t=t: trainer(t))
    bad = sum(sam.wait_one_child() for _ in pids)
    assert bad == 0, f"{bad} process(es) failed"
    # every Task of every epoch reached exactly one trainer
    keys = []
    for t in range(num_trainer):
        fn = os.path.join(d, f"arch5_keys_{os.getpid()}_{t}.txt")
        keys += [int(x) for x in open(fn).read().split()]
        os.remove(fn)
    num_step = -(-(-(-len(g["train"]) // num_sampler)) // BATCH) * num_sampler
    assert sorted(keys) == list(range(EPOCHS * num_step)), "Tasks lost or duplicated across trainers"
    print("DRIVER OK")


def main():
    mode, d, sample_type = sys.argv[1], sys.argv[2], int(sys.argv[3])
    import xgnn_b200.samgraph.torch as sam
    g = load(d)
    if mode == "single":
        background = len(sys.argv) > 4 and sys.argv[4] == "bg"
        fan = run_config(sam, d, sample_type, g, sam.kArch3, {"sampler_ctx": "cuda:0", "trainer_ctx": "cuda:0"})
        sam.init()
        k = int(os.environ.get("SAMGRAPH_EMPTY_FEAT", "0"))  # mock features: the table has 2^k rows (engine.cc:198-208)
        assert tuple(sam.get_dataset_feat().shape) == ((1 << k) if k else g["V"], g["dim"])
        assert np.array_equal(sam.get_dataset_label().numpy(), g["label"])
        consume(sam, g, fan, sample_type, 0, 1, background)
        print("DRIVER OK")
        return
    if mode == "arch5":
        run_arch5(sam, g, d, sample_type, int(sys.argv[4]), int(sys.argv[5]))
        return
    num_worker = int(sys.argv[4])
    import torch
    ngpu = int(os.environ.get("XG_TEST_NUM_GPU", "1"))
    uniform = sample_type in (sam.kKHop0, sam.kKHop2, sam.kKHop3, sam.kKHop1, sam.kRandomWalk)
    extra = {"num_worker": num_worker, "part_cache": "True", "gpu_extract": "True"}
    if uniform:
        extra["use_dist_graph"] = 0.7  # weighted tables need the flat CSR
    if os.environ.get("XG_TEST_ICS22"):  # "clique_size,alpha": the ICS'22 clique placement instead of the GGMS partition cache
        clique, alpha = os.environ["XG_TEST_ICS22"].split(",")
        del extra["part_cache"]
        extra.update(use_ics22_song_solver="True", clique_size=int(clique), ics22_song_alpha=float(alpha), use_dist_graph=0.7)
    fan = run_config(sam, d, sample_type, g, sam.kArch6, extra)
    sam.data_init()  # before fork, no CUDA in the parent (train_graphsage.py:386-402)
    pids = []
    for w in range(num_worker):
        pid = os.fork()
        if pid == 0:
            try:
                ctx = f"cuda:{w % ngpu}"
                sam.sample_init(w, ctx)
                sam.train_init(w, ctx)
                consume(sam, g, fan, sample_type, w, num_worker, background=(w % 2 == 1))
                os._exit(0)
            except BaseException as e:  # noqa: BLE001
                import traceback
                traceback.print_exc()
                os._exit(1)
        pids.append(pid)
    bad = sum(sam.wait_one_child() for _ in pids)
    assert bad == 0, f"{bad} worker(s) failed"
    print("DRIVER OK")


if __name__ == "__main__":
    main()
