"""World-size-2 (and 3) CPU tests of the host-side N > 1 logic over the gloo backend: the work split
(DistAlignedShuffler slices, dist/dist_shuffler_aligned.cc:37-121) and the data split (GGMS placement:
topology rows v % P, cache ranks r % P, the replicated hot tier) must tile the whole job exactly once
across ranks.  No GPU, no data-path collective: gloo only carries the cross-rank CHECK."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gather_arrays(a: np.ndarray, world: int):
    """all_gather of variable-length uint32 arrays over gloo"""
    n = torch.tensor([len(a)], dtype=torch.int64)
    ns = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(ns, n)
    m = int(max(x.item() for x in ns))
    buf = torch.zeros(m, dtype=torch.int64)
    buf[: len(a)] = torch.from_numpy(a.astype(np.int64))
    out = [torch.zeros(m, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(out, buf)
    return [o[: int(k.item())].numpy().astype(np.uint32) for o, k in zip(out, ns)]


def _worker(rank, world, port, V, n_train, batch, replicate):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        from xgnn_b200 import dataset as D
        from xgnn_b200._clib import check, lib

        # ---- work split: bench.AlignedShuffler (what bench.py's epochs use) and the engine's routine (C-ABI) ----
        train = D.train_set(V, n_train)
        sh = bench.AlignedShuffler(train, rank, world, batch)
        for _ in range(3):
            epoch_batches = sh.next_epoch()
        mine = np.concatenate(epoch_batches)
        assert all(len(b) == batch for b in epoch_batches[:-1]) and len(epoch_batches) == sh.steps
        parts = _gather_arrays(mine, world)
        assert len({len(p) for p in parts}) == 1, "slices must be equal (aligned shuffler pads to a multiple of P)"
        allv = np.concatenate(parts)
        pad = (-n_train) % world
        assert len(allv) == n_train + pad
        assert set(allv.tolist()) == set(train.tolist())  # every train node is somebody's seed; only padding repeats
        # the reference's own shuffle: identical permutation in every process, contiguous slice per worker
        data = np.concatenate([train, train[:pad]]) if pad else train.copy()
        for epoch in range(2):
            check(lib.xg_aligned_shuffle_epoch(data.ctypes.data, len(data), epoch))
        n_local = len(data) // world
        slices = _gather_arrays(data[rank * n_local:(rank + 1) * n_local], world)
        assert np.array_equal(np.concatenate(slices), data)
        steps = -(-n_local // batch)
        t = torch.tensor([steps], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert int(t.item()) == steps  # num_local_step agrees on every worker (dist_shuffler_aligned.h:38-40)

        # ---- data split: topology rows and cache ranks ----
        rows = np.arange(rank, V, world, dtype=np.uint32)  # part `rank` holds v % P == rank, row v / P
        assert len(rows) == lib.xg_part_num_rows(V, rank, world)
        owned = _gather_arrays(rows, world)
        assert np.array_equal(np.sort(np.concatenate(owned)), np.arange(V, dtype=np.uint32))
        n_cached = int(V * 0.8)
        n_rep = int(V * replicate)
        ranks = D.cache_rank_nodes(V, n_cached, n_rep)
        every = _gather_arrays(ranks, world)
        assert all(np.array_equal(e, ranks) for e in every), "the rank order must be identical in every process"
        assert np.array_equal(ranks[:n_rep], np.arange(n_rep, dtype=np.uint32))  # hot tier keeps the policy order
        assert sorted(ranks.tolist()) == list(range(n_cached))
        my_part = ranks[n_rep:][rank::world]
        allp = _gather_arrays(my_part, world)
        assert sum(len(p) for p in allp) == n_cached - n_rep
        assert np.array_equal(np.sort(np.concatenate(allp)), np.sort(ranks[n_rep:]))
        # routing rule: rank r >= n_rep lives in part (r - n_rep) % P at row (r - n_rep) / P
        table = np.full(V, 0xFFFFFFFF, np.uint32)
        table[ranks] = np.arange(n_cached, dtype=np.uint32)
        for node in my_part[:: max(1, len(my_part) // 50)]:
            r = int(table[node]) - n_rep
            assert r % world == rank and my_part[r // world] == node
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,replicate", [(2, 0.0), (2, 0.3), (3, 0.25)])
def test_work_and_data_split_tile_the_job(world, replicate):
    mp.spawn(_worker, args=(world, _free_port(), 20011, 1001, 64, replicate), nprocs=world, join=True)


def test_arch5_queue_protocol_model_is_deadlock_free():
    """Thread model of the arch5 Task hand-over (tools/arch5_queue_model.py = the semaphore protocol of engine.cc:
    stage ring per sampler, ticketed descriptor ring, one synchronous and one prefetching trainer): with the shipped rule
    - a trainer releases every stage it holds before a blocking dequeue - every Task is delivered exactly once and nothing
    ever waits forever, across random sampler / consumer speeds (2 samplers x 2 trainers, the configuration in which the
    first version of the extract loop deadlocked on a B200 box)."""
    sys.path.insert(0, str(ROOT / "tools"))
    import arch5_queue_model as M
    M.TIMEOUT = 3.0
    for seed in range(40):
        assert M.run_trial(seed, hold_stage_while_dequeuing=False), f"seed {seed}"


def _ics22_worker(rank, world, port, V, n_cached, clique, alpha):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ctypes

        import oracle_lib as O
        from xgnn_b200._clib import check, lib
        rng = np.random.default_rng(99)  # same graph in every process, as the shared dataset is
        indices = np.minimum((rng.pareto(1.1, V * 10) * 4).astype(np.int64), V - 1).astype(np.uint32)
        rank_nodes, rank_w = O.ics22_in_degree_rank(indices, V)
        table, slots, tot = np.empty(V, np.uint32), np.empty(n_cached, np.uint32), ctypes.c_uint32(0)
        check(lib.xg_ics22_placement(rank_nodes.ctypes.data, rank_w.ctypes.data, V, n_cached, world, clique, alpha, rank, 3,
                                     table.ctypes.data, slots.ctypes.data, ctypes.addressof(tot)))
        # what FeatureLoad shares over IPC: every worker's slot list; what each worker keeps private: its map
        all_slots = _gather_arrays(slots, world)
        totals = _gather_arrays(np.array([tot.value], np.uint32), world)
        assert len({int(t[0]) for t in totals}) == 1, "every worker must see the same number of cached nodes"
        cached = np.nonzero(table != 0xFFFFFFFF)[0]
        assert len(cached) == tot.value
        owner, row = table[cached] >> 29, table[cached] & ((1 << 29) - 1)
        first = rank // clique * clique
        assert np.all((owner >= first) & (owner < first + clique)), "reads stay inside the clique"
        for v, o, r in zip(cached, owner, row):
            assert all_slots[o][r] == v, "my map names a slot that holds another node"
        # the hottest nodes that were not split away are served from my own HBM
        assert table[rank_nodes[0]] >> 29 == rank or tot.value > n_cached
        # distinct nodes held by the clique == total (nothing is cached twice once its slot was split)
        held = np.unique(np.concatenate(all_slots[first:first + clique]))
        assert len(held) == tot.value
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,clique,alpha", [(2, 2, 0.0), (2, 1, 0.0), (4, 2, 0.4)])
def test_ics22_views_of_all_workers_agree(world, clique, alpha):
    """Each worker derives only ITS view of the ICS'22 clique placement (xg_ics22_placement); together the views must
    describe one consistent layout: a (owner, row) named by any worker's map holds that node in the owner's slot list."""
    mp.spawn(_ics22_worker, args=(world, _free_port(), 4001, 300, clique, alpha), nprocs=world, join=True)
