"""KHop1, weighted KHop (alias / prefix / hash-dedup) and random-walk top-K.

The reference has no CPU twin and no test for these ("parity unpinned"); the oracle
restates the CUDA files and takes the random words as an input, so the GPU tests replay the kernel's own
Philox words through it and compare BIT FOR BIT; the CPU tests pin the oracle's law on small cases.
"""
import ctypes

import numpy as np
import pytest

import oracle_lib as O
import philox
import synth


@pytest.fixture(scope="module")
def graph():
    V, E = 3000, 30000
    indptr, indices = synth.make_graph(V, E, seed=21, max_deg=40)
    w = synth.edge_weights(indptr, indices)
    prob, alias = synth.alias_tables(indptr, indices, w)
    return dict(V=V, indptr=indptr, indices=indices, w=w, prob=prob, alias=alias, prefix=synth.prefix_table(indptr, w))


# ------------------------------------------------------------------------------------------------ CPU
def test_philox_numpy_matches_library_and_known_answers():
    from xgnn_b200._clib import lib
    assert [hex(x) for x in philox.philox4x32(0, 0, 0, 0, 0, 0)] == ['0x6627e8d5', '0xe169c58d', '0xbc57ac4c', '0x9b00dbd8']
    assert lib.xg_sample_salt() == philox.SALT
    out = (ctypes.c_uint32 * 4)()
    for seed, key, c in [(0, 0, (0, 0, 0, 0)), (7, 3, (5, 1, philox.SALT, 0)), (2**63 + 5, 2**40, (2**32 - 1, 9, 77, 1))]:
        lib.xg_philox4x32(seed, key, *c, out)
        lo, hi = philox.batch_key(seed, key)
        assert list(out) == list(philox.philox4x32(lo, hi, *c))


def test_oracle_khop1_sorted_by_src_and_adjacent_dedup(graph):
    inp = np.array([50, 7, 7, 1200, 3], np.uint32)  # unsorted, with a duplicate seed
    rnd = philox.words_slots(1, 0, len(inp), 6)
    src, dst = O.sample_with_replacement(1, graph["indptr"], graph["indices"], inp, 6, rnd)
    assert np.all(np.diff(src.astype(np.int64)) >= 0)
    for v in np.unique(src):
        b, e = graph["indptr"][v], graph["indptr"][v + 1]
        assert set(dst[src == v]) <= set(graph["indices"][b:e])
    pairs = list(zip(src, dst))
    assert all(pairs[i] != pairs[i + 1] for i in range(len(pairs) - 1))


def test_oracle_alias_law(graph):
    """alias draws follow the edge weights: chi-square over many draws of one node"""
    indptr, indices = graph["indptr"], graph["indices"]
    v = int(np.argmax(np.diff(indptr.astype(np.int64))))
    b, e = int(indptr[v]), int(indptr[v + 1])
    n = 40000
    rng = np.random.default_rng(5)
    rnd = rng.integers(0, 2**32, size=(n, 4), dtype=np.uint32)
    # one draw per "node copy": fanout 1, n copies of v -> no adjacent-dedup interference except equal neighbours
    src, dst = O.sample_with_replacement(2, indptr, indices, np.full(1, v, np.uint32), n, rnd,
                                         prob=graph["prob"], alias=graph["alias"])
    # adjacent duplicates collapse, so count through an undeduplicated replay instead
    k = rnd[:, 0] % (e - b)
    r = ((rnd[:, 1] >> 8).astype(np.float32) + 1) / np.float32(16777216.0)
    pick = np.where(r < graph["prob"][b + k], indices[b + k], graph["alias"][b + k])
    want = {}
    for j in range(b, e):
        want[indices[j]] = want.get(indices[j], 0.0) + graph["w"][j]
    tot = sum(want.values())
    chi = sum((np.sum(pick == u) - n * wu / tot) ** 2 / (n * wu / tot) for u, wu in want.items())
    assert chi < 2.5 * len(want) + 30, chi
    assert set(dst) <= set(want)


def test_oracle_random_walk_deterministic_graphs():
    # chain 0 -> 1 -> 2 -> 3 (3 has no neighbours): every walk visits 1, 2, 3 once when restart = 0
    indptr = np.array([0, 1, 2, 3, 3], np.uint32)
    indices = np.array([1, 2, 3], np.uint32)
    rnd = philox.words_walks(1, 0, 2, 4, 3)
    src, dst, cnt = O.random_walk_topk(indptr, indices, np.array([0, 2], np.uint32), 3, 0.0, 4, 5, rnd)
    assert list(src) == [0, 0, 0, 2] and list(dst) == [1, 2, 3, 3] and list(cnt) == [4, 4, 4, 4]
    # restart = 1: one step per walk
    src, dst, cnt = O.random_walk_topk(indptr, indices, np.array([0], np.uint32), 3, 1.0, 4, 5, rnd)
    assert list(dst) == [1] and list(cnt) == [4]
    # star: K limits the output, counts descending
    indptr = np.array([0, 6, 6, 6, 6, 6, 6, 6], np.uint32)
    indices = np.array([1, 2, 3, 4, 5, 6], np.uint32)
    rnd = philox.words_walks(3, 1, 1, 64, 1)
    src, dst, cnt = O.random_walk_topk(indptr, indices, np.array([0], np.uint32), 1, 0.5, 64, 3, rnd)
    assert len(dst) == 3 and np.all(np.diff(cnt.astype(np.int64)) <= 0) and len(set(dst)) == 3


# ------------------------------------------------------------------------------------------------ GPU
gpu = pytest.mark.gpu


def _dev(a):
    import torch
    a = np.ascontiguousarray(a)
    return torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a).cuda()


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.fixture(scope="module")
def dgraph(graph):
    from xgnn_b200 import ops
    return ops, ops.make_graph(_dev(graph["indptr"]), _dev(graph["indices"]))


def _inputs(graph, n, seed, dup=True):
    rng = np.random.default_rng(seed)
    inp = rng.permutation(graph["V"])[:n].astype(np.uint32)
    if dup and n > 10:
        inp[5] = inp[1]
    return inp


@gpu
@pytest.mark.parametrize("n,fanout", [(1, 1), (37, 5), (1000, 10), (2500, 25)])
def test_khop1_bit_exact_against_oracle_replay(graph, dgraph, n, fanout):
    ops, g = dgraph
    inp = _inputs(graph, n, n)
    src, dst = ops.GPUSampleKHop1(g, _dev(inp), fanout, seed=9, batch_key=4)
    osrc, odst = O.sample_with_replacement(1, graph["indptr"], graph["indices"], inp, fanout,
                                           philox.words_slots(9, 4, n, fanout))
    assert np.array_equal(_u32(src), osrc) and np.array_equal(_u32(dst), odst)


@gpu
@pytest.mark.parametrize("V,n", [(3000, 4096), (3000, 4097), (300_000, 20001), (300_000, 8192), (70_000_000, 9000)])
def test_khop1_sort_across_tiles_passes_and_heavy_duplicates(V, n):
    """the stable radix sort in front of the draws (sampling_ext.cu): inputs longer than one 4096-element tile, ids drawn
    WITH replacement from a small pool (long runs of equal keys: stability decides the output), and id ranges that need
    2 (12 bits), 3 (19 bits) and 3 full (27 bits) passes - the result must still be the oracle's, bit for bit"""
    from xgnn_b200 import ops
    rng = np.random.default_rng(V + n)
    if V > 1_000_000:  # a huge, almost empty graph: only the pool's nodes have neighbours
        pool = np.unique(rng.integers(0, V, size=600))[:500].astype(np.int64)
        deg = np.zeros(V, np.int64)
        deg[pool] = rng.integers(1, 9, size=len(pool))
        indptr = np.zeros(V + 1, np.uint32)
        np.cumsum(deg, out=indptr[1:])
        indices = rng.integers(0, V, size=int(indptr[-1])).astype(np.uint32)
    else:
        indptr, indices = synth.make_graph(V, 4 * V, seed=5, max_deg=30)
        pool = rng.permutation(V)[:700]
    g = ops.make_graph(_dev(indptr), _dev(indices))
    inp = pool[rng.integers(0, len(pool), size=n)].astype(np.uint32)
    fanout = 3
    src, dst = ops.GPUSampleKHop1(g, _dev(inp), fanout, seed=11, batch_key=6)
    osrc, odst = O.sample_with_replacement(1, indptr, indices, inp, fanout, philox.words_slots(11, 6, n, fanout))
    assert np.array_equal(_u32(src), osrc) and np.array_equal(_u32(dst), odst)
    if V == 3000:  # a handle that does not say how many nodes there are: all 32 key bits are sorted (4 passes)
        g.num_node = 0
        src4, dst4 = ops.GPUSampleKHop1(g, _dev(inp), fanout, seed=11, batch_key=6)
        assert np.array_equal(_u32(src4), osrc) and np.array_equal(_u32(dst4), odst)


@gpu
@pytest.mark.parametrize("mode", [2, 4])
@pytest.mark.parametrize("n,fanout", [(64, 3), (2000, 10)])
def test_weighted_bit_exact_against_oracle_replay(graph, dgraph, mode, n, fanout):
    ops, g = dgraph
    inp = _inputs(graph, n, n + mode)
    kw = dict(prob_table=_dev(graph["prob"]), alias_table=_dev(graph["alias"])) if mode == 2 else \
        dict(prob_prefix_table=_dev(graph["prefix"]))
    src, dst = ops.GPUSampleWeightedKHop(g, _dev(inp), fanout, sample_type=mode, seed=2, batch_key=8, **kw)
    osrc, odst = O.sample_with_replacement(mode, graph["indptr"], graph["indices"], inp, fanout,
                                           philox.words_slots(2, 8, n, fanout), prob=graph["prob"],
                                           alias=graph["alias"], prefix=graph["prefix"])
    assert np.array_equal(_u32(src), osrc) and np.array_equal(_u32(dst), odst)


@gpu
@pytest.mark.parametrize("n,fanout", [(50, 4), (1500, 10), (300, 40)])
def test_weighted_hash_dedup_bit_exact_against_oracle_replay(graph, dgraph, n, fanout):
    ops, g = dgraph
    inp = _inputs(graph, n, 3 * n)
    src, dst = ops.GPUSampleWeightedKHop(g, _dev(inp), fanout, sample_type=6, prob_table=_dev(graph["prob"]),
                                         alias_table=_dev(graph["alias"]), seed=6, batch_key=1)
    rnd = philox.words_slots(6, 1, n, 16 * fanout, c3=1)
    osrc, odst = O.sample_weighted_hash_dedup(graph["indptr"], graph["indices"], inp, fanout, rnd, graph["prob"],
                                              graph["alias"])
    assert np.array_equal(_u32(src), osrc) and np.array_equal(_u32(dst), odst)
    # distinct neighbours per node
    s, d = _u32(src), _u32(dst)
    for v in np.unique(s)[:50]:
        if np.sum(inp == v) == 1:
            assert len(set(d[s == v])) == np.sum(s == v)


@gpu
def test_hash_dedup_terminates_when_fewer_distinct_outcomes_than_fanout(dgraph):
    """the reference spins forever here"""
    ops, _ = dgraph
    indptr = np.array([0, 6], np.uint32)
    indices = np.array([0, 0, 0, 0, 0, 0], np.uint32)  # six parallel edges to one node, fanout 3 < degree
    g = ops.make_graph(_dev(indptr), _dev(indices))
    src, dst = ops.GPUSampleWeightedKHop(g, _dev(np.array([0], np.uint32)), 3, sample_type=6,
                                         prob_table=_dev(np.ones(6, np.float32)), alias_table=_dev(indices))
    assert list(_u32(dst)) == [0]


@gpu
@pytest.mark.parametrize("n,W,L,K,p", [(1, 4, 3, 5, 0.5), (500, 4, 3, 5, 0.5), (2000, 10, 5, 8, 0.2), (77, 40, 3, 1, 0.0),
                                       (64, 33, 7, 64, 0.9)])
def test_random_walk_topk_bit_exact_against_oracle_replay(graph, dgraph, n, W, L, K, p):
    ops, g = dgraph
    inp = _inputs(graph, n, 7 * n + W)
    src, dst, cnt = ops.GPUSampleRandomWalk(g, _dev(inp), L, p, W, K, seed=12, batch_key=5)
    osrc, odst, ocnt = O.random_walk_topk(graph["indptr"], graph["indices"], inp, L, p, W, K,
                                          philox.words_walks(12, 5, n, W, L))
    assert np.array_equal(_u32(src), osrc) and np.array_equal(_u32(dst), odst) and np.array_equal(_u32(cnt), ocnt)


@gpu
def test_ext_samplers_empty_input_and_ggms_routing(graph, dgraph):
    import torch
    ops, g = dgraph
    e = torch.empty(0, dtype=torch.int32, device="cuda")
    assert ops.GPUSampleKHop1(g, e, 5)[0].numel() == 0
    assert ops.GPUSampleRandomWalk(g, e, 3, 0.5, 4, 5)[0].numel() == 0
    # KHop1 and random walk over a partitioned topology read the same lists -> same result as flat
    P, indptr, indices = 3, graph["indptr"], graph["indices"]
    ncn = O.num_cache_node(indptr, 0.6)
    parts = [O.partition_csr(indptr, indices, ncn, p, P) for p in range(P)]
    keep = [(_dev(pi), _dev(px if len(px) else np.zeros(1, np.uint32))) for pi, px in parts]
    gd = ops.make_graph(part_indptr=[k[0] for k in keep], part_indices=[k[1] for k in keep], num_cache_node=ncn,
                        num_node=graph["V"], host_indptr=_dev(indptr), host_indices=_dev(indices))
    inp = _dev(_inputs(graph, 800, 99))
    a, b = ops.GPUSampleKHop1(g, inp, 7, seed=3), ops.GPUSampleKHop1(gd, inp, 7, seed=3)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    a, b = ops.GPUSampleRandomWalk(g, inp, 3, 0.5, 4, 5, seed=3), ops.GPUSampleRandomWalk(gd, inp, 3, 0.5, 4, 5, seed=3)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
