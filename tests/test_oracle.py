"""CPU: pins the oracle (oracle/oracle.c) against
  (1) the committed golden fixtures generated from the reference's own CPU code (tests/golden/make_golden.py),
  (2) the reference's own code when oracle/_ref/libxgnn_ref.so is present (this container, or shipped prebuilt),
  (3) the reference's gtest vectors for OrderedHashTable (samgraph/unittest/test_hashmap.cc:84-276).
"""
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as O
import synth

GOLD = np.load(Path(__file__).parent / "golden" / "ref_path_small.npz")
needs_ref = pytest.mark.skipif(O.ref is None, reason="oracle/_ref not built (no /root/reference here)")


# ---- (1) golden fixtures ----------------------------------------------------------------------
def test_khop0_full_neighbourhood_matches_golden():
    src, dst = O.sample_khop0(GOLD["indptr"], GOLD["indices"], GOLD["inp"], int(GOLD["fanout"]))
    assert np.array_equal(src, GOLD["khop0_src"])
    assert np.array_equal(dst, GOLD["khop0_dst"])


def test_hashtable0_matches_golden():
    t = O.Table(4000)
    t.populate(GOLD["fill1"])
    assert np.array_equal(t.unique(), GOLD["uniq1"])
    t.populate(GOLD["fill2"])
    assert np.array_equal(t.unique(), GOLD["uniq2"])
    ms, md = t.map_edges(GOLD["fill2"], GOLD["fill2"][::-1].copy())
    assert np.array_equal(ms, GOLD["map_src"])
    assert np.array_equal(md, GOLD["map_dst"])


def test_extract_matches_golden():
    V = len(GOLD["indptr"]) - 1
    idx = GOLD["idx"]
    assert np.array_equal(O.extract(synth.make_features(V, 100, np.float32), idx), GOLD["x32"])
    assert np.array_equal(O.extract(synth.make_features(V, 24, np.float16), idx).view(np.uint16),
                          GOLD["x16"].view(np.uint16))
    assert np.array_equal(O.extract(synth.make_labels(V), idx), GOLD["xl"])


def test_minibatch_matches_golden():
    f = int(GOLD["b_fanout"])
    mb = O.minibatch(GOLD["b_indptr"], GOLD["b_indices"], GOLD["b_seeds"], [f, f, f], seed=3)
    counts = GOLD["b_counts"].reshape(3, 3)
    for l in range(3):
        assert (mb["num_src"][l], mb["num_dst"][l], mb["num_edge"][l]) == tuple(int(x) for x in counts[l])
        assert np.array_equal(mb["row"][l], GOLD[f"b_row{l}"])
        assert np.array_equal(mb["col"][l], GOLD[f"b_col{l}"])
    assert np.array_equal(mb["input_nodes"], GOLD["b_input_nodes"])
    feat = O.extract(synth.make_features(4000, 16, np.float32), mb["input_nodes"])
    assert np.array_equal(feat, GOLD["b_feat"])
    assert np.array_equal(O.extract(synth.make_labels(4000), GOLD["b_seeds"]), GOLD["b_label"])
    assert sum(mb["num_edge"]) == int(GOLD["b_total"])


# ---- (2) live reference ------------------------------------------------------------------------
@needs_ref
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_oracle_vs_reference_live(seed):
    rng = np.random.default_rng(seed)
    indptr, indices = synth.make_graph(5000, 40000, seed=seed, max_deg=40)
    inp = rng.integers(0, 5000, size=900).astype(np.uint32)
    fan = int(np.diff(indptr.astype(np.int64)).max())
    rs, rd = O.ref_sample_khop0(indptr, indices, inp, fan)
    os_, od = O.sample_khop0(indptr, indices, inp, fan)
    assert np.array_equal(rs, os_) and np.array_equal(rd, od)
    # small fanout: sizes and src agree; dst must be real neighbours
    rs, rd = O.ref_sample_khop0(indptr, indices, inp, 5)
    os_, od = O.sample_khop0(indptr, indices, inp, 5)
    assert np.array_equal(rs, os_)
    rt, ot = O.RefTable(0, 6000), O.Table(6000)
    for _ in range(3):
        a = rng.integers(0, 5000, size=2000).astype(np.uint32)
        rt.populate(a)
        ot.populate(a)
        assert np.array_equal(rt.unique(), ot.unique())
    a = rt.unique()[rng.integers(0, rt.num_items(), size=3000)]
    b = rt.unique()[rng.integers(0, rt.num_items(), size=3000)]
    r1, r2 = rt.map_edges(a, b)
    o1, o2 = ot.map_edges(a, b)
    assert np.array_equal(r1, o1) and np.array_equal(r2, o2)
    for dt, dim in ((np.float32, 100), (np.float16, 768), (np.int64, 1), (np.uint8, 7)):
        tab = (rng.integers(0, 100, size=(5000, dim)) if dim > 1 else rng.integers(0, 100, size=5000)).astype(dt)
        idx = rng.integers(0, 5000, size=333).astype(np.uint32)
        assert np.array_equal(O.ref_extract(tab, idx), O.extract(tab, idx))


@needs_ref
def test_reference_hashtable2_is_set_equal():
    """CPUHashTable2 (arch0 default) has a racy winner: only the SET and the prefix property hold."""
    rng = np.random.default_rng(5)
    rt, ot = O.RefTable(2, 6000), O.Table(6000)
    a = rng.integers(0, 5000, size=4000).astype(np.uint32)
    rt.populate(a)
    ot.populate(a)
    assert np.array_equal(np.sort(rt.unique()), np.sort(ot.unique()))


# ---- (3) reference gtest vectors (unittest/test_hashmap.cc:98-117,168-195) ----------------------
def _check_unique_props(t, inputs_so_far, prev_unique):
    u = t.unique()
    assert set(u.tolist()) == set(np.concatenate(inputs_so_far).tolist())
    assert len(set(u.tolist())) == len(u)
    assert np.array_equal(u[: len(prev_unique)], prev_unique)
    assert np.array_equal(t.lookup(u), np.arange(len(u), dtype=np.uint32))
    return u.copy()


def test_reference_gtest_vectors_on_oracle():
    t = O.Table(2048)
    v1 = np.array([1, 2, 3, 4, 5, 2, 5, 1, 10, 233], np.uint32)
    v2 = np.array([1, 2, 3, 6, 9, 2, 8, 1, 3, 7, 1023], np.uint32)
    t.populate(v1)
    u1 =_check_unique_props(t, [v1], np.empty(0, np.uint32))
    t.populate(v2)
    _check_unique_props(t, [v1, v2], u1)
    t.reset()
    v3 = np.array([1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 1, 2, 3, 4, 5], np.uint32)
    t.populate(v3)
    u3 = _check_unique_props(t, [v3], np.empty(0, np.uint32))
    assert len(u3) == 13
    v4 = np.arange(1, 513, dtype=np.uint32)
    t.populate(v4)
    _check_unique_props(t, [v3, v4], u3)


# ---- routing / partition restatements: self-consistency ------------------------------------------
@pytest.mark.parametrize("P", [1, 2, 3, 8])
def test_partitioned_topology_reassembles(P):
    indptr, indices = synth.make_graph(2000, 16000, seed=11, max_deg=50)
    for ratio in (0.0, 0.37, 1.0):
        ncn = O.num_cache_node(indptr, ratio)
        want = int(int(indptr[-1]) * ratio)
        assert ncn == 0 or indptr[ncn - 1] < want
        assert indptr[ncn] >= want or ncn == 2000
        parts = [O.partition_csr(indptr, indices, ncn, p, P) for p in range(P)]
        for v in range(0, 2000, 7):
            p, r = O.route_node(v, P, ncn)
            lst = indices[indptr[v]: indptr[v + 1]]
            if v < ncn:
                assert (p, r) == (v % P, v // P)
                pi, px = parts[p]
                assert np.array_equal(px[pi[r]: pi[r + 1]], lst)
            else:
                assert (p, r) == (P, v)


@pytest.mark.parametrize("P", [1, 4])
def test_partitioned_features_reassemble(P):
    V, dim = 3000, 12
    feat = synth.make_features(V, dim)
    indptr, _ = synth.make_graph(V, 20000, seed=4)
    rank = synth.degree_rank(indptr)
    ncached = 1700
    table = O.build_cache_table(rank, ncached, V)
    parts = [O.partition_features(feat, rank, ncached, p, P) for p in range(P)]
    nodes = np.random.default_rng(0).integers(0, V, size=5000).astype(np.uint32)
    out = O.extract_ggms(feat, nodes, table, parts)
    assert np.array_equal(out, feat[nodes])
    ms, md, cs, cd = O.miss_cache_index(table, nodes)
    assert len(ms) + len(cs) == len(nodes)
    assert np.array_equal(nodes[md], ms) and np.all(table[ms] == O.EMPTY)
    assert np.array_equal(table[nodes[cd]], cs)
    assert np.all(np.diff(md.astype(np.int64)) > 0) and np.all(np.diff(cd.astype(np.int64)) > 0)


def test_two_tier_placement_oracle_reads_the_same_rows():
    """The replicated + partitioned placement (B200 extension) must never change WHICH row a node resolves to."""
    import synth
    V, dim = 5000, 12
    feat = synth.make_features(V, dim)
    indptr, _ = synth.make_graph(V, 40000, seed=9)
    rank = synth.degree_rank(indptr)
    nodes = np.random.default_rng(3).integers(0, V, size=4001).astype(np.uint32)
    for ncached, nrep, P in [(5000, 0, 4), (5000, 1250, 8), (3000, 3000, 2), (2000, 100, 1), (0, 0, 3)]:
        table = O.build_cache_table(rank, ncached, V)
        replica, parts = O.hybrid_placement(feat, rank, ncached, nrep, P)
        assert len(replica) == nrep and sum(len(p) for p in parts) == ncached - nrep
        got = O.extract_hybrid(feat, nodes, table, replica, parts)
        assert np.array_equal(got.view(np.uint32), O.extract(feat, nodes).view(np.uint32))
        if nrep == 0 and ncached:  # degenerates to the reference's single-tier routing
            ref_parts = [O.partition_features(feat, rank, ncached, p, P) for p in range(P)]
            assert all(np.array_equal(a, b) for a, b in zip(parts, ref_parts))


def test_csc_indptr_oracle_matches_scipy_conversion():
    """coo_to_csc_indptr restates the caller-side COO -> CSC conversion: check it against scipy's."""
    import scipy.sparse as sp
    rng = np.random.default_rng(2)
    num_dst, num_src = 50, 400
    counts = rng.integers(0, 7, num_dst)
    col = np.repeat(np.arange(num_dst), counts).astype(np.uint32)  # node-major
    row = rng.integers(0, num_src, len(col)).astype(np.uint32)
    indptr = O.coo_to_csc_indptr(col, num_dst)
    m = sp.coo_matrix((np.ones(len(col)), (row, col)), shape=(num_src, num_dst)).tocsc()
    assert np.array_equal(indptr, m.indptr.astype(np.uint32))
    for d in range(num_dst):  # same neighbour multiset per destination
        assert sorted(row[indptr[d]:indptr[d + 1]].tolist()) == sorted(row[col == d].tolist())


def test_presample_rank_oracle_order():
    """freq descending, ties by node id descending (std::greater on {low: id, high: freq}, pre_sampler.cc:113-119)"""
    nodes, freq = O.presample_rank(6, [np.array([1, 3], np.uint32), np.array([3, 4], np.uint32), np.array([3, 1], np.uint32)])
    assert nodes.tolist() == [3, 1, 4, 5, 2, 0] and freq.tolist() == [3, 2, 1, 0, 0, 0]


@pytest.mark.parametrize("V,n_cached,num_dev,clique,alpha", [
    (200, 40, 4, 4, 0.0), (200, 40, 4, 2, 0.5), (300, 25, 8, 4, 0.9), (120, 30, 2, 2, 0.3), (64, 64, 4, 4, 0.0),
    (150, 10, 6, 3, 0.0), (100, 20, 3, 1, 0.5), (90, 0, 2, 2, 0.0), (97, 5, 8, 8, 0.05)])
def test_ics22_placement_matches_the_all_device_restatement(V, n_cached, num_dev, clique, alpha):
    """xg_ics22_placement (one worker's view, single pass) against the oracle that keeps the reference's per-device
    maps (cuda/ics22_song_dist_feature.cc:22-167): same compact table, same slot contents, same total; and the
    layout is consistent - whatever GPU/row a worker's table names holds that node."""
    import ctypes
    from xgnn_b200._clib import check, lib
    rng = np.random.default_rng(V * 31 + num_dev)
    E = V * 12
    indices = np.minimum((rng.pareto(1.2, E) * 3).astype(np.int64), V - 1).astype(np.uint32)  # skewed, many ties
    rank_nodes, rank_w = O.ics22_in_degree_rank(indices, V)
    weight = np.bincount(indices, minlength=V).astype(np.float64)
    ranking, dev_map, idx_map, cached, total = O.ics22_placement(weight, n_cached, num_dev, clique, alpha)
    assert np.array_equal(ranking, rank_nodes)
    tables, slots = [], []
    for w in range(num_dev):
        table = np.empty(V, np.uint32)
        slot_nodes = np.empty(max(n_cached, 1), np.uint32)
        tot = ctypes.c_uint32(0)
        check(lib.xg_ics22_placement(rank_nodes.ctypes.data, rank_w.ctypes.data, V, n_cached, num_dev, clique, alpha, w, 3,
                                     table.ctypes.data, slot_nodes.ctypes.data, ctypes.addressof(tot)))
        assert tot.value == total
        assert np.array_equal(table, O.ics22_compact_table(ranking, dev_map, idx_map, total, w))
        assert np.array_equal(slot_nodes[:n_cached], cached[w])
        tables.append(table)
        slots.append(slot_nodes[:n_cached])
    assert int((tables[0] != 0xFFFFFFFF).sum()) == total
    if clique > 1 and alpha == 0.0 and 0 < n_cached and n_cached * clique <= V:
        assert total == n_cached * clique  # every slot split: the clique caches c x as many distinct nodes
    for w in range(num_dev):
        cached_nodes = np.nonzero(tables[w] != 0xFFFFFFFF)[0]
        owner, row = tables[w][cached_nodes] >> 29, tables[w][cached_nodes] & ((1 << 29) - 1)
        assert np.all(owner // clique == w // clique), "a worker only reads inside its clique"
        for v, o, r in zip(cached_nodes, owner, row):
            assert slots[o][r] == v


def test_ics22_placement_rejects_bad_arguments():
    import ctypes
    from xgnn_b200._clib import lib
    nodes, w = np.arange(8, dtype=np.uint32)[::-1].copy(), np.arange(8, dtype=np.uint32)[::-1].copy()
    table, slots = np.zeros(8, np.uint32), np.zeros(8, np.uint32)
    tot = ctypes.c_uint32(0)
    args = lambda **k: (nodes.ctypes.data, w.ctypes.data, 8, k.get("n", 4), k.get("dev", 4), k.get("c", 2), 0.0,
                        k.get("w", 0), k.get("b", 3), table.ctypes.data, slots.ctypes.data, ctypes.addressof(tot))
    assert lib.xg_ics22_placement(*args()) == 0
    assert lib.xg_ics22_placement(*args(c=3)) != 0   # 4 GPUs are not a multiple of 3
    assert lib.xg_ics22_placement(*args(w=4)) != 0   # worker out of range
    assert lib.xg_ics22_placement(*args(b=1)) != 0   # 1 bit cannot name 4 GPUs
    assert lib.xg_ics22_placement(*args(n=9)) != 0   # more slots than nodes


# ---- (4) the cache-rank routines pinned against the reference's own code ----------------------------------------------
# ICS22SongPlacementSolver (cuda/ics22_song_dist_feature.cc:22-167) and dist::PreSampler (dist/pre_sampler.cc:40-165)
# compiled in place (oracle/ref_host_shim.cc): golden fixtures generated from them (tests/golden/make_golden_rank.py)
# and, where oracle/_ref/libxgnn_ref_cuda.so is present, the live code.
RANK_GOLD = np.load(Path(__file__).parent / "golden" / "ref_rank_small.npz")
needs_refcu = pytest.mark.skipif(O.refcu is None, reason="oracle/_ref/libxgnn_ref_cuda.so not built")
sys_path_golden = str(Path(__file__).parent / "golden")


def _rank_cases():
    import sys
    if sys_path_golden not in sys.path:
        sys.path.insert(0, sys_path_golden)
    import make_golden_rank as G
    return G


def _product_ics22_views(rank_nodes, rank_w, V, n_cached, num_dev, clique, alpha):
    import ctypes
    from xgnn_b200._clib import check, lib
    views = []
    for w in range(num_dev):
        table = np.empty(V, np.uint32)
        slot_nodes = np.empty(max(n_cached, 1), np.uint32)
        tot = ctypes.c_uint32(0)
        check(lib.xg_ics22_placement(rank_nodes.ctypes.data, rank_w.ctypes.data, V, n_cached, num_dev, clique, alpha, w, 3,
                                     table.ctypes.data, slot_nodes.ctypes.data, ctypes.addressof(tot)))
        views.append((table, slot_nodes[:n_cached], tot.value))
    return views


@pytest.mark.parametrize("case", range(8))
def test_ics22_oracle_and_product_match_reference_golden(case):
    """restatement == reference solver output, and xg_ics22_placement's per-worker view == the compact table built
    from the REFERENCE's maps (cuda_cache_manager_host.cc:203-222), for every worker"""
    G = _rank_cases()
    V, n_cached, num_dev, clique, alpha = G.ICS22_CASES[case]
    indices, weight = G.ics22_weight(V, num_dev)
    g = {k: RANK_GOLD[f"ics22_{case}_{k}"] for k in ("ranking", "dev_map", "idx_map", "cached", "total")}
    ranking, dev_map, idx_map, cached, total = O.ics22_placement(weight, n_cached, num_dev, clique, alpha)
    assert np.array_equal(ranking, g["ranking"]) and total == int(g["total"])
    assert np.array_equal(dev_map, g["dev_map"]) and np.array_equal(idx_map, g["idx_map"])
    assert np.array_equal(cached, g["cached"])
    rank_nodes, rank_w = O.ics22_in_degree_rank(indices, V)
    assert np.array_equal(rank_nodes, g["ranking"])
    for w, (table, slots, tot) in enumerate(_product_ics22_views(rank_nodes, rank_w, V, n_cached, num_dev, clique, alpha)):
        assert tot == int(g["total"])
        assert np.array_equal(table, O.ics22_compact_table(g["ranking"], g["dev_map"], g["idx_map"], int(g["total"]), w))
        assert np.array_equal(slots, g["cached"][w])


@needs_refcu
@pytest.mark.parametrize("V,n_cached,num_dev,clique,alpha", [
    (257, 33, 4, 2, 0.0), (257, 33, 4, 4, 0.25), (500, 100, 8, 8, 0.0), (500, 499, 8, 2, 0.0), (80, 1, 2, 2, 0.0),
    (1000, 300, 6, 3, 0.6), (333, 50, 5, 1, 0.0), (90, 0, 2, 2, 0.0)])
def test_ics22_product_matches_reference_solver_live(V, n_cached, num_dev, clique, alpha):
    rng = np.random.default_rng(V + n_cached)
    indices = np.minimum((rng.pareto(1.1, V * 9) * 4).astype(np.int64), V - 1).astype(np.uint32)
    weight = np.bincount(indices, minlength=V).astype(np.float64)
    ranking, dev_map, idx_map, cached, total = O.ref_ics22_solve(weight, n_cached, num_dev, clique, alpha)
    r2 = O.ics22_placement(weight, n_cached, num_dev, clique, alpha)
    assert all(np.array_equal(a, b) for a, b in zip((ranking, dev_map, idx_map, cached), r2[:4])) and total == r2[4]
    rank_nodes, rank_w = O.ics22_in_degree_rank(indices, V)
    assert np.array_equal(rank_nodes, ranking)
    for w, (table, slots, tot) in enumerate(_product_ics22_views(rank_nodes, rank_w, V, n_cached, num_dev, clique, alpha)):
        assert tot == total
        assert np.array_equal(table, O.ics22_compact_table(ranking, dev_map, idx_map, total, w))
        assert np.array_equal(slots, cached[w])


@pytest.mark.parametrize("case,V,n,epochs", [(0, 500, 6, 1), (1, 1000, 8, 2), (2, 64, 3, 3)])
def test_presample_rank_oracle_matches_reference_golden(case, V, n, epochs):
    G = _rank_cases()
    lists = G.presample_lists(V, n * epochs, 77 + case)
    nodes, freq = O.presample_rank(V, lists)
    assert np.array_equal(nodes, RANK_GOLD[f"presample_{case}_nodes"])
    assert np.array_equal(freq, RANK_GOLD[f"presample_{case}_freq"])


@needs_refcu
def test_presample_rank_oracle_matches_reference_live():
    G = _rank_cases()
    for seed, (V, n, epochs) in enumerate([(2000, 5, 2), (31, 4, 1), (4096, 12, 1)]):
        lists = G.presample_lists(V, n * epochs, 1000 + seed)
        want_nodes, want_freq = O.ref_presample_rank(V, lists, epochs)
        nodes, freq = O.presample_rank(V, lists)
        assert np.array_equal(nodes, want_nodes) and np.array_equal(freq, want_freq)


@needs_ref
@pytest.mark.parametrize("dtype,dim,bits", [(np.float32, 100, 6), (np.float16, 768, 3), (np.uint8, 7, 10), (np.int64, 1, 1)])
def test_mock_extract_oracle_matches_reference(dtype, dim, bits):
    """SAMGRAPH_EMPTY_FEAT: CPUMockExtract (cpu/cpu_extraction.cc:47-64, 92-118) masks the ROW INDEX to 2^bits rows"""
    rng = np.random.default_rng(bits)
    table = rng.integers(0, 100, size=(1 << bits, dim)).astype(dtype)
    idx = rng.integers(0, 1 << 20, size=777).astype(np.uint32)
    want = O.ref_mock_extract(table, idx, bits)
    got = O.mock_extract(table, idx, bits)
    assert got.tobytes() == want.tobytes()
    assert got.tobytes() == table[idx & ((1 << bits) - 1)].tobytes()
