"""GPU parity tests of the single operators, through the C-ABI (xgnn_b200.ops -> ctypes -> libxgnn_b200.so),
against the CPU oracle (oracle/oracle.c) and the committed reference fixtures (tests/golden/)."""
from pathlib import Path

import numpy as np
import pytest
import torch

import oracle_lib as O
import synth

pytestmark = pytest.mark.gpu
GOLD = np.load(Path(__file__).parent / "golden" / "ref_path_small.npz")


def dev(a):
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(a).cuda()


def host(t):
    a = t.detach().cpu().numpy()
    return a.view(np.uint32) if a.dtype == np.int32 else a


@pytest.fixture(scope="module")
def ops():
    from xgnn_b200 import ops as _ops
    return _ops


# ---- hash table ------------------------------------------------------------------------------
def test_hashtable_golden(ops):
    t = ops.OrderedHashTable(4000)
    u1 = t.FillWithDuplicates(dev(GOLD["fill1"]))
    assert np.array_equal(host(u1), GOLD["uniq1"])
    u2 = t.FillWithDuplicates(dev(GOLD["fill2"]))
    assert np.array_equal(host(u2), GOLD["uniq2"])
    ns, nd = ops.GPUMapEdges(dev(GOLD["fill2"]), dev(GOLD["fill2"][::-1].copy()), t)
    assert np.array_equal(host(ns), GOLD["map_src"])
    assert np.array_equal(host(nd), GOLD["map_dst"])


@pytest.mark.parametrize("fill", ["FillWithDupRevised", "FillWithDuplicates", "FillWithDupMutable"])
@pytest.mark.parametrize("reuse", [False, True])
def test_hashtable_reference_gtest_vectors(ops, fill, reuse):
    """samgraph/unittest/test_hashmap.cc:84-276: set equality, prefix, no duplicates, SearchO2N bijection."""
    t = ops.OrderedHashTable(4096)
    if reuse:
        t.FillWithDupRevised(dev(np.arange(100, 900, dtype=np.uint32)))
        t.Reset()
    seen = []
    prev = np.empty(0, np.uint32)
    orc = O.Table(4096)
    for v in ([1, 2, 3, 4, 5, 2, 5, 1, 10, 233], [1, 2, 3, 6, 9, 2, 8, 1, 3, 7, 1023],
              [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 1, 2, 3, 4, 5], list(range(1, 513))):
        a = np.array(v, np.uint32)
        seen.append(a)
        getattr(t, fill)(dev(a)) if fill != "FillWithDuplicates" else t.FillWithDuplicates(dev(a))
        orc.populate(a)
        u = host(t.RefUnique())
        assert t.NumItems() == len(u)
        assert set(u.tolist()) == set(np.concatenate(seen).tolist())
        assert len(set(u.tolist())) == len(u)
        assert np.array_equal(u[: len(prev)], prev)
        assert np.array_equal(host(t.SearchO2N(dev(u))), np.arange(len(u), dtype=np.uint32))
        assert np.array_equal(u, orc.unique())  # stronger than the reference: first-occurrence order
        prev = u


@pytest.mark.parametrize("n,space", [(1, 10), (255, 50), (1025, 100000), (1_000_000, 300_000), (1_000_000, 1 << 31)])
def test_hashtable_random_vs_oracle(ops, n, space):
    rng = np.random.default_rng(n + space)
    t = ops.OrderedHashTable(3 * n + 16)
    orc = O.Table(3 * n + 16)
    for rep in range(3):  # 3 x 1M random ids as in test_hashmap.cc:223-273
        a = rng.integers(0, space, size=n).astype(np.uint32)
        u = t.FillWithDuplicates(dev(a))
        orc.populate(a)
        assert np.array_equal(host(u), orc.unique())
    src = orc.unique()[rng.integers(0, orc.num_items(), size=min(n, 200000))]
    dst = orc.unique()[rng.integers(0, orc.num_items(), size=min(n, 200000))]
    ns, nd = ops.GPUMapEdges(dev(src), dev(dst), t)
    os_, od = orc.map_edges(src, dst)
    assert np.array_equal(host(ns), os_) and np.array_equal(host(nd), od)
    # absent ids are reported, not spun on
    absent = np.array([space + 5, space + 77], np.uint32) if space < (1 << 31) else np.array([0xFFFFFFF0], np.uint32)
    if not np.isin(absent, orc.unique()).any():
        assert np.all(host(t.SearchO2N(dev(absent))) == O.EMPTY)


def test_hashtable_fill_with_unique_and_empty(ops):
    t = ops.OrderedHashTable(1000)
    a = np.random.default_rng(0).permutation(500).astype(np.uint32)
    t.FillWithUnique(dev(a))
    assert t.NumItems() == 500
    assert np.array_equal(host(t.RefUnique()), a)
    u = t.FillWithDuplicates(dev(np.empty(0, np.uint32)))
    assert np.array_equal(host(u), a)
    b = np.array([7, 900, 7, 901], np.uint32)
    u = t.FillWithDuplicates(dev(b))
    assert np.array_equal(host(u), np.concatenate([a, [900, 901]]).astype(np.uint32))


def test_hashtable_capacity_error(ops):
    from xgnn_b200._clib import XgError
    t = ops.OrderedHashTable(64)
    with pytest.raises(XgError):
        t.FillWithDuplicates(dev(np.arange(1000, dtype=np.uint32)))


def test_hashtable_many_fills_fold(ops):
    t = ops.OrderedHashTable(10000)
    orc = O.Table(10000)
    rng = np.random.default_rng(3)
    for _ in range(150):  # more fills than the device-side counter history holds
        a = rng.integers(0, 5000, size=40).astype(np.uint32)
        t.FillWithDupRevised(dev(a))
        orc.populate(a)
    assert np.array_equal(host(t.RefUnique()), orc.unique())


# ---- uniform sampler ---------------------------------------------------------------------------
def test_khop_full_neighbourhood_golden(ops):
    g = ops.make_graph(dev(GOLD["indptr"]), dev(GOLD["indices"]))
    for st in (ops.kKHop0, ops.kKHop2, ops.kKHop3):
        src, dst = ops.GPUSampleKHop(g, dev(GOLD["inp"]), int(GOLD["fanout"]), sample_type=st)
        assert np.array_equal(host(src), GOLD["khop0_src"])
        assert np.array_equal(host(dst), GOLD["khop0_dst"])


@pytest.mark.parametrize("fanout", [1, 3, 5, 10, 15, 25, 40, 100, 200])
def test_khop_structure_vs_oracle(ops, fanout):
    """deg <= fanout: bit-exact CSR order; deg > fanout: exactly `fanout` DISTINCT positions of the list."""
    indptr, indices = synth.make_graph(20000, 400000, seed=5, max_deg=300)
    # make neighbour ids unique per list position so that distinct positions are observable
    rng = np.random.default_rng(1)
    inp = rng.integers(0, 20000, size=7001).astype(np.uint32)
    g = ops.make_graph(dev(indptr), dev(indices))
    src, dst = ops.GPUSampleKHop(g, dev(inp), fanout, seed=9, batch_key=4)
    src, dst = host(src), host(dst)
    counts = np.minimum(np.diff(indptr.astype(np.int64))[inp], fanout)
    assert len(src) == counts.sum()
    assert np.array_equal(src, np.repeat(inp, counts))
    osrc, odst = O.sample_khop0(indptr, indices, inp, fanout)
    assert np.array_equal(src, osrc)
    off = np.concatenate([[0], np.cumsum(counts)])
    deg = np.diff(indptr.astype(np.int64))
    for i in range(0, len(inp), 13):
        v = inp[i]
        lst = indices[indptr[v]: indptr[v + 1]]
        got = dst[off[i]: off[i + 1]]
        if deg[v] <= fanout:
            assert np.array_equal(got, lst)
            assert np.array_equal(got, odst[off[i]: off[i + 1]])
        else:
            # multiset inclusion: got must be a sub-multiset of lst
            lu, lc = np.unique(lst, return_counts=True)
            gu, gc = np.unique(got, return_counts=True)
            assert np.isin(gu, lu).all()
            assert (gc <= lc[np.searchsorted(lu, gu)]).all()
    # reproducible for the same (seed, batch_key), different otherwise
    s2, d2 = ops.GPUSampleKHop(g, dev(inp), fanout, seed=9, batch_key=4)
    assert np.array_equal(host(d2), dst)
    if (deg[inp] > fanout).sum() > 50:
        s3, d3 = ops.GPUSampleKHop(g, dev(inp), fanout, seed=9, batch_key=5)
        assert not np.array_equal(host(d3), dst)


@pytest.mark.parametrize("deg,fanout", [(7, 3), (20, 5), (40, 15), (33, 32), (200, 25), (70, 40)])
def test_khop_selection_frequency_chi_square(ops, deg, fanout):
    """Per-edge inclusion frequency is uniform f/deg (north_star: chi-square when fanout < degree)."""
    reps = 20000
    indptr = np.array([0, deg], np.uint32)
    indices = np.arange(100, 100 + deg, dtype=np.uint32)
    g = ops.make_graph(dev(indptr), dev(indices))
    inp = np.zeros(reps, np.uint32)  # the same node `reps` times: every position has its own RNG counter
    src, dst = ops.GPUSampleKHop(g, dev(inp), fanout, seed=123, batch_key=1)
    dst = host(dst).reshape(reps, fanout) - 100
    # no repeats inside a sample
    assert (np.sort(dst, axis=1)[:, 1:] != np.sort(dst, axis=1)[:, :-1]).all()
    obs = np.bincount(dst.ravel(), minlength=deg).astype(np.float64)
    p = fanout / deg
    # each position is included with prob p independently across reps: z-score per edge, chi2 over edges
    var = reps * p * (1 - p)
    chi2 = ((obs - reps * p) ** 2 / var).sum()
    # inclusion indicators are negatively correlated with sum fixed -> chi2 ~ (deg-1) dof scaled by deg/(deg-1)
    from scipy import stats
    crit = stats.chi2.ppf(1 - 1e-6, deg - 1) * deg / (deg - 1)
    assert chi2 < crit, (chi2, crit)
    # pairwise: joint inclusion of (0,1) should be f(f-1)/(deg(deg-1))
    both = ((dst == 0).any(axis=1) & (dst == 1).any(axis=1)).mean()
    want = fanout * (fanout - 1) / (deg * (deg - 1))
    assert abs(both - want) < 6 * np.sqrt(max(want * (1 - want), 1e-9) / reps) + 1e-9


def test_khop_empty_and_isolated(ops):
    indptr = np.array([0, 0, 3, 3], np.uint32)
    indices = np.array([5, 6, 7], np.uint32)
    g = ops.make_graph(dev(indptr), dev(indices))
    src, dst = ops.GPUSampleKHop(g, dev(np.array([0, 2, 0], np.uint32)), 4)
    assert len(src) == 0
    src, dst = ops.GPUSampleKHop(g, dev(np.empty(0, np.uint32)), 4)
    assert len(src) == 0
    src, dst = ops.GPUSampleKHop(g, dev(np.array([0, 1, 2, 1], np.uint32)), 4)
    assert host(src).tolist() == [1, 1, 1, 1, 1, 1] and host(dst).tolist() == [5, 6, 7, 5, 6, 7]


@pytest.mark.parametrize("P", [1, 2, 3, 8])
@pytest.mark.parametrize("ratio", [0.0, 0.4, 1.0])
def test_khop_ggms_routed_equals_flat(ops, P, ratio):
    """DeviceDistGraph routing (cuda/dist_graph.h:141-151): partitioned store returns the flat graph's lists."""
    indptr, indices = synth.make_graph(6000, 60000, seed=8, max_deg=80)
    ncn = O.num_cache_node(indptr, ratio)
    parts = [O.partition_csr(indptr, indices, ncn, p, P) for p in range(P)]
    d_parts = [(dev(pi), dev(px) if len(px) else torch.zeros(1, dtype=torch.int32, device="cuda")) for pi, px in parts]
    # host slot: pinned whole graph, read through UVA
    h_indptr = torch.from_numpy(indptr.view(np.int32)).pin_memory()
    h_indices = torch.from_numpy(indices.view(np.int32)).pin_memory()
    g = ops.make_graph(part_indptr=[a for a, _ in d_parts], part_indices=[b for _, b in d_parts], num_cache_node=ncn,
                       num_node=6000, host_indptr=h_indptr, host_indices=h_indices)
    flat = ops.make_graph(dev(indptr), dev(indices))
    inp = np.random.default_rng(2).integers(0, 6000, size=3000).astype(np.uint32)
    for fanout in (5, 100):
        a = ops.GPUSampleKHop(g, dev(inp), fanout, seed=1, batch_key=2)
        b = ops.GPUSampleKHop(flat, dev(inp), fanout, seed=1, batch_key=2)
        assert np.array_equal(host(a[0]), host(b[0])) and np.array_equal(host(a[1]), host(b[1]))
    torch.cuda.synchronize()


# ---- extraction ------------------------------------------------------------------------------
def test_extract_golden(ops):
    V = len(GOLD["indptr"]) - 1
    idx = dev(GOLD["idx"])
    assert np.array_equal(host(ops.GPUExtract(dev(synth.make_features(V, 100, np.float32)), idx)), GOLD["x32"])
    x16 = ops.GPUExtract(dev(synth.make_features(V, 24, np.float16)), idx)
    assert np.array_equal(host(x16).view(np.uint16), GOLD["x16"].view(np.uint16))
    assert np.array_equal(host(ops.GPUExtract(dev(synth.make_labels(V)), idx)), GOLD["xl"])


@pytest.mark.parametrize("dtype,dim", [(np.float32, 100), (np.float32, 128), (np.float32, 256), (np.float16, 768),
                                       (np.float16, 3), (np.uint8, 7), (np.int64, 1), (np.float64, 5), (np.int32, 1),
                                       (np.float32, 1), (np.float32, 602)])
def test_extract_bit_exact_all_dtypes(ops, dtype, dim):
    rng = np.random.default_rng(dim)
    V = 50000
    tab = rng.integers(0, 255, size=(V, dim)).astype(dtype)
    idx = rng.integers(0, V, size=33333).astype(np.uint32)
    out = ops.GPUExtract(dev(tab), dev(idx))
    want = O.extract(tab, idx)
    assert np.array_equal(host(out).view(np.uint8), want.view(np.uint8))
    assert ops.GPUExtract(dev(tab), dev(np.empty(0, np.uint32))).shape[0] == 0


def test_extract_from_pinned_host_labels(ops):
    """DoGPULabelExtract: dim=1 i64 gather straight from host memory (dist/dist_loops.cc:938-974)."""
    lab = synth.make_labels(100000)
    h = torch.from_numpy(lab).pin_memory()
    idx = np.random.default_rng(0).integers(0, 100000, size=8000).astype(np.uint32)
    out = ops.GPUExtract(h.data_ptr(), dev(idx), dim=1, dtype=torch.int64)
    assert np.array_equal(host(out), lab[idx])


@pytest.mark.parametrize("P", [1, 2, 8])
@pytest.mark.parametrize("pct", [0.0, 0.3, 1.0])
def test_extract_ggms_vs_oracle(ops, P, pct):
    V, dim = 40000, 128
    feat = synth.make_features(V, dim)
    indptr, _ = synth.make_graph(V, 300000, seed=4)
    rank = synth.degree_rank(indptr)
    ncached = int(V * pct)
    table = O.build_cache_table(rank, ncached, V)
    parts = [O.partition_features(feat, rank, ncached, p, P) for p in range(P)]
    d_parts = [dev(p) if len(p) else torch.zeros((1, dim), device="cuda") for p in parts]
    h_feat = torch.from_numpy(feat).pin_memory()
    store = ops.make_feature_store(dim * 4, h_feat, dev(table), d_parts, local_parts=(0,))
    nodes = np.random.default_rng(7).integers(0, V, size=50001).astype(np.uint32)
    stats = torch.zeros(4, dtype=torch.int64, device="cuda")
    out = ops.ExtractGGMS(store, dev(nodes), dim, torch.float32, stats=stats)
    want = O.extract_ggms(feat, nodes, table, parts)
    assert np.array_equal(host(out).view(np.uint32), want.view(np.uint32))
    ms, md, cs, cd = O.miss_cache_index(table, nodes)
    st = stats.cpu().numpy()
    assert st[2] == len(ms) and st[0] + st[1] == len(cs)
    assert st[0] == (cs % P == 0).sum()
    # the reference's three-step form gives the same rows
    gms, gmd, gcs, gcd = ops.GetMissCacheIndex(dev(table), dev(nodes))
    assert np.array_equal(host(gms), ms) and np.array_equal(host(gmd), md)
    assert np.array_equal(host(gcs), cs) and np.array_equal(host(gcd), cd)
    out2 = torch.empty_like(out)
    ops.GPUExtractMissData(out2, h_feat, gms, gmd, dim * 4)
    ops.CombineCacheData(out2, store, gcs, gcd)
    assert torch.equal(out, out2)
    torch.cuda.synchronize()


@pytest.mark.parametrize("P", [1, 2, 8])
@pytest.mark.parametrize("pct,rep", [(1.0, 0.25), (0.6, 0.6), (0.5, 0.1), (1.0, 1.0)])
def test_extract_two_tier_store_vs_oracle(ops, P, pct, rep):
    """Hot ranks replicated, the rest partitioned, misses from pinned host: same rows as CPUExtract."""
    V, dim = 30000, 100
    feat = synth.make_features(V, dim)
    indptr, _ = synth.make_graph(V, 200000, seed=5)
    rank = synth.degree_rank(indptr)
    ncached, nrep = int(V * pct), int(V * rep)
    table = O.build_cache_table(rank, ncached, V)
    replica, parts = O.hybrid_placement(feat, rank, ncached, nrep, P)
    d_parts = [dev(p) if len(p) else torch.zeros((1, dim), device="cuda") for p in parts]
    d_rep = dev(replica) if nrep else None
    h_feat = torch.from_numpy(feat).pin_memory()
    store = ops.make_feature_store(dim * 4, h_feat, dev(table), d_parts, local_parts=(0,), replica=d_rep,
                                   num_replicated=nrep)
    nodes = np.random.default_rng(11).integers(0, V, size=20011).astype(np.uint32)
    stats = torch.zeros(4, dtype=torch.int64, device="cuda")
    out = ops.ExtractGGMS(store, dev(nodes), dim, torch.float32, stats=stats)   # register-staged kernel (stats)
    out_ring = ops.ExtractGGMS(store, dev(nodes), dim, torch.float32)            # shared-memory ring kernel
    want = O.extract_hybrid(feat, nodes, table, replica, parts)
    assert np.array_equal(want.view(np.uint32), O.extract(feat, nodes).view(np.uint32))  # placement never changes rows
    assert np.array_equal(host(out).view(np.uint32), want.view(np.uint32))
    assert np.array_equal(host(out_ring).view(np.uint32), want.view(np.uint32))
    r = table[nodes]
    hit = r != 0xFFFFFFFF
    local = hit & ((r < nrep) | (((r.astype(np.int64) - nrep) % P) == 0))
    st = stats.cpu().numpy()
    assert st[2] == (~hit).sum() and st[0] == local.sum() and st[1] == (hit & ~local).sum()


def test_sanity_check_counts_violations():
    """xg_sanity_check_ids / _edges count exactly the violations the reference's device asserts would trip on
    (cuda/cuda_sanity_check.cu:28-88): kEmptyKey, ids out of range, ids repeated inside an epoch, edge endpoints outside
    their block."""
    from xgnn_b200 import ops
    V = 1000
    rng = np.random.default_rng(3)
    ids = rng.permutation(V)[:600].astype(np.uint32)
    d = torch.from_numpy(ids.view(np.int32)).cuda()
    emap = torch.zeros(V, dtype=torch.int32, device="cuda")
    assert ops.sanity_check_ids(d, V) == 0
    assert ops.sanity_check_ids(d[:300], V, emap) == 0
    assert ops.sanity_check_ids(d[300:], V, emap) == 0  # same epoch, disjoint batch
    assert int(emap.sum()) == 600
    assert ops.sanity_check_ids(d[100:150], V, emap) == 50  # 50 seeds seen twice in one epoch
    bad = ids.copy()
    bad[7] = 0xFFFFFFFF
    bad[9] = V
    bad[11] = bad[12]
    db = torch.from_numpy(bad.view(np.int32)).cuda()
    assert ops.sanity_check_ids(db, V) == 2
    assert ops.sanity_check_ids(db, V, torch.zeros(V, dtype=torch.int32, device="cuda")) == 3
    assert ops.sanity_check_ids(d[:0], V) == 0
    row = torch.from_numpy(rng.integers(0, 50, 4000).astype(np.int32)).cuda()
    col = torch.from_numpy(rng.integers(0, 20, 4000).astype(np.int32)).cuda()
    assert ops.sanity_check_edges(row, col, 50, 20) == 0
    want = int(((row >= 40) | (col >= 15)).sum())
    assert ops.sanity_check_edges(row, col, 40, 15) == want and want > 0


@pytest.mark.parametrize("num_dev,clique,alpha,pct", [(4, 4, 0.0, 0.1), (4, 2, 0.5, 0.2), (8, 8, 0.2, 0.05), (2, 1, 0.0, 0.3)])
def test_extract_through_ics22_compact_map(num_dev, clique, alpha, pct):
    """ICS'22 clique placement end to end at the operator level: rank by in-degree on the GPU (xg_presample_accumulate
    over `indices` + xg_presample_rank) == the oracle's ranking; every worker's xg_ics22_placement view; one gather launch
    through the compact (owner << 29 | row) map == DeviceICS22SongDistFeature::Get for cached nodes and host rows for the
    rest (cuda/ics22_song_dist_feature.h:28-47), bit for bit.  All `GPUs` are buffers of this one device."""
    import ctypes
    from xgnn_b200 import ops
    from xgnn_b200._clib import check, lib
    V, E, dim = 6000, 90000, 24
    indptr, indices = synth.make_graph(V, E, seed=11, max_deg=400)
    feat = synth.make_features(V, dim)
    want_nodes, want_w = O.ics22_in_degree_rank(indices, V)
    d_idx = torch.from_numpy(indices.view(np.int32)).cuda()
    freq = torch.zeros(V, dtype=torch.int32, device="cuda")
    check(lib.xg_presample_accumulate(freq.data_ptr(), V, d_idx.data_ptr(), d_idx.numel(), None, None))
    rank_nodes = torch.empty(V, dtype=torch.int32, device="cuda")
    rank_w = torch.empty(V, dtype=torch.int32, device="cuda")
    check(lib.xg_presample_rank(freq.data_ptr(), V, rank_nodes.data_ptr(), rank_w.data_ptr(), None))
    h_nodes, h_w = rank_nodes.cpu().numpy().view(np.uint32), rank_w.cpu().numpy().view(np.uint32)
    assert np.array_equal(h_nodes, want_nodes) and np.array_equal(h_w, want_w)
    n_cached = int(V * pct)
    weight = np.bincount(indices, minlength=V).astype(np.float64)
    ranking, dev_map, idx_map, cached, total = O.ics22_placement(weight, n_cached, num_dev, clique, alpha)
    parts_np = [feat[cached[d]] for d in range(num_dev)]  # FeatureLoad: CPUExtract of the device's slot list
    d_parts = []
    for w in range(num_dev):
        table, slots, tot = np.empty(V, np.uint32), np.empty(n_cached, np.uint32), ctypes.c_uint32(0)
        check(lib.xg_ics22_placement(h_nodes.ctypes.data, h_w.ctypes.data, V, n_cached, num_dev, clique, alpha, w, 3,
                                     table.ctypes.data, slots.ctypes.data, ctypes.addressof(tot)))
        assert tot.value == total and np.array_equal(slots, cached[w])
        d_parts.append((torch.from_numpy(feat[slots]).cuda(), torch.from_numpy(table.view(np.int32)).cuda(), table))
    host = torch.from_numpy(feat).cuda()  # the "host" slot of the handle (device memory stands in for pinned host rows)
    rng = np.random.default_rng(1)
    nodes = rng.integers(0, V, 5000).astype(np.uint32)
    nodes[:50] = ranking[:50]  # the hottest nodes are in every batch
    d_nodes = torch.from_numpy(nodes.view(np.int32)).cuda()
    for w in range(num_dev):
        st = ops.make_feature_store(dim * 4, host, d_parts[w][1], parts=[p[0] for p in d_parts], local_parts=(w,),
                                    compact_bitwidth=3)
        got = ops.ExtractGGMS(st, d_nodes, dim, torch.float32).cpu().numpy()
        want = O.extract_ics22(parts_np, d_parts[w][2], feat, nodes)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        assert np.array_equal(got.view(np.uint32), feat[nodes].view(np.uint32))  # placement never changes WHAT is read
        stats = torch.zeros(4, dtype=torch.int64, device="cuda")
        ops.ExtractGGMS(st, d_nodes, dim, torch.float32, stats=stats)
        t = d_parts[w][2][nodes]
        n_host = int((t == 0xFFFFFFFF).sum())
        n_local = int(((t != 0xFFFFFFFF) & ((t >> 29) == w)).sum())
        assert stats.cpu().tolist()[:3] == [n_local, len(nodes) - n_host - n_local, n_host]
        # the reference's three-step form (GetMissCacheIndex -> GPUExtractMissData -> CombineCacheData with the compact
        # values as cache src, cuda_cache_manager_device.cu:355-441,573-731) gives the same rows
        gms, gmd, gcs, gcd = ops.GetMissCacheIndex(d_parts[w][1], d_nodes)
        out2 = torch.empty((len(nodes), dim), dtype=torch.float32, device="cuda")
        ops.GPUExtractMissData(out2, host, gms, gmd, dim * 4)
        ops.CombineCacheData(out2, st, gcs, gcd)
        assert np.array_equal(out2.cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("P,pct,rep", [(1, 0.5, 0.2), (4, 0.8, 0.3), (8, 1.0, 0.6), (2, 0.3, 0.3)])
def test_extract_two_tier_rank_select_words(P, pct, rep):
    """Replicated tier laid out by node id and resolved through the rank/select words (xg_build_replica_words) instead of
    the direct-address table: the words must reproduce the table for every replicated node, the gather must return
    CPUExtract's rows bit for bit, and a replica that is NOT in node-id order must be refused."""
    from xgnn_b200 import ops
    from xgnn_b200._clib import lib
    V, dim = 30011, 100
    feat = synth.make_features(V, dim)
    indptr, _ = synth.make_graph(V, 200000, seed=5)
    rank = synth.degree_rank(indptr).copy()
    ncached, nrep = int(V * pct), int(V * rep)
    policy_table = O.build_cache_table(rank, ncached, V)
    rank[:nrep] = np.sort(rank[:nrep])
    if pct == 1.0:
        rank[nrep:] = np.sort(rank[nrep:])  # everything cached, sharded tier by node id too: the words cover every node
    table = O.build_cache_table(rank, ncached, V)
    replica, parts = O.hybrid_placement(feat, rank, ncached, nrep, P)
    d_table = dev(table)
    words = ops.build_replica_words(d_table, nrep)
    assert words.cover_all == (pct == 1.0)
    w = words.cpu().numpy().view(np.uint32)
    member = np.zeros(((V + 31) // 32) * 32, bool)
    member[:V] = table < nrep
    bits = np.packbits(member.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1)
    assert np.array_equal(w[:, 0], bits)
    pc = member.reshape(-1, 32).sum(1)
    assert np.array_equal(w[:, 1], np.concatenate([[0], np.cumsum(pc)[:-1]]).astype(np.uint32))
    d_parts = [dev(p) if len(p) else torch.zeros((1, dim), device="cuda") for p in parts]
    h_feat = torch.from_numpy(feat).pin_memory()
    store = ops.make_feature_store(dim * 4, h_feat, d_table, d_parts, local_parts=(0,), replica=dev(replica),
                                   num_replicated=nrep, replica_words=words)
    nodes = np.random.default_rng(12).integers(0, V, size=20011).astype(np.uint32)
    stats = torch.zeros(4, dtype=torch.int64, device="cuda")
    out = ops.ExtractGGMS(store, dev(nodes), dim, torch.float32, stats=stats)
    out_ring = ops.ExtractGGMS(store, dev(nodes), dim, torch.float32)
    want = O.extract(feat, nodes)
    assert np.array_equal(host(out).view(np.uint32), want.view(np.uint32))
    assert np.array_equal(host(out_ring).view(np.uint32), want.view(np.uint32))
    assert np.array_equal(O.extract_hybrid(feat, nodes, table, replica, parts).view(np.uint32), want.view(np.uint32))
    r = table[nodes]
    st = stats.cpu().numpy()
    assert st[2] == (r == 0xFFFFFFFF).sum() and st[0] + st[1] == (r != 0xFFFFFFFF).sum()
    if words.cover_all:
        # the table is never read in this mode: a poisoned one (every node -> replica row 0) must not change a single byte
        poisoned = ops.make_feature_store(dim * 4, h_feat, torch.zeros_like(d_table), d_parts, local_parts=(0,),
                                          replica=dev(replica), num_replicated=nrep, replica_words=words)
        assert np.array_equal(host(ops.ExtractGGMS(poisoned, dev(nodes), dim, torch.float32)).view(np.uint32), want.view(np.uint32))
    if nrep > 1 and not np.array_equal(policy_table, table):
        bad = torch.empty_like(words)
        assert lib.xg_build_replica_words(dev(policy_table).data_ptr(), V, nrep, bad.data_ptr(), None, None) != 0


@pytest.mark.parametrize("dtype,dim,bits", [(np.float32, 128, 8), (np.float16, 768, 4), (np.float32, 100, 1), (np.uint8, 16, 12)])
def test_mock_extract_matches_oracle(ops, dtype, dim, bits):
    """GPUMockExtract (cuda/cuda_extraction.cu:51-72,118-160) == CPUMockExtract restatement, byte for byte; and through
    the GGMS store: host-resident rows are masked, cached rows are whatever the cache was built with"""
    import torch
    rng = np.random.default_rng(dim + bits)
    table = rng.integers(0, 100, size=(1 << bits, dim)).astype(dtype)
    idx = rng.integers(0, 1 << 22, size=5000).astype(np.uint32)
    d_table = torch.from_numpy(table).cuda()
    got = ops.GPUMockExtract(d_table, dev(idx), bits)
    assert got.cpu().numpy().tobytes() == O.mock_extract(table, idx, bits).tobytes()
    # a store whose host tier is the mock table: nothing cached -> every row is a masked host row
    store = ops.make_feature_store(table.strides[0], d_table, None, empty_feat=bits)
    out = ops.ExtractGGMS(store, dev(idx), dim, d_table.dtype)
    assert out.cpu().numpy().tobytes() == table[idx & ((1 << bits) - 1)].tobytes()
