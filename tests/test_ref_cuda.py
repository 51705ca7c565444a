"""The repo's sampler kernels against the REFERENCE's own CUDA samplers, both running on this GPU.

oracle/_ref/libxgnn_ref_cuda.so = the reference's cuda_sampling_khop{0,1,2,3}.cu, cuda_sampling_weighted_khop*.cu,
cuda_sampling_random_walk.cu, cuda_frequency_hashmap.cu, cuda_random_states.cu, cuda_cache_manager_*.cu, cuda_hashtable.cu
and the device accessors of dist_graph.h / ics22_song_dist_feature.h compiled in place for sm_100
(oracle/Makefile `refcuda`, oracle/ref_cuda_shim.cu).  These are the only files that define KHop1 / weighted /
random-walk behaviour (the CPU twins are empty stubs), so this is where SURVEY 8a rows a5, a6, a7 are pinned:

  * the reference seeds thread t with curand_init(wall_clock + t, 0, 0) (cuda_random_states.cu:36-46,106-108):
    XORWOW generators with consecutive seeds are strongly correlated over their first outputs (measured below: a
    degree-16 node NEVER repeats a neighbour in 6 draws, where independent draws do 28 % of the time), and decorrelate
    as the states advance.  The reference keeps its states for the life of the engine and stores them back after
    every kernel, so its steady state is what the repo is compared with: WARMUP discarded calls first.
    test_reference_cold_start_correlation keeps the cold-start measurement;
  * exact equality wherever the reference's result does not depend on its curand draws:
    degree <= fanout copies, degree-1 / degree-0 nodes, walks on chains / cycles with restart 0;
  * a two-sample chi-square (homogeneity, p > 1e-5) between the reference kernel's and the repo kernel's output
    frequencies everywhere else: per (degree class, neighbour position) draw counts, edges-per-node histograms (the
    adjacent-duplicate law of KHop1 / weighted), per (seed, node, visit count) for walks, sorted top-K counts;
  * KHop0 against 20 k runs of the reference's CPU reservoir (cpu/cpu_sampling_khop0.cc:56-68): its inclusive
    RandomID(0, j + 1) makes it measurably non-uniform; the test fits the reference to that derived law, the repo
    kernel to the uniform law north_star asks for, and prints the gap.
"""
import numpy as np
import pytest

import oracle_lib as O
import synth
from stats import one_sample_chi2, two_sample_chi2

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(O.refcu is None, reason="oracle/_ref/libxgnn_ref_cuda.so not built")]

P_MIN = 1e-5
WARMUP = 48  # discarded reference calls per (sampler, fanout, batch) before its output is used (see the module docstring)


def statistical(attempts=3):
    """The reference seeds its curand states from the wall clock, so its half of a two-sample test is a fresh random
    sample on every run, and a correct pair of samplers can still fail a p > 1e-5 bar once in a while.  (The one failure seen
    so far, p = 1.3e-7 and later p = 0 on a walk test, had another cause: see _warm_set.)  A test marked @statistical() is repeated with NEW draws when it fails and passes if any attempt
    does: a real difference in law fails every attempt (its p-values are ~0), a fluke does not repeat.  The round-end run
    uses pytest -x, where a fluke would hide every test behind it."""
    import functools

    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*a, **kw):
            for k in range(attempts):
                try:
                    return fn(*a, **kw)
                except AssertionError as e:
                    if k == attempts - 1:
                        raise
                    print(f"[statistical] attempt {k + 1} of {fn.__name__} failed ({str(e)[:120]}); drawing again")
        return wrapper
    return deco


def _warm_set(rg):
    """per OBJECT, not per id(): a test-local RefCudaGraph can be freed and its address handed to the next one, whose fresh
    (cold, correlated) curand states would then be taken for warmed ones - seen as p = 0 on one walk test in one run"""
    if not hasattr(rg, "_warmed_keys"):
        rg._warmed_keys = set()
    return rg._warmed_keys


def _ref_sample(rg, st, inp, fanout, **kw):
    key = (st, fanout, len(inp))
    warmed = _warm_set(rg)
    if key not in warmed:
        for _ in range(WARMUP):
            rg.sample(st, inp, fanout, **kw)
        warmed.add(key)
    return rg.sample(st, inp, fanout, **kw)

DEGS = [0, 1, 2, 3, 7, 16, 40]
M = 500  # replicas per degree class
T0 = len(DEGS) * M  # first target node
V = T0 + max(DEGS)


def _dev(a):
    import torch
    a = np.ascontiguousarray(a)
    return torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a).cuda()


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.fixture(scope="module")
def gadget():
    """DEGS degree classes x M identical replicas; a hub of class t lists the target nodes T0 .. T0 + DEGS[t] - 1 with
    position-dependent weights, so (class, position) pools every replica's draws into one category."""
    deg = np.concatenate([np.repeat(DEGS, M), np.zeros(max(DEGS), np.int64)])
    indptr = np.zeros(V + 1, np.uint32)
    np.cumsum(deg, out=indptr[1:])
    indices = np.concatenate([T0 + np.arange(d) for d in np.repeat(DEGS, M)] + [np.zeros(0, np.int64)]).astype(np.uint32)
    pos = indices.astype(np.int64) - T0
    w = 0.2 + ((pos * 7) % 5).astype(np.float64) + (pos % 3 == 0) * 3.0
    prob, alias = synth.alias_tables(indptr, indices, w)
    prefix = synth.prefix_table(indptr, w)
    from xgnn_b200 import ops
    g = ops.make_graph(_dev(indptr), _dev(indices))
    rg = O.RefCudaGraph(indptr, indices, prob, alias, prefix)
    rng = np.random.default_rng(3)
    hubs = rng.permutation(T0).astype(np.uint32)
    return dict(ops=ops, g=g, rg=rg, indptr=indptr, indices=indices, prob=prob, alias=alias, prefix=prefix, hubs=hubs, w=w)


def _cls(src):
    return src.astype(np.int64) // M


def _draw_table(src, dst):
    """counts per (degree class, neighbour position)"""
    t = np.zeros((len(DEGS), max(DEGS)), np.int64)
    np.add.at(t, (_cls(src), dst.astype(np.int64) - T0), 1)
    return t


def _edges_per_node_hist(src, fanout):
    """per degree class: histogram of the number of output edges per input node"""
    per_node = np.bincount(src.astype(np.int64), minlength=T0)[:T0]
    h = np.zeros((len(DEGS), fanout + 1), np.int64)
    np.add.at(h, (np.arange(T0) // M, per_node), 1)
    return h


def _repo_sample(gd, st, inp, fanout, key):
    ops, g = gd["ops"], gd["g"]
    if st == 1:
        s, d = ops.GPUSampleKHop1(g, _dev(inp), fanout, seed=77, batch_key=key)
    elif st in (2, 6):
        s, d = ops.GPUSampleWeightedKHop(g, _dev(inp), fanout, sample_type=st, prob_table=_dev(gd["prob"]),
                                         alias_table=_dev(gd["alias"]), seed=77, batch_key=key)
    elif st == 4:
        s, d = ops.GPUSampleWeightedKHop(g, _dev(inp), fanout, sample_type=4, prob_prefix_table=_dev(gd["prefix"]),
                                         seed=77, batch_key=key)
    else:
        s, d = ops.GPUSampleKHop(g, _dev(inp), fanout, seed=77, batch_key=key, sample_type=st)
    return _u32(s), _u32(d)


def _check_structure(gd, st, src, dst, inp, fanout):
    """what every legal output satisfies, reference and repo alike"""
    indptr, indices = gd["indptr"], gd["indices"]
    assert np.all(dst.astype(np.int64) - T0 < np.array(DEGS)[_cls(src)]), "a neighbour outside the node's list"
    assert not np.any(_cls(src) == 0), "degree-0 nodes emit nothing"
    per_node = np.bincount(src.astype(np.int64), minlength=T0)[:T0]
    deg = np.diff(indptr.astype(np.int64))[:T0]
    if st in (1, 2, 4):  # with replacement, grouped by ascending global src, adjacent duplicates removed
        assert np.all(np.diff(src.astype(np.int64)) >= 0)
        same = (src[1:] == src[:-1]) & (dst[1:] == dst[:-1])
        assert not same.any()
        assert np.all(per_node[deg == 1] == 1) and np.all(per_node[deg > 0] >= 1) and np.all(per_node <= fanout)
    else:  # without replacement, node-major in input order
        assert np.array_equal(per_node, np.minimum(deg, fanout))
        order = np.repeat(inp, np.minimum(deg[inp], fanout))
        assert np.array_equal(src, order)
        pairs = src.astype(np.int64) * (1 << 20) + dst
        assert len(np.unique(pairs)) == len(pairs), "a neighbour chosen twice"
        full = np.isin(src, np.nonzero(deg <= fanout)[0])  # degree <= fanout: the whole list in CSR order
        want = np.concatenate([indices[indptr[v]:indptr[v + 1]] for v in inp if deg[v] <= fanout] + [np.zeros(0, np.uint32)])
        assert np.array_equal(dst[full], want)


@pytest.mark.parametrize("st,fanout,reps", [(1, 6, 5), (2, 6, 5), (4, 6, 5), (1, 25, 2), (2, 3, 6), (4, 12, 3)])
@statistical()
def test_with_replacement_samplers_match_reference_kernels(gadget, st, fanout, reps):
    """KHop1 (cuda_sampling_khop1.cu:42-236), alias (weighted_khop.cu:41-76), prefix (weighted_khop_prefix.cu:41-92)"""
    inp = gadget["hubs"]
    tabs, hists = [], []
    for who in ("ref", "repo"):
        t, h = np.zeros((len(DEGS), max(DEGS)), np.int64), np.zeros((len(DEGS), fanout + 1), np.int64)
        for r in range(reps):
            if who == "ref":
                s, d, _ = _ref_sample(gadget["rg"], st, inp, fanout)
            else:
                s, d = _repo_sample(gadget, st, inp, fanout, key=r)
            _check_structure(gadget, st, s, d, inp, fanout)
            t += _draw_table(s, d)
            h += _edges_per_node_hist(s, fanout)
        tabs.append(t)
        hists.append(h)
    # degree-1 nodes: one edge each, to the only neighbour -> identical output restricted to them
    for c, dcls in enumerate(DEGS):
        if dcls < 2:
            assert np.array_equal(tabs[0][c], tabs[1][c]) and np.array_equal(hists[0][c], hists[1][c])
            continue
        stat, dof, p = two_sample_chi2(tabs[0][c, :dcls], tabs[1][c, :dcls])
        assert p > P_MIN, f"type {st} degree {dcls}: per-neighbour draw counts differ (chi2 {stat:.1f}, dof {dof}, p {p:.2e})"
        stat, dof, p = two_sample_chi2(hists[0][c], hists[1][c])
        assert p > P_MIN, f"type {st} degree {dcls}: edges-per-node histograms differ (chi2 {stat:.1f}, dof {dof}, p {p:.2e})"
        # and both follow `fanout` independent draws of the marginal law with adjacent repeats collapsed
        sim = _simulate_runs_hist(gadget, st, dcls, fanout, 400000)
        for who, hh in (("reference", hists[0][c]), ("repo", hists[1][c])):
            _, _, p = two_sample_chi2(hh, sim)
            assert p > P_MIN, f"type {st} degree {dcls}: {who} edges-per-node law is not that of independent draws (p {p:.2e})"


def _simulate_runs_hist(gd, st, d, fanout, n):
    """edges per node after adjacent-duplicate removal for `fanout` INDEPENDENT draws of the sampler's marginal law"""
    rng = np.random.default_rng(d * 1000 + fanout)
    if st == 1:
        p = np.full(d, 1.0 / d)
    else:
        cls = DEGS.index(d)
        b = int(gd["indptr"][cls * M])
        p = gd["w"][b:b + d] / gd["w"][b:b + d].sum()
    x = rng.choice(d, size=(n, fanout), p=p)
    runs = 1 + (x[:, 1:] != x[:, :-1]).sum(axis=1)
    return np.bincount(runs, minlength=fanout + 1)[:fanout + 1]


@pytest.mark.parametrize("st,fanout,reps", [(7, 5, 6), (7, 10, 4), (6, 5, 6), (6, 10, 4), (7, 39, 2), (0, 5, 1), (0, 16, 1), (5, 5, 6),
                                            (5, 10, 4)])
@statistical()
def test_without_replacement_samplers_match_reference_kernels(gadget, st, fanout, reps):
    """KHop3 (cuda_sampling_khop3.cu:78-146), KHop2 (cuda_sampling_khop2.cu), weighted hash-dedup
    (weighted_khop_hash_dedup.cu:59-116): exact for
    degree <= fanout, two-sample chi-square on per-edge inclusion otherwise.  KHop0's GPU kernel re-seeds every launch
    with curand_init(thread + num_input) and applies its lane-parallel replacements in racy order
    (cuda_sampling_khop0.cu:114-117): only its exact part is compared here; its law is covered by
    test_khop0_against_the_reference_cpu_reservoir."""
    inp = gadget["hubs"]
    # KHop2 (cuda_sampling_khop2.cu:46-375) shuffles the lists of the nodes it samples IN PLACE: it gets a graph copy of
    # its own, so that the order it leaves behind cannot reach the other samplers (or another fanout) through the fixture
    rg = O.RefCudaGraph(gadget["indptr"], gadget["indices"]) if st == 5 else gadget["rg"]
    tabs = []
    for who in ("ref", "repo"):
        t = np.zeros((len(DEGS), max(DEGS)), np.int64)
        for r in range(reps):
            if who == "ref":
                s, d, _ = _ref_sample(rg, st, inp, fanout)
            else:
                s, d = _repo_sample(gadget, st, inp, fanout, key=r)
            _check_structure(gadget, st, s, d, inp, fanout)
            t += _draw_table(s, d)
        tabs.append(t)
    for c, dcls in enumerate(DEGS):
        if dcls <= fanout:
            assert np.array_equal(tabs[0][c], tabs[1][c])
        elif st != 0:
            stat, dof, p = two_sample_chi2(tabs[0][c, :dcls], tabs[1][c, :dcls])
            assert p > P_MIN, f"type {st} degree {dcls}: inclusion counts differ (chi2 {stat:.1f}, dof {dof}, p {p:.2e})"


@statistical()
def test_reference_cold_start_correlation(gadget, capsys):
    """The measurement behind WARMUP: on freshly seeded states the reference's KHop1 draws of one node (consecutive
    threads = consecutive seeds) are far from independent; after WARMUP calls they are, and the repo always is."""
    inp, fanout, cls = gadget["hubs"], 6, DEGS.index(16)
    rg = gadget["rg"]
    cold = np.zeros(fanout + 1, np.int64)
    for _ in range(4):
        rg.reset_states()
        _warm_set(rg).clear()  # every cached state of this handle is gone: nothing is warm any more
        s, _, _ = rg.sample(1, inp, fanout)
        cold += _edges_per_node_hist(s, fanout)[cls]
    warm, repo = np.zeros(fanout + 1, np.int64), np.zeros(fanout + 1, np.int64)
    for _ in range(WARMUP):
        rg.sample(1, inp, fanout)
    _warm_set(rg).add((1, fanout, len(inp)))
    for r in range(4):
        warm += _edges_per_node_hist(rg.sample(1, inp, fanout)[0], fanout)[cls]
        repo += _edges_per_node_hist(_repo_sample(gadget, 1, inp, fanout, key=50 + r)[0], fanout)[cls]
    sim = _simulate_runs_hist(gadget, 1, 16, fanout, 400000)
    p_cold, p_warm, p_repo = (two_sample_chi2(x, sim)[2] for x in (cold, warm, repo))
    with capsys.disabled():
        print(f"\n[KHop1, degree 16, 6 draws, edges per node 0..6] reference cold {cold.tolist()} (p vs independent draws "
              f"{p_cold:.1e}); reference after {WARMUP} calls {warm.tolist()} (p {p_warm:.2f}); repo {repo.tolist()} (p {p_repo:.2f}); "
              f"independent draws x{sim.sum() // warm.sum()} {sim.tolist()}")
    assert p_cold < 1e-6 and p_warm > P_MIN and p_repo > P_MIN


@pytest.mark.parametrize("P,ratio", [(2, 1.0), (3, 0.6), (8, 0.35)])
def test_ggms_routing_equals_the_reference_device_dist_graph(P, ratio):
    """DeviceDistGraph::_GetRealPartId / NumEdge / operator[] (cuda/dist_graph.h:114-151) - the reference's own class, over
    partitions built by the oracle's partitioner - against the repo's route() over the same partitions: with fanout >=
    every degree both samplers return whole neighbour lists, so the outputs must be EQUAL arrays (and equal to the flat
    CSR's); node ids >= num_cache_node take the host slot in both."""
    from xgnn_b200 import ops
    nv, ne = 6000, 50000
    indptr, indices = synth.make_graph(nv, ne, seed=31, max_deg=30)
    fan = int(np.diff(indptr.astype(np.int64)).max())
    ncn = O.num_cache_node(indptr, ratio)
    parts = [O.partition_csr(indptr, indices, ncn, p, P) for p in range(P)]
    rg = O.RefCudaGraph(indptr, indices)
    inp = np.random.default_rng(P).permutation(nv)[:3000].astype(np.uint32)
    flat_s, flat_d, _ = rg.sample(7, inp, fan)
    rg.set_parts(parts, ncn)
    keep = [(_dev(pi), _dev(px if len(px) else np.zeros(1, np.uint32))) for pi, px in parts]
    gd = ops.make_graph(part_indptr=[k[0] for k in keep], part_indices=[k[1] for k in keep], num_cache_node=ncn,
                        num_node=nv, host_indptr=_dev(indptr), host_indices=_dev(indices))
    for st in (7, 0):
        rs, rd, _ = rg.sample(st, inp, fan)
        s, d = ops.GPUSampleKHop(gd, _dev(inp), fan, seed=1, batch_key=st, sample_type=st)
        assert np.array_equal(rs, flat_s) and np.array_equal(rd, flat_d), "the reference's own routing changed its result"
        assert np.array_equal(_u32(s), rs) and np.array_equal(_u32(d), rd)
    assert (inp >= ncn).any() == (ratio < 1.0)


@pytest.mark.parametrize("dtype,dim,pct", [(np.float32, 128, 0.3), (np.float16, 768, 0.6), (np.float32, 100, 1.0), (np.float32, 64, 0.0)])
def test_feature_cache_path_equals_the_reference_cache_manager(dtype, dim, pct):
    """The reference's own GPUCacheManager - its constructor builds the direct-address table and the cached rows
    (cuda_cache_manager_host.cc:62-131), GetMissCacheIndex splits a node list (cuda_cache_manager_device.cu:355-441),
    GPUExtractMissData + CombineCacheData fill the rows (:573-731) - against the repo's operators on the same inputs:
    the four index lists equal element for element, the rows equal byte for byte, in the three-call form and in the
    single fused launch."""
    import torch
    from xgnn_b200 import ops
    nv = 30000
    feat = synth.make_features(nv, dim, dtype)
    indptr, _ = synth.make_graph(nv, 200000, seed=8)
    rank = synth.degree_rank(indptr)
    rc = O.RefCudaCache(feat, rank, pct)
    # a minibatch's input nodes are unique (the reference CHECKs num_cache <= num_cached_nodes on that assumption)
    nodes = np.random.default_rng(dim).permutation(nv)[:20001].astype(np.uint32)
    want, ms, md, cs, cd = rc.extract(nodes)
    assert want.tobytes() == feat[nodes].tobytes(), "the reference's own path must return feat[nodes]"
    ncached = int(nv * pct)
    table = O.build_cache_table(rank, ncached, nv)
    cache_rows = feat[rank[:ncached]] if ncached else np.zeros((1, dim), dtype)
    h_feat = torch.from_numpy(feat).pin_memory()
    store = ops.make_feature_store(feat.strides[0], h_feat, _dev(table), [torch.from_numpy(cache_rows).cuda()], local_parts=(0,))
    gms, gmd, gcs, gcd = ops.GetMissCacheIndex(_dev(table), _dev(nodes))
    assert np.array_equal(_u32(gms), ms) and np.array_equal(_u32(gmd), md)
    assert np.array_equal(_u32(gcs), cs) and np.array_equal(_u32(gcd), cd)
    out = ops.ExtractGGMS(store, _dev(nodes), dim, h_feat.dtype)
    assert out.cpu().numpy().tobytes() == want.tobytes()
    out2 = torch.empty_like(out)
    ops.GPUExtractMissData(out2, h_feat, gms, gmd, feat.strides[0])
    ops.CombineCacheData(out2, store, gcs, gcd)
    assert out2.cpu().numpy().tobytes() == want.tobytes()


@pytest.mark.parametrize("P,pct,dim", [(1, 0.5, 16), (2, 0.3, 128), (3, 1.0, 25), (8, 0.6, 100)])
def test_partitioned_cache_routing_equals_the_reference_device_dist_feature(P, pct, dim):
    """where a cache rank lives.  The reference's own DeviceDistFeature::Get (cuda/dist_graph.h:193-207;
    what combine_cache_data_for_partition reads, cuda_cache_manager_device.cu:277-299) run on this GPU over parts laid out
    as DistGraph::FeatureLoad leaves them (dist_graph.cu:492-505), against the repo's CombineCacheData and fused gather over
    the same parts: byte for byte, and both equal feat[rank_node[rank]]."""
    import torch
    from xgnn_b200 import ops
    nv = 20011
    feat = synth.make_features(nv, dim)
    indptr, _ = synth.make_graph(nv, 150000, seed=6)
    rank = synth.degree_rank(indptr)
    ncached = int(nv * pct)
    parts = [O.partition_features(feat, rank, ncached, p, P) for p in range(P)]
    rng = np.random.default_rng(P)
    ranks = rng.permutation(ncached)[: min(ncached, 9001)].astype(np.uint32)
    want = O.ref_dist_feature_get(parts, ncached, ranks)
    assert want.tobytes() == feat[rank[ranks]].tobytes(), "the reference accessor over FeatureLoad's layout returns the ranked node's row"
    table = O.build_cache_table(rank, ncached, nv)
    h_feat = torch.from_numpy(feat).pin_memory()
    d_parts = [torch.from_numpy(p).cuda() if len(p) else torch.zeros((1, dim), device="cuda") for p in parts]
    d_table = _dev(table)  # the store keeps raw pointers: the tensors must outlive it
    store = ops.make_feature_store(dim * 4, h_feat, d_table, d_parts, local_parts=(0,))
    out = torch.full((len(ranks), dim), -1.0, device="cuda")
    dst = np.arange(len(ranks), dtype=np.uint32)
    ops.CombineCacheData(out, store, _dev(ranks), _dev(dst))
    assert out.cpu().numpy().tobytes() == want.tobytes()
    # by node id through the table (the fused launch): cached nodes resolve to the same (part, row) the reference reads
    nodes = rank[ranks]
    got = ops.ExtractGGMS(store, _dev(nodes), dim, torch.float32)
    assert got.cpu().numpy().tobytes() == want.tobytes()


@pytest.mark.parametrize("num_dev,clique,alpha,pct", [(4, 4, 0.0, 0.1), (4, 2, 0.5, 0.2), (8, 8, 0.2, 0.05)])
def test_ics22_compact_routing_equals_the_reference_device_accessor(num_dev, clique, alpha, pct):
    """DeviceICS22SongDistFeature::Get (cuda/ics22_song_dist_feature.h:27-47), the reference's own, in both of its forms -
    the (device_map, new_idx_map) pair and the compact `owner << (32 - bitwidth) | row` ids - over the placement the
    reference's own solver produced (ref_ics22_solve, compiled in place): the repo's gather through the compact table reads
    the same rows, for every worker's view."""
    import ctypes
    import torch
    from xgnn_b200 import ops
    from xgnn_b200._clib import check, lib
    V, E, dim, bw = 6000, 90000, 24, 3
    indptr, indices = synth.make_graph(V, E, seed=11, max_deg=400)
    feat = synth.make_features(V, dim)
    weight = np.bincount(indices, minlength=V).astype(np.float64)
    n_cached = int(V * pct)
    ranking, dev_map, idx_map, cached, total = O.ref_ics22_solve(weight, n_cached, num_dev, clique, alpha)
    parts_np = [feat[cached[d]] for d in range(num_dev)]  # ICS22SongDistGraph::FeatureLoad: CPUExtract of the slot list
    h_nodes, h_w = O.ics22_in_degree_rank(indices, V)
    d_parts = [torch.from_numpy(p).cuda() for p in parts_np]
    host = torch.from_numpy(feat).cuda()
    rng = np.random.default_rng(2)
    for w in range(num_dev):
        table, slots, tot = np.empty(V, np.uint32), np.empty(n_cached, np.uint32), ctypes.c_uint32(0)
        check(lib.xg_ics22_placement(h_nodes.ctypes.data, h_w.ctypes.data, V, n_cached, num_dev, clique, alpha, w, bw,
                                     table.ctypes.data, slots.ctypes.data, ctypes.addressof(tot)))
        reachable = np.nonzero(dev_map[w] != 0xFFFFFFFF)[0].astype(np.uint32)
        assert np.array_equal(np.nonzero(table != 0xFFFFFFFF)[0], reachable)
        nodes = rng.permutation(reachable)[:3000].astype(np.uint32)
        by_maps = O.ref_dist_feature_get(parts_np, V, nodes, device_map=dev_map[w], new_idx_map=idx_map[w])
        by_compact = O.ref_dist_feature_get(parts_np, V, table[nodes], compact_bitwidth=bw)
        assert by_maps.tobytes() == feat[nodes].tobytes() == by_compact.tobytes()
        d_table = _dev(table)
        st = ops.make_feature_store(dim * 4, host, d_table, parts=d_parts, local_parts=(w,), compact_bitwidth=bw)
        got = ops.ExtractGGMS(st, _dev(nodes), dim, torch.float32)
        assert got.cpu().numpy().tobytes() == by_compact.tobytes()
        out = torch.full((len(nodes), dim), -1.0, device="cuda")
        ops.CombineCacheData(out, st, _dev(table[nodes]), _dev(np.arange(len(nodes), dtype=np.uint32)))
        assert out.cpu().numpy().tobytes() == by_compact.tobytes()


def test_hashtable_against_the_reference_gpu_table():
    """cuda::OrderedHashTable + GPUMapEdges, the reference's own (cuda_hashtable.cu:699-1064, cuda_mapping.cu:50-80).
    Its winner among duplicate occurrences is whichever CAS lands first, so: EXACT equality of the unique list and of
    the remapped edges wherever a fill's input has no duplicates (FillWithUnique, and FillWithDuplicates of fresh
    ids: the reference's numbering is then the input order, i.e. ours); with duplicates: equal sets, equal counts,
    the reference's earlier list is a prefix of its later one exactly as ours is of ours, and both remaps invert to
    the same global ids."""
    from xgnn_b200 import ops
    rng = np.random.default_rng(11)
    space, cap = 200000, 60000
    seeds = rng.permutation(space)[:5000].astype(np.uint32)
    fresh = np.setdiff1d(rng.permutation(space)[:20000].astype(np.uint32), seeds)[:9000]
    rng.shuffle(fresh)
    dup = rng.integers(0, space, size=30000).astype(np.uint32)
    rt, t = O.RefCudaTable(cap), ops.OrderedHashTable(cap)
    t.Reset()
    # fill 1: unique seeds -> identical lists
    u_ref = rt.fill(seeds, unique_input=True)
    t.FillWithUnique(_dev(seeds))
    assert np.array_equal(u_ref, seeds) and np.array_equal(_u32(t.RefUnique()), seeds)
    # fill 2: duplicate-free fresh ids through the duplicate-tolerant path -> still identical, seeds stay the prefix
    u_ref = rt.fill(fresh)
    u_own = _u32(t.FillWithDuplicates(_dev(fresh)))
    assert np.array_equal(u_ref, u_own) and np.array_equal(u_own[:len(seeds)], seeds)
    rs, rd = rt.map_edges(seeds[:4000], fresh[:4000])
    s, d = ops.GPUMapEdges(_dev(seeds[:4000]), _dev(fresh[:4000]), t)
    assert np.array_equal(_u32(s), rs) and np.array_equal(_u32(d), rd)
    # fill 3: heavy duplication -> order inside the NEW part may differ (the reference is racy), everything else may not
    before = len(u_own)
    u_ref = rt.fill(dup)
    u_own = _u32(t.FillWithDuplicates(_dev(dup)))
    assert len(u_ref) == len(u_own) and np.array_equal(u_ref[:before], u_own[:before])
    assert np.array_equal(np.sort(u_ref[before:]), np.sort(u_own[before:]))
    assert len(np.unique(u_ref)) == len(u_ref)
    rs, rd = rt.map_edges(dup[::-1].copy(), dup)
    s, d = ops.GPUMapEdges(_dev(dup[::-1].copy()), _dev(dup), t)
    assert np.array_equal(u_ref[rd], dup) and np.array_equal(u_own[_u32(d)], dup)
    assert np.array_equal(u_ref[rs], dup[::-1]) and np.array_equal(u_own[_u32(s)], dup[::-1])
    # ours is the deterministic legal outcome: first occurrence order (CPUHashTable0)
    first = dup[np.sort(np.unique(dup, return_index=True)[1])]
    assert np.array_equal(u_own[before:], first[~np.isin(first, u_own[:before])])


# ---------------------------------------------------------------------------------------------- random walk + top-K
def _walk_sets(src, dst, cnt):
    return sorted(zip(src.tolist(), dst.tolist(), cnt.tolist()))


@pytest.mark.parametrize("name", ["chain", "cycle", "two_cycles"])
def test_random_walk_deterministic_graphs_equal_reference(name):
    """restart 0 on graphs where every node has one out-edge: the visit counts do not depend on the draws, so the
    (seed, node, count) sets must be EQUAL to the reference's (order inside equal counts is arbitrary in the reference:
    its sort payload is a hash-table position, cuda_frequency_hashmap.cu:500-532)."""
    from xgnn_b200 import ops
    if name == "chain":  # 0 -> 1 -> 2 -> 3 -> 4 (4 has no out-edge)
        indptr, indices = np.array([0, 1, 2, 3, 4, 4], np.uint32), np.array([1, 2, 3, 4], np.uint32)
    elif name == "cycle":  # 0 -> 1 -> 2 -> 0
        indptr, indices = np.array([0, 1, 2, 3], np.uint32), np.array([1, 2, 0], np.uint32)
    else:  # 0 <-> 1 and 2 -> 3 -> 4 -> 2, plus an isolated node 5
        indptr, indices = np.array([0, 1, 2, 3, 4, 5, 5], np.uint32), np.array([1, 0, 3, 4, 2], np.uint32)
    nv = len(indptr) - 1
    inp = np.arange(nv, dtype=np.uint32).repeat(3)
    g = ops.make_graph(_dev(indptr), _dev(indices))
    rg = O.RefCudaGraph(indptr, indices)
    for W, L, K in [(4, 3, 5), (3, 7, 2), (8, 5, 40)]:
        rs, rd, rc = rg.sample(3, inp, K, walk_length=L, restart_prob=0.0, num_walk=W)
        s, d, c = ops.GPUSampleRandomWalk(g, _dev(inp), L, 0.0, W, K, seed=5, batch_key=1)
        s, d, c = _u32(s), _u32(d), _u32(c)
        assert np.array_equal(s, rs), "grouped by seed in input order"
        assert np.array_equal(c, rc), "count-descending inside a seed"
        if K >= L:  # nothing truncated: same sets
            assert _walk_sets(s, d, c) == _walk_sets(rs, rd, rc)


@pytest.mark.parametrize("W,L,K,p", [(4, 3, 12, 0.5), (4, 3, 5, 0.5), (10, 4, 3, 0.2), (6, 2, 12, 0.0)])
@statistical()
def test_random_walk_topk_matches_reference_kernels(W, L, K, p):
    """cuda_sampling_random_walk.cu:44-165 + FrequencyHashmap::GetTopK (cuda_frequency_hashmap.cu:643-841): 8 seeds
    x 400 copies (each input position is its own seed slot in both implementations).  Untruncated (K >= W * L):
    two-sample chi-square over (seed, node, count) categories.  Always: over the (seed, count) histogram of the
    emitted top-K and over the number of edges per seed, neither of which depends on how ties are ordered."""
    from xgnn_b200 import ops
    nv, ne = 400, 2400
    indptr, indices = synth.make_graph(nv, ne, seed=5, max_deg=12)
    deg = np.diff(indptr.astype(np.int64))
    seeds = np.nonzero(deg >= 2)[0][:8].astype(np.uint32)
    copies = 400
    inp = np.repeat(seeds, copies)
    g = ops.make_graph(_dev(indptr), _dev(indices))
    rg = O.RefCudaGraph(indptr, indices)
    maxc = W * L + 1

    res = {}
    for who in ("ref", "repo"):
        acc_node = np.zeros((len(seeds), nv, maxc), np.int64)
        acc_total = np.zeros((len(seeds), maxc), np.int64)
        n_edges = np.zeros(len(seeds), np.int64)
        for r in range(3):
            if who == "ref":
                s, d, c = _ref_sample(rg, 3, inp, K, walk_length=L, restart_prob=p, num_walk=W)
            else:
                s, d, c = (_u32(x) for x in ops.GPUSampleRandomWalk(g, _dev(inp), L, p, W, K, seed=31, batch_key=r))
            assert np.all(np.diff(np.searchsorted(seeds, s)) >= 0), "grouped by seed in input order"
            assert np.all(c >= 1) and np.all(c <= W * L)
            sid = np.searchsorted(seeds, s)
            np.add.at(acc_node, (sid, d.astype(np.int64), c.astype(np.int64)), 1)
            np.add.at(acc_total, (sid, c.astype(np.int64)), 1)
            np.add.at(n_edges, sid, 1)
        res[who] = (acc_node, acc_total, n_edges)
    untruncated = K >= W * L
    for i in range(len(seeds)):
        if untruncated:
            a, b = res["ref"][0][i].ravel(), res["repo"][0][i].ravel()
            stat, dof, pv = two_sample_chi2(a, b)
            assert pv > P_MIN, f"seed {seeds[i]}: (node, count) frequencies differ (chi2 {stat:.1f}, dof {dof}, p {pv:.2e})"
        a, b = res["ref"][1][i], res["repo"][1][i]
        stat, dof, pv = two_sample_chi2(a, b)
        assert pv > P_MIN, f"seed {seeds[i]}: top-K count histograms differ (chi2 {stat:.1f}, dof {dof}, p {pv:.2e})"
    # number of output edges per seed: a two-sample test on the totals
    stat, dof, pv = two_sample_chi2(res["ref"][2], res["repo"][2])
    assert pv > P_MIN, f"edges per seed differ (chi2 {stat:.1f}, p {pv:.2e})"


# ---------------------------------------------------------------------------------------------- KHop0 vs the CPU reservoir
@pytest.mark.skipif(O.ref is None, reason="oracle/_ref/libxgnn_ref.so not built")
@pytest.mark.parametrize("cls,fanout", [(5, 5), (6, 10), (4, 3)])
def test_khop0_against_the_reference_cpu_reservoir(gadget, cls, fanout, capsys):
    """north_star: 'per-edge selection-frequency chi-square against the reference's own sampler'.  20 000 runs of
    CPUSampleKHop0 (one thread: its generators are default-seeded per thread, cpu_random.cc:26-30) against 20 000
    draws of the repo kernel.  The reference draws RandomID(0, j + 1) INCLUSIVE (cpu_sampling_khop0.cc:63) where
    Algorithm R needs [0, j]: item j >= f replaces with probability f / (j + 2), which works out to inclusion
    (f + 1) / (d + 1) for the first f neighbours and f / (d + 1) for the others instead of f / d for all.  The test
    fits the reference to that derived law, the repo kernel to the uniform law, and prints the measured gap."""
    d = DEGS[cls]
    nodes = np.arange(cls * M, (cls + 1) * M, dtype=np.uint32)
    runs = 40  # x M = 20 000 samples of a degree-d node
    O.ref.ref_set_threads(1)
    ref_t, repo_t = np.zeros(d, np.int64), np.zeros(d, np.int64)
    for r in range(runs):
        s, dd = O.ref_sample_khop0(gadget["indptr"], gadget["indices"], nodes, fanout)
        ref_t += np.bincount(dd.astype(np.int64) - T0, minlength=d)[:d]
        s, dd = _repo_sample(gadget, 0, nodes, fanout, key=1000 + r)
        repo_t += np.bincount(dd.astype(np.int64) - T0, minlength=d)[:d]
    n = runs * M
    assert ref_t.sum() == n * fanout and repo_t.sum() == n * fanout
    law_ref = np.r_[np.full(fanout, (fanout + 1) / (d + 1)), np.full(d - fanout, fanout / (d + 1))]
    law_uni = np.full(d, fanout / d)
    _, _, p_ref_own = one_sample_chi2(ref_t, n * law_ref)
    _, _, p_ref_uni = one_sample_chi2(ref_t, n * law_uni)
    _, _, p_repo_uni = one_sample_chi2(repo_t, n * law_uni)
    _, _, p_two = two_sample_chi2(ref_t, repo_t)
    gap = np.abs(ref_t / n - repo_t / n)
    with capsys.disabled():
        print(f"\n[khop0 d={d} f={fanout}] reference CPU reservoir: first-{fanout} inclusion {ref_t[:fanout].mean() / n:.4f} "
              f"(derived {(fanout + 1) / (d + 1):.4f}), rest {ref_t[fanout:].mean() / n:.4f} (derived {fanout / (d + 1):.4f}); "
              f"repo kernel {repo_t.mean() / n:.4f} everywhere (uniform {fanout / d:.4f}); max |gap| {gap.max():.4f}; "
              f"p(ref ~ its derived law) {p_ref_own:.3f}, p(ref ~ uniform) {p_ref_uni:.1e}, p(repo ~ uniform) {p_repo_uni:.3f}, "
              f"two-sample p {p_two:.1e}")
    assert p_ref_own > P_MIN, "the reference no longer follows the law derived from its inclusive RandomID"
    assert p_repo_uni > P_MIN, "the repo kernel is not uniform without replacement"
    assert p_ref_uni < 1e-6, "expected: the reference's reservoir is measurably non-uniform"
    # the gap between the two samplers is the derived one, not something else
    want_gap = np.abs(law_ref - law_uni)
    assert np.all(np.abs(gap - want_gap) < 5 * np.sqrt(0.25 / n) * 2)
