"""Minibatch engine (pipeline.cu) against the CPU oracle's DoCPUSample restatement, through the C-ABI."""
import numpy as np
import pytest

import oracle_lib as O
import synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    a = np.ascontiguousarray(a)
    return torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a).cuda()


def u32(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.fixture(scope="module")
def ops():
    from xgnn_b200 import ops
    return ops


@pytest.fixture(scope="module")
def small():
    V, E, dim = 30000, 240000, 100
    indptr, indices = synth.make_graph(V, E, seed=7, max_deg=48)
    return dict(V=V, indptr=indptr, indices=indices, feat=synth.make_features(V, dim), dim=dim,
                label=synth.make_labels(V), maxdeg=int(np.diff(indptr.astype(np.int64)).max()))


def check_against_oracle(mb, ref, feat=None, label=None, seeds=None):
    L = mb.num_layer
    assert np.array_equal(u32(mb.input_nodes), ref["input_nodes"])
    for l in range(L):
        assert mb.num_src[l] == ref["num_src"][l] and mb.num_dst[l] == ref["num_dst"][l]
        assert mb.num_edge[l] == ref["num_edge"][l]
        assert np.array_equal(u32(mb.row(l)), ref["row"][l])
        assert np.array_equal(u32(mb.col(l)), ref["col"][l])
        ip = mb.csc_indptr(l)
        c = ref["col"][l]
        if len(c) and not np.all(c[1:] >= c[:-1]):
            assert ip is None  # duplicate seeds / grouped-by-global-src samplers: no destination-major form
        elif ip is not None:   # uniform samplers: the block's CSC offsets come with the batch
            assert np.array_equal(u32(ip), O.coo_to_csc_indptr(c, ref["num_dst"][l]))
    if feat is not None:
        got = mb.feat.cpu().numpy()
        assert got.view(np.uint32 if got.itemsize == 4 else np.uint16).tobytes() == feat[ref["input_nodes"]].tobytes()
    if label is not None:
        assert np.array_equal(mb.label.cpu().numpy(), label[seeds])


@pytest.mark.parametrize("use_graph", [True, False])
@pytest.mark.parametrize("L", [1, 2, 3])
def test_minibatch_bit_exact_when_fanout_covers_degree(ops, small, L, use_graph):
    g = ops.make_graph(dev(small["indptr"]), dev(small["indices"]))
    feat = dev(small["feat"])
    store = ops.make_feature_store(small["dim"] * 4, feat)
    fan = [small["maxdeg"]] * L
    B = 64 if L == 3 else 300
    sm = ops.MinibatchSampler(g, fan, B, store=store, feat_dim=small["dim"], label=dev(small["label"]),
                              seed=5, use_graph=use_graph, max_input_nodes=small["V"] + B)
    train = synth.train_set(small["V"], 0.2)
    for step in range(4):  # several batches through the same slots: Reset + graph replay
        n = B if step != 2 else B // 3  # a short batch in between
        seeds = train[step * B: step * B + n]
        mb = sm.sample(seeds, batch_key=step)
        ref = O.minibatch(small["indptr"], small["indices"], seeds, fan, seed=1)
        check_against_oracle(mb, ref, small["feat"], small["label"], seeds)
        mb.release()
    assert sm.launches_per_batch >= 2 * L + 5


def test_minibatch_structure_with_real_fanouts(ops, small):
    """fanout < degree: edges are a random subset; check everything that is still determined."""
    indptr, indices = small["indptr"], small["indices"]
    g = ops.make_graph(dev(indptr), dev(indices))
    store = ops.make_feature_store(small["dim"] * 4, dev(small["feat"]))
    fan = [5, 10, 15]
    B = 500
    sm = ops.MinibatchSampler(g, fan, B, store=store, feat_dim=small["dim"], label=dev(small["label"]), seed=11, num_slots=3)
    seeds = synth.train_set(small["V"], 0.2)[:B]
    mb = sm.sample(seeds, batch_key=3)
    inp = u32(mb.input_nodes)
    assert len(set(inp.tolist())) == len(inp)
    assert np.array_equal(inp[:B], seeds)  # prefix property (unique seeds here)
    deg = np.diff(indptr.astype(np.int64))
    cur = seeds
    for l in (2, 1, 0):
        row, col = u32(mb.row(l)), u32(mb.col(l))
        assert mb.num_dst[l] == len(cur)
        want = np.minimum(deg[cur], fan[l])
        assert mb.num_edge[l] == want.sum()
        assert np.array_equal(col, np.repeat(np.arange(len(cur), dtype=np.uint32), want))  # node-major
        # every edge exists in the graph, no neighbour position sampled twice when deg > fanout
        src_g, dst_g = inp[col], inp[row]
        for e in range(0, len(row), max(1, len(row) // 2000)):
            v = src_g[e]
            assert dst_g[e] in indices[indptr[v]:indptr[v + 1]]
        nxt = inp[: mb.num_src[l]]
        assert np.array_equal(nxt[: len(cur)], cur)
        assert set(dst_g.tolist()) <= set(nxt.tolist())
        cur = nxt
    assert np.array_equal(cur, inp)
    assert mb.feat.cpu().numpy().tobytes() == small["feat"][inp].tobytes()
    # same key -> same batch (counter-based RNG), different key -> different draws
    mb2 = sm.sample(seeds, batch_key=3)
    assert np.array_equal(u32(mb2.row(0)), u32(mb.row(0))) and np.array_equal(u32(mb2.input_nodes), inp)
    mb3 = sm.sample(seeds, batch_key=4)
    assert not np.array_equal(u32(mb3.input_nodes), inp)


def test_minibatch_duplicate_seeds_and_device_seeds(ops, small):
    """the aligned shuffler pads the train set with wrapped-around ids (dist_shuffler_aligned.cc:53-55)"""
    g = ops.make_graph(dev(small["indptr"]), dev(small["indices"]))
    fan = [small["maxdeg"], small["maxdeg"]]
    seeds = np.array([5, 9, 5, 17, 9, 23, 100, 5], np.uint32)
    sm = ops.MinibatchSampler(g, fan, 16, seed=1)
    ref = O.minibatch(small["indptr"], small["indices"], seeds, fan, seed=1)
    check_against_oracle(sm.sample(seeds), ref)
    check_against_oracle(sm.sample(dev(seeds)), ref)  # seeds already on the device
    mb = sm.sample(np.zeros(0, np.uint32))             # empty batch
    assert mb.num_input == 0 and mb.num_edge == [0, 0]


def test_minibatch_capacity_error_then_recovery(ops, small):
    g = ops.make_graph(dev(small["indptr"]), dev(small["indices"]))
    fan = [small["maxdeg"], small["maxdeg"]]
    seeds = synth.train_set(small["V"], 0.2)[:200]
    sm = ops.MinibatchSampler(g, fan, 200, seed=1, max_input_nodes=400)
    from xgnn_b200._clib import XgError
    with pytest.raises(XgError):
        sm.sample(seeds)
    few = seeds[:2]
    ref = O.minibatch(small["indptr"], small["indices"], few, fan, seed=1)
    if len(ref["input_nodes"]) <= 400:
        check_against_oracle(sm.sample(few), ref)


def test_slots_exhausted_is_reported(ops, small):
    g = ops.make_graph(dev(small["indptr"]), dev(small["indices"]))
    sm = ops.MinibatchSampler(g, [3], 8, num_slots=2)
    seeds = np.arange(8, dtype=np.uint32)
    a, b = sm.sample(seeds), sm.sample(seeds)
    from xgnn_b200._clib import XgError
    with pytest.raises(XgError):
        sm.sample(seeds)
    a.release()
    sm.sample(seeds)
    del b


# ---- the engine with the non-uniform samplers: replay the Philox words through the oracle, layer by layer ----
oracle_minibatch_ext = O.minibatch_ext


@pytest.fixture(scope="module")
def wgraph():
    V, E = 4000, 36000
    indptr, indices = synth.make_graph(V, E, seed=33, max_deg=40)
    w = synth.edge_weights(indptr, indices)
    prob, alias = synth.alias_tables(indptr, indices, w)
    return dict(V=V, indptr=indptr, indices=indices, prob=prob, alias=alias, prefix=synth.prefix_table(indptr, w),
                feat=synth.make_features(V, 16), label=synth.make_labels(V))


@pytest.mark.parametrize("use_graph", [True, False])
@pytest.mark.parametrize("sample_type", [1, 2, 4, 6, 3])
def test_minibatch_other_samplers_bit_exact_vs_oracle_replay(ops, wgraph, sample_type, use_graph):
    g = ops.make_graph(dev(wgraph["indptr"]), dev(wgraph["indices"]))
    store = ops.make_feature_store(16 * 4, dev(wgraph["feat"]))
    fan, B = ([5, 5] if sample_type == 3 else [4, 6]), 200
    kw = {}
    if sample_type in (2, 6):
        kw = dict(prob_table=dev(wgraph["prob"]), alias_table=dev(wgraph["alias"]))
    if sample_type == 4:
        kw = dict(prob_prefix_table=dev(wgraph["prefix"]))
    if sample_type == 3:
        kw = dict(random_walk_length=3, num_random_walk=4, random_walk_restart_prob=0.5)
    sm = ops.MinibatchSampler(g, fan, B, store=store, feat_dim=16, label=dev(wgraph["label"]), seed=77,
                              sample_type=sample_type, use_graph=use_graph, **kw)
    train = synth.train_set(wgraph["V"], 0.3)
    for step in range(3):
        seeds = train[step * B:(step + 1) * B].copy()
        if step == 1:
            seeds[3] = seeds[0]  # arch6 seeds may hold wrapped-around duplicates
        mb = sm.sample(seeds, batch_key=step)
        ref = oracle_minibatch_ext(wgraph, seeds, fan, sample_type, 77, step, rw=(3, 0.5, 4))
        assert np.array_equal(u32(mb.input_nodes), ref["input_nodes"])
        for l in range(len(fan)):
            assert np.array_equal(u32(mb.col(l)), ref["col"][l]) and np.array_equal(u32(mb.row(l)), ref["row"][l])
            assert mb.num_src[l] == ref["num_src"][l] and mb.num_dst[l] == ref["num_dst"][l]
            if sample_type == 3:
                assert np.array_equal(u32(mb.data(l)), ref["data"][l])
        assert mb.feat.cpu().numpy().tobytes() == wgraph["feat"][ref["input_nodes"]].tobytes()
        assert np.array_equal(mb.label.cpu().numpy(), wgraph["label"][seeds])
        mb.release()


def test_presample_rank_bit_exact_vs_oracle(ops, small):
    """kCacheByPreSample ranking (dist/pre_sampler.cc): with fanout >= degree the batches are deterministic, so the
    frequency table and the (freq desc, id desc) order must equal the oracle's bit for bit."""
    g = ops.make_graph(dev(small["indptr"]), dev(small["indices"]))
    fan = [small["maxdeg"], small["maxdeg"]]
    B = 200
    sm = ops.MinibatchSampler(g, fan, B, seed=5, max_input_nodes=small["V"] + B)
    train = synth.train_set(small["V"], 0.1)
    batches = [train[o:o + B] for o in range(0, len(train), B)]
    rank_nodes, rank_freq = ops.presample_rank(sm, small["V"], batches)
    lists = [O.minibatch(small["indptr"], small["indices"], b, fan, seed=1)["input_nodes"] for b in batches]
    want_nodes, want_freq = O.presample_rank(small["V"], lists)
    assert np.array_equal(u32(rank_freq), want_freq)
    assert np.array_equal(u32(rank_nodes), want_nodes)
    assert want_freq[0] >= want_freq[-1] and sorted(want_nodes.tolist()) == list(range(small["V"]))
    if O.refcu is not None:  # the reference's own PreSampler (dist/pre_sampler.cc:40-165, compiled in place) on the same lists
        ref_nodes, ref_freq = O.ref_presample_rank(small["V"], lists)
        assert np.array_equal(u32(rank_nodes), ref_nodes) and np.array_equal(u32(rank_freq), ref_freq)
