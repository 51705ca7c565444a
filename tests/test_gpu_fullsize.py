"""BASELINE.json configs at (or near) their full sizes on one B200: size-independent properties.

The oracle cannot replay a 1.6 B-edge graph in seconds, so these check what is determined without it:
  * every feature row equals the generator's closed form for its node id (0 ULP, byte compare) - whichever
    tier (local HBM / pinned host) served it; labels likewise;
  * unique/remap: input_nodes duplicate-free, seeds are its prefix in order (prefix property),
    every layer's previous node list is a prefix of the next, all local ids in range;
  * sampling: per-node edge count == min(degree, fanout) for EVERY node of every layer, edges node-major in
    input order, and a sample of edges checked for membership in the CSR neighbour list.
Graphs come from the device-side generator (csrc/synth.cu); its numpy twin is pinned in tests/synth.py.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

FANOUT = [5, 10, 15]


def _lazy():
    from xgnn_b200 import dataset as D
    from xgnn_b200 import ops
    return D, ops


def _free_gb():
    free, _ = torch.cuda.mem_get_info()
    return free / 1e9


def _u32(t):  # int32 view -> int64 values of the uint32 bits
    return t.to(torch.int64) & 0xFFFFFFFF


def check_minibatch_properties(mb, st, D, seeds_np, fanout, uniform=True):
    from xgnn_b200.selfcheck import check_minibatch_properties as check  # shared with bench.py's per-rank self-check
    check(mb, st, D, seeds_np, fanout, uniform=uniform)


def _run(shape, scale, batches=3, cache_pct=1.0, host_rows=False, need_gb=0.0, flat=True, **sampler_kw):
    D, ops = _lazy()
    if _free_gb() < need_gb:
        pytest.skip(f"needs {need_gb:.0f} GB of free HBM")
    st = D.build_store(shape, 0, 0, 1, scale=scale, cache_pct=cache_pct, flat_when_full=flat)
    host = None
    if host_rows:  # cold rows served from pinned host memory over UVA (GPUExtractMissData path)
        host = torch.empty((st.num_node, st.dim), dtype=st.dtype).pin_memory()
        step = 1 << 21
        for o in range(0, st.num_node, step):
            n = min(step, st.num_node - o)
            host[o:o + n].copy_(D.gen_features(n, st.dim, st.dtype, torch.device("cuda", 0), begin=o))
        torch.cuda.synchronize()
    D.attach(st, host_feat=host)
    sm = ops.MinibatchSampler(st.graph, FANOUT, 8000, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                              label=st.label, seed=99, num_slots=2, device=0, **sampler_kw)
    rng = np.random.default_rng(5)
    for step in range(batches):
        seeds = rng.choice(st.train, size=8000 if step else 7999, replace=False).astype(np.uint32)
        mb = sm.sample(seeds, batch_key=step)
        check_minibatch_properties(mb, st, D, seeds, FANOUT, uniform=sampler_kw.get("sample_type", 0) in (0, 5, 7))
        mb.release()
    del sm, st
    torch.cuda.empty_cache()


def test_cfg2_papers100M_full_size():
    """configs[1]: 111 M nodes, 1.6 B edges, 128 x fp32 (57 GB of rows) in one B200."""
    _run("papers100M", 1, need_gb=90)


def test_cfg3_twitter_full_size():
    """configs[2]: 41.7 M nodes, 1.5 B edges, 256 x fp32 (42.6 GB of rows)."""
    _run("twitter", 1, need_gb=80, flat=False)  # through the direct-address cache table (rank-ordered part)


def test_cfg1_products_full_size_khop3():
    """configs[0] shape on the GPU path (2.4 M nodes, 62 M edges, 100 x fp32 = 400-byte rows: 25 chunks per row)."""
    _run("products", 1, sample_type=7)


def test_cfg5_mag240M_shape_hot_in_hbm_cold_in_pinned_host():
    """configs[4] scaled 1/16 (15.3 M nodes, 768 x fp16 = 1536-byte rows): 30 % of the rows cached in HBM, the
    rest read from pinned host memory over UVA inside the same gather launch."""
    _run("mag240M-homo", 16, batches=2, cache_pct=0.3, host_rows=True, need_gb=40)


def test_cfg4_friendster_shape_random_walk():
    """configs[3] scaled 1/2 (32.8 M nodes, 1.8 B edges): PinSAGE random-walk top-K through the engine; the walk's
    structure (grouped by seed, <= K neighbours, counts descending) is checked here, its law in test_samplers_ext."""
    D, ops = _lazy()
    if _free_gb() < 60:
        pytest.skip("needs 60 GB of free HBM")
    st = D.attach(D.build_store("friendster", 0, 0, 1, scale=2))
    K = 5
    sm = ops.MinibatchSampler(st.graph, [K, K, K], 8000, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                              label=st.label, seed=7, num_slots=2, device=0, sample_type=3, random_walk_length=3,
                              num_random_walk=4, random_walk_restart_prob=0.5)
    seeds = st.train[:8000].astype(np.uint32)
    mb = sm.sample(seeds, batch_key=0)
    inp = _u32(mb.input_nodes)
    assert torch.unique(inp).numel() == inp.numel()
    assert torch.equal(inp[:8000], torch.from_numpy(seeds.astype(np.int64)).cuda())
    for l in range(3):
        col, cnt = _u32(mb.col(l)), _u32(mb.data(l))
        assert bool((col[1:] >= col[:-1]).all())  # grouped by seed in input order
        per_seed = torch.bincount(col, minlength=mb.num_dst[l])
        assert int(per_seed.max()) <= K
        same = col[1:] == col[:-1]
        assert bool((cnt[1:][same] <= cnt[:-1][same]).all())  # visit counts descending inside a seed
        assert int(cnt.min()) >= 1 and int(cnt.max()) <= 4 * 3
    want = D.gen_features(mb.num_input, st.dim, st.dtype, inp.device, row_ids=mb.input_nodes)
    assert torch.equal(mb.feat.view(torch.int32), want.view(torch.int32))
    mb.release()
    del sm, st
    torch.cuda.empty_cache()
