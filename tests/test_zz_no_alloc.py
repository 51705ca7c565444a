"""xg_get_miss_cache_index_noalloc: the GetMissCacheIndex split for callers that cannot call cudaMalloc (scan scratch in
a file-scope __device__ array, include/xgnn_b200.h).  No counterpart in the reference; checked against the existing
entry point xg_get_miss_cache_index (itself pinned against the oracle and the reference's GPUCacheManager elsewhere) -
bit-exact, the bar of every index test here - on calls IN A ROW: the second call of each pair finds the first one's
leftovers in the shared scratch (a longer call before a shorter one, and the other way round).
Sorted last on purpose: a failure here hides nothing after it.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _existing(lib, check, torch, table, nodes):
    n = nodes.numel()
    bufs = [torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda") for _ in range(4)]
    nm, nc = ctypes.c_size_t(0), ctypes.c_size_t(0)
    check(lib.xg_get_miss_cache_index(_ptr(table), _ptr(nodes), n, _ptr(bufs[0]), _ptr(bufs[1]), ctypes.byref(nm),
                                      _ptr(bufs[2]), _ptr(bufs[3]), ctypes.byref(nc), None))
    return [bufs[0][: nm.value], bufs[1][: nm.value], bufs[2][: nc.value], bufs[3][: nc.value]]


def _noalloc(lib, check, torch, table, nodes, bufs, counts):
    """every buffer preallocated by the caller; the call itself must not allocate"""
    n = nodes.numel()
    check(lib.xg_get_miss_cache_index_noalloc(_ptr(table), _ptr(nodes), n, _ptr(bufs[0]), _ptr(bufs[1]), _ptr(bufs[2]),
                                              _ptr(bufs[3]), _ptr(counts), None))


def test_noalloc_split_matches_existing_entry_point_on_calls_in_a_row():
    import torch
    from xgnn_b200._clib import check, lib
    torch.cuda.set_device(0)
    rng = np.random.default_rng(11)
    V = 300_000
    table = rng.permutation(V).astype(np.uint32)
    table[rng.random(V) < 0.37] = 0xFFFFFFFF  # 37 % of the nodes are not cached
    d_table = torch.from_numpy(table.view(np.int32)).cuda()
    sizes = [1_000_003, 4097, 1, 2_500_000, 1024, 1025]  # long before short and back: stale scan words beyond / below
    cap = max(sizes)
    bufs = [torch.empty(cap, dtype=torch.int32, device="cuda") for _ in range(4)]
    counts = torch.empty(2, dtype=torch.int64, device="cuda")
    for n in sizes:
        nodes = torch.from_numpy(rng.integers(0, V, size=n).astype(np.uint32).view(np.int32)).cuda()
        want = _existing(lib, check, torch, d_table, nodes)
        for call in range(2):  # two calls in a row on the same inputs, then the next size
            for b in bufs:
                b.fill_(-1)
            counts.fill_(-1)
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            _noalloc(lib, check, torch, d_table, nodes, bufs, counts)
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            nm, nc = counts.tolist()
            assert (nm, nc) == (want[0].numel(), want[2].numel()) and nm + nc == n, (n, call)
            got = [bufs[0][:nm], bufs[1][:nm], bufs[2][:nc], bufs[3][:nc]]
            for g, w in zip(got, want):
                assert torch.equal(g, w), (n, call)
            for b, used in zip(bufs, (nm, nm, nc, nc)):
                assert bool((b[used:n] == -1).all()), "wrote beyond its count"


def test_noalloc_split_empty_input_and_size_limit():
    import torch
    from xgnn_b200._clib import lib
    torch.cuda.set_device(0)
    d_table = torch.zeros(16, dtype=torch.int32, device="cuda")
    counts = torch.full((2,), -1, dtype=torch.int64, device="cuda")
    assert lib.xg_get_miss_cache_index_noalloc(_ptr(d_table), None, 0, None, None, None, None, _ptr(counts), None) == 0
    torch.cuda.synchronize()
    assert counts.tolist() == [0, 0]
    # one node beyond the stated maximum: refused with a clear error before anything is launched or touched
    too_many = (1 << 24) + 1
    one = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.xg_get_miss_cache_index_noalloc(_ptr(d_table), _ptr(one), too_many, _ptr(one), _ptr(one), _ptr(one), _ptr(one),
                                             _ptr(counts), None)
    assert rc != 0
    msg = lib.xg_last_error().decode()
    assert "XG_NOALLOC_MAX_NODES" in msg and str(too_many) in msg
    torch.cuda.synchronize()
    assert counts.tolist() == [0, 0]


def test_noalloc_split_at_the_stated_maximum():
    """num_nodes == XG_NOALLOC_MAX_NODES uses the whole static array: still the existing entry point's result"""
    import torch
    from xgnn_b200._clib import check, lib
    torch.cuda.set_device(0)
    n, V = 1 << 24, 1 << 20
    g = torch.Generator(device="cuda").manual_seed(5)
    d_table = torch.randint(0, V, (V,), generator=g, device="cuda", dtype=torch.int32)
    d_table[torch.rand(V, generator=g, device="cuda") < 0.5] = -1
    nodes = torch.randint(0, V, (n,), generator=g, device="cuda", dtype=torch.int32)
    want = _existing(lib, check, torch, d_table, nodes)
    bufs = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4)]
    counts = torch.empty(2, dtype=torch.int64, device="cuda")
    for call in range(2):
        _noalloc(lib, check, torch, d_table, nodes, bufs, counts)
        nm, nc = counts.tolist()
        assert (nm, nc) == (want[0].numel(), want[2].numel())
        for gq, w in zip([bufs[0][:nm], bufs[1][:nm], bufs[2][:nc], bufs[3][:nc]], want):
            assert torch.equal(gq, w)
