"""ctypes access to the CPU oracle (oracle/liboracle.so) and, when prebuilt, to the reference's own
CPU code (oracle/_ref/libxgnn_ref.so).  TEST INFRASTRUCTURE -- never imported by xgnn_b200/."""
from __future__ import annotations

import ctypes
import sys
from ctypes import POINTER, c_double, c_int, c_size_t, c_uint32, c_uint64, c_void_p
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_SO = ROOT / "oracle" / "liboracle.so"
REF_SO = ROOT / "oracle" / "_ref" / "libxgnn_ref.so"
EMPTY = 0xFFFFFFFF

orc = ctypes.CDLL(str(ORACLE_SO))
ref = ctypes.CDLL(str(REF_SO)) if REF_SO.exists() else None
# the reference's CUDA samplers + ICS'22 solver + PreSampler compiled in place (oracle/Makefile `refcuda`)
REF_CUDA_SO = ROOT / "oracle" / "_ref" / "libxgnn_ref_cuda.so"
refcu = ctypes.CDLL(str(REF_CUDA_SO)) if REF_CUDA_SO.exists() else None

u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")


def _sig(lib, name, restype, *args):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(args)


_sig(orc, "orc_sample_khop0", c_size_t, u32p, u32p, u32p, c_size_t, c_size_t, u32p, u32p, c_uint64)
_sig(orc, "orc_sample_counts", None, u32p, u32p, c_size_t, c_size_t, u32p)
_sig(orc, "orc_table_create", c_void_p, c_size_t)
_sig(orc, "orc_table_destroy", None, c_void_p)
_sig(orc, "orc_table_reset", None, c_void_p)
_sig(orc, "orc_table_num_items", c_size_t, c_void_p)
_sig(orc, "orc_table_populate", c_int, c_void_p, u32p, c_size_t)
_sig(orc, "orc_table_map_nodes", None, c_void_p, u32p, c_size_t)
_sig(orc, "orc_table_map_edges", c_int, c_void_p, u32p, u32p, c_size_t, u32p, u32p)
_sig(orc, "orc_table_lookup", None, c_void_p, u32p, c_size_t, u32p)
_sig(orc, "orc_extract", None, c_void_p, c_void_p, u32p, c_size_t, c_size_t)
_sig(orc, "orc_mock_extract", None, c_void_p, c_void_p, u32p, c_size_t, c_size_t, ctypes.c_uint)
_sig(orc, "orc_route_node", None, c_uint32, c_uint32, c_uint32, POINTER(c_uint32), POINTER(c_uint32))
_sig(orc, "orc_num_cache_node", c_uint32, u32p, c_uint32, c_double)
_sig(orc, "orc_partition_csr", c_size_t, u32p, u32p, c_uint32, c_uint32, c_uint32, c_void_p, c_void_p,
     POINTER(c_size_t))
_sig(orc, "orc_build_cache_table", None, u32p, c_size_t, u32p, c_size_t)
_sig(orc, "orc_miss_cache_index", None, u32p, u32p, c_size_t, u32p, u32p, POINTER(c_size_t), u32p, u32p,
     POINTER(c_size_t))
_sig(orc, "orc_partition_features", c_size_t, c_void_p, c_size_t, u32p, c_size_t, c_uint32, c_uint32, c_void_p)
_sig(orc, "orc_extract_ggms", None, c_void_p, c_size_t, u32p, c_size_t, c_void_p, c_void_p, c_void_p, c_uint32)
f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_sig(orc, "orc_sample_with_replacement", c_size_t, c_int, u32p, u32p, c_void_p, c_void_p, c_void_p, u32p, c_size_t,
     c_size_t, u32p, u32p, u32p)
_sig(orc, "orc_sample_weighted_hash_dedup", c_size_t, u32p, u32p, f32p, u32p, u32p, c_size_t, c_size_t, c_size_t, u32p,
     u32p, u32p)
_sig(orc, "orc_random_walk_topk", c_size_t, u32p, u32p, u32p, c_size_t, c_size_t, c_double, c_size_t, c_size_t, u32p,
     u32p, u32p, u32p)


def u32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a).astype(np.uint32, copy=False))


# ---- oracle wrappers --------------------------------------------------------------------------
def sample_khop0(indptr, indices, inp, fanout, seed=1):
    inp = u32(inp)
    src = np.empty(max(len(inp) * fanout, 1), np.uint32)
    dst = np.empty(max(len(inp) * fanout, 1), np.uint32)
    n = orc.orc_sample_khop0(u32(indptr), u32(indices), inp, len(inp), fanout, src, dst, seed)
    return src[:n].copy(), dst[:n].copy()


def _p(a):
    return None if a is None else a.ctypes.data


def sample_with_replacement(mode, indptr, indices, inp, fanout, rnd, prob=None, alias=None, prefix=None):
    """mode 1 = KHop1, 2 = alias, 4 = prefix (SampleType numbering); rnd = uint32 [n * fanout, 4]"""
    inp = u32(inp)
    n = max(len(inp) * fanout, 1)
    src, dst = np.empty(n, np.uint32), np.empty(n, np.uint32)
    keep = [np.ascontiguousarray(x) if x is not None else None for x in (prob, alias, prefix)]
    m = orc.orc_sample_with_replacement(mode, u32(indptr), u32(indices), _p(keep[0]), _p(keep[1]), _p(keep[2]), inp,
                                        len(inp), fanout, np.ascontiguousarray(rnd).reshape(-1), src, dst)
    return src[:m].copy(), dst[:m].copy()


def sample_weighted_hash_dedup(indptr, indices, inp, fanout, rnd, prob, alias, tries_per_slot=16):
    inp = u32(inp)
    n = max(len(inp) * fanout, 1)
    src, dst = np.empty(n, np.uint32), np.empty(n, np.uint32)
    m = orc.orc_sample_weighted_hash_dedup(u32(indptr), u32(indices), np.ascontiguousarray(prob, np.float32),
                                           u32(alias), inp, len(inp), fanout, tries_per_slot,
                                           np.ascontiguousarray(rnd).reshape(-1), src, dst)
    return src[:m].copy(), dst[:m].copy()


def random_walk_topk(indptr, indices, inp, walk_len, restart_prob, num_walk, K, rnd):
    inp = u32(inp)
    n = max(len(inp) * K, 1)
    outs = [np.empty(n, np.uint32) for _ in range(3)]
    m = orc.orc_random_walk_topk(u32(indptr), u32(indices), inp, len(inp), walk_len, restart_prob, num_walk, K,
                                 np.ascontiguousarray(rnd).reshape(-1), *outs)
    return tuple(o[:m].copy() for o in outs)


class Table:
    """CPUHashTable0 restatement (first-occurrence ids)."""

    def __init__(self, max_items):
        self.h = orc.orc_table_create(max_items)

    def __del__(self):
        orc.orc_table_destroy(self.h)

    def reset(self):
        orc.orc_table_reset(self.h)

    def populate(self, a):
        a = u32(a)
        assert orc.orc_table_populate(self.h, a, len(a)) == 0

    def num_items(self):
        return orc.orc_table_num_items(self.h)

    def unique(self):
        out = np.empty(max(self.num_items(), 1), np.uint32)
        orc.orc_table_map_nodes(self.h, out, self.num_items())
        return out[: self.num_items()]

    def map_edges(self, src, dst):
        src, dst = u32(src), u32(dst)
        ns, nd = np.empty(max(len(src), 1), np.uint32), np.empty(max(len(src), 1), np.uint32)
        assert orc.orc_table_map_edges(self.h, src, dst, len(src), ns, nd) == 0
        return ns[: len(src)], nd[: len(src)]

    def lookup(self, keys):
        keys = u32(keys)
        out = np.empty(max(len(keys), 1), np.uint32)
        orc.orc_table_lookup(self.h, keys, len(keys), out)
        return out[: len(keys)]


def extract(src: np.ndarray, index) -> np.ndarray:
    index = u32(index)
    src = np.ascontiguousarray(src)
    row_bytes = src.strides[0] if src.ndim == 2 else src.itemsize
    out = np.empty((len(index),) + src.shape[1:], src.dtype)
    orc.orc_extract(out.ctypes.data, src.ctypes.data, index, len(index), row_bytes)
    return out


def mock_extract(src: np.ndarray, index, empty_feat_bits: int) -> np.ndarray:
    """CPUMockExtract (cpu/cpu_extraction.cc:47-64): out[i] = src[index[i] & (2^bits - 1)]"""
    index = u32(index)
    src = np.ascontiguousarray(src)
    out = np.empty((len(index),) + src.shape[1:], src.dtype)
    row_bytes = src.strides[0]
    orc.orc_mock_extract(out.ctypes.data, src.ctypes.data, index, len(index), row_bytes, empty_feat_bits)
    return out


def route_node(v, num_part, num_cache_node):
    p, r = c_uint32(), c_uint32()
    orc.orc_route_node(v, num_part, num_cache_node, ctypes.byref(p), ctypes.byref(r))
    return p.value, r.value


def num_cache_node(indptr, ratio):
    indptr = u32(indptr)
    return orc.orc_num_cache_node(indptr, len(indptr) - 1, ratio)


def partition_csr(indptr, indices, ncn, part, num_part):
    indptr, indices = u32(indptr), u32(indices)
    ln = c_size_t()
    e = orc.orc_partition_csr(indptr, indices, ncn, part, num_part, None, None, ctypes.byref(ln))
    pi = np.zeros(ln.value, np.uint32)
    px = np.zeros(max(e, 1), np.uint32)
    orc.orc_partition_csr(indptr, indices, ncn, part, num_part, pi.ctypes.data, px.ctypes.data, ctypes.byref(ln))
    return pi, px[:e]


def build_cache_table(rank_nodes, num_cached, num_node):
    t = np.empty(num_node, np.uint32)
    orc.orc_build_cache_table(u32(rank_nodes), num_cached, t, num_node)
    return t


def miss_cache_index(table, nodes):
    nodes = u32(nodes)
    n = len(nodes)
    bufs = [np.empty(max(n, 1), np.uint32) for _ in range(4)]
    nm, nc = c_size_t(), c_size_t()
    orc.orc_miss_cache_index(u32(table), nodes, n, bufs[0], bufs[1], ctypes.byref(nm), bufs[2], bufs[3],
                             ctypes.byref(nc))
    return bufs[0][: nm.value], bufs[1][: nm.value], bufs[2][: nc.value], bufs[3][: nc.value]


def partition_features(feat, rank_nodes, num_cached, part, num_part):
    feat = np.ascontiguousarray(feat)
    rb = feat.strides[0]
    rows = orc.orc_partition_features(feat.ctypes.data, rb, u32(rank_nodes), num_cached, part, num_part, None)
    out = np.empty((max(rows, 1),) + feat.shape[1:], feat.dtype)
    orc.orc_partition_features(feat.ctypes.data, rb, u32(rank_nodes), num_cached, part, num_part, out.ctypes.data)
    return out[:rows]


def extract_ggms(feat, nodes, table, parts):
    feat = np.ascontiguousarray(feat)
    nodes = u32(nodes)
    rb = feat.strides[0]
    out = np.empty((len(nodes),) + feat.shape[1:], feat.dtype)
    arr = (c_void_p * max(len(parts), 1))(*[p.ctypes.data for p in parts])
    orc.orc_extract_ggms(out.ctypes.data, rb, nodes, len(nodes), u32(table).ctypes.data if table is not None else None,
                         feat.ctypes.data, arr, len(parts))
    return out


def presample_rank(num_node, input_node_lists):
    """dist::PreSampler (dist/pre_sampler.cc:76-152): freq[v] = number of batches whose input_nodes contain v;
    ranking = sort of the 64-bit words {low: id, high: freq} with std::greater (freq desc, then id desc)."""
    freq = np.zeros(num_node, np.uint64)
    for nodes in input_node_lists:
        freq[u32(nodes)] += 1
    words = (freq << np.uint64(32)) | np.arange(num_node, dtype=np.uint64)
    words = np.sort(words)[::-1]
    return (words & np.uint64(0xFFFFFFFF)).astype(np.uint32), (words >> np.uint64(32)).astype(np.uint32)


def coo_to_csc_indptr(col, num_dst):
    """What the caller-side conversion computes from the COO the reference hands over (torch/adapter.py:98-103,
    create_unitgraph_from_coo(..., ['coo','csr','csc']) -> DGL's COOToCSR on the destination axis): a counting
    sort's offsets.  For node-major edges the sort is the identity, so (indptr, row) IS the CSC form."""
    col = u32(col)
    counts = np.bincount(col, minlength=num_dst).astype(np.uint32)
    indptr = np.zeros(num_dst + 1, np.uint32)
    np.cumsum(counts, out=indptr[1:])
    assert len(col) == 0 or np.all(col[1:] >= col[:-1]), "edges are not destination-major"
    return indptr


def hybrid_placement(feat, rank_nodes, num_cached, num_replicated, num_part):
    """Two-tier GGMS placement (B200 extension, include/xgnn_b200.h xg_feature_store): cache ranks
    [0, num_replicated) form the replica every GPU holds; rank r >= num_replicated goes to part
    (r - num_replicated) % P, row (r - num_replicated) / P, i.e. DeviceDistFeature's rule
    (cuda/dist_graph.h:203-207, cuda/dist_graph.cu:493-505) applied to the remaining ranks."""
    ranked = u32(rank_nodes)[:num_cached]
    replica = np.ascontiguousarray(feat[ranked[:num_replicated]])
    rest = ranked[num_replicated:]
    parts = [np.ascontiguousarray(feat[rest[p::num_part]]) for p in range(num_part)]
    return replica, parts


def extract_hybrid(feat, nodes, table, replica, parts):
    """Row-by-row routing oracle of the two-tier store (numpy restatement of route_row in csrc/extract.cu,
    which follows DeviceDistFeature::Get, cuda/dist_graph.h:203-207, after removing the replicated ranks)."""
    nodes = u32(nodes)
    out = np.empty((len(nodes),) + feat.shape[1:], feat.dtype)
    P, nrep = len(parts), len(replica)
    for i, v in enumerate(nodes):
        r = int(table[v])
        if r == 0xFFFFFFFF:
            out[i] = feat[v]
        elif r < nrep:
            out[i] = replica[r]
        else:
            r -= nrep
            out[i] = parts[r % P][r // P]
    return out


_sig(orc, "orc_minibatch", c_size_t, u32p, u32p, u32p, c_size_t, POINTER(c_size_t), c_int, c_uint64,
     POINTER(c_void_p), POINTER(c_void_p), POINTER(c_size_t), u32p, c_size_t)


def predict_num_nodes(batch, fanouts):
    """PredictNumNodes, common.cc:488-497: worst-case |input_nodes| of a batch (fanouts consumed last -> first)."""
    cur, total = batch, batch
    for f in reversed(list(fanouts)):
        cur = cur * f
        total += cur
    return total


def minibatch(indptr, indices, seeds, fanouts, seed=1):
    """DoCPUSample call order (cpu/cpu_loops.cc:55-192) on the oracle's table + khop0."""
    seeds = u32(seeds)
    L = len(fanouts)
    max_nodes = min(predict_num_nodes(len(seeds), fanouts), len(indptr) - 1 + len(seeds)) + 16
    # worst-case edges per layer
    caps, cur = [0] * L, len(seeds)
    for l in range(L - 1, -1, -1):
        caps[l] = cur * fanouts[l] + 1
        cur = min(cur + caps[l], max_nodes)
    rows = [np.empty(c, np.uint32) for c in caps]
    cols = [np.empty(c, np.uint32) for c in caps]
    rp = (c_void_p * L)(*[r.ctypes.data for r in rows])
    cp = (c_void_p * L)(*[c.ctypes.data for c in cols])
    counts = (c_size_t * (3 * L))()
    fan = (c_size_t * L)(*fanouts)
    inp = np.empty(max_nodes, np.uint32)
    n = orc.orc_minibatch(u32(indptr), u32(indices), seeds, len(seeds), fan, L, seed, rp, cp, counts, inp, max_nodes)
    return {
        "row": [rows[l][: counts[3 * l + 2]].copy() for l in range(L)],
        "col": [cols[l][: counts[3 * l + 2]].copy() for l in range(L)],
        "num_src": [counts[3 * l] for l in range(L)],
        "num_dst": [counts[3 * l + 1] for l in range(L)],
        "num_edge": [counts[3 * l + 2] for l in range(L)],
        "input_nodes": inp[:n].copy(),
    }


# ---- reference (oracle/_ref) wrappers -----------------------------------------------------------
if ref is not None:
    _sig(ref, "ref_set_threads", None, c_int)
    _sig(ref, "ref_cpu_sample_khop0", None, u32p, u32p, u32p, c_size_t, u32p, u32p, POINTER(c_size_t), c_size_t)
    _sig(ref, "ref_cpu_extract", None, c_void_p, c_void_p, u32p, c_size_t, c_size_t, c_int)
    _sig(ref, "ref_hashtable_create", c_void_p, c_int, c_size_t)
    _sig(ref, "ref_hashtable_destroy", None, c_void_p)
    _sig(ref, "ref_hashtable_reset", None, c_void_p)
    _sig(ref, "ref_hashtable_populate", None, c_void_p, u32p, c_size_t)
    _sig(ref, "ref_hashtable_num_items", c_size_t, c_void_p)
    _sig(ref, "ref_hashtable_map_nodes", None, c_void_p, u32p, c_size_t)
    _sig(ref, "ref_hashtable_map_edges", None, c_void_p, u32p, u32p, c_size_t, u32p, u32p)
    for _n in ("f32", "f16", "i64", "u8", "i32", "f64"):
        _sig(ref, f"ref_dtype_{_n}", c_int)
    ref.ref_set_threads(4)

    REF_DTYPE = {np.dtype(np.float32): ref.ref_dtype_f32(), np.dtype(np.float16): ref.ref_dtype_f16(),
                 np.dtype(np.int64): ref.ref_dtype_i64(), np.dtype(np.uint8): ref.ref_dtype_u8(),
                 np.dtype(np.int32): ref.ref_dtype_i32(), np.dtype(np.float64): ref.ref_dtype_f64()}

    def ref_sample_khop0(indptr, indices, inp, fanout):
        inp = u32(inp)
        src = np.empty(max(len(inp) * fanout, 1), np.uint32)
        dst = np.empty(max(len(inp) * fanout, 1), np.uint32)
        n = c_size_t()
        ref.ref_cpu_sample_khop0(u32(indptr), u32(indices), inp, len(inp), src, dst, ctypes.byref(n), fanout)
        return src[: n.value].copy(), dst[: n.value].copy()

    def ref_mock_extract(src: np.ndarray, index, bits: int):
        index = u32(index)
        src = np.ascontiguousarray(src)
        dim = src.shape[1] if src.ndim == 2 else 1
        out = np.empty((len(index),) + src.shape[1:], src.dtype)
        code = {np.dtype(np.float32): ref.ref_dtype_f32(), np.dtype(np.float16): ref.ref_dtype_f16(),
                np.dtype(np.int64): ref.ref_dtype_i64(), np.dtype(np.uint8): ref.ref_dtype_u8()}[src.dtype]
        ref.ref_cpu_mock_extract(ctypes.c_void_p(out.ctypes.data), ctypes.c_void_p(src.ctypes.data),
                                 index.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(len(index)), ctypes.c_size_t(dim),
                                 code, bits)
        return out

    def ref_extract(src: np.ndarray, index):
        index = u32(index)
        src = np.ascontiguousarray(src)
        dim = src.shape[1] if src.ndim == 2 else 1
        out = np.empty((len(index),) + src.shape[1:], src.dtype)
        ref.ref_cpu_extract(out.ctypes.data, src.ctypes.data, index, len(index), dim, REF_DTYPE[src.dtype])
        return out

    class RefTable:
        def __init__(self, kind, max_items):
            self.h = ref.ref_hashtable_create(kind, max_items)

        def __del__(self):
            ref.ref_hashtable_destroy(self.h)

        def reset(self):
            ref.ref_hashtable_reset(self.h)

        def populate(self, a):
            a = u32(a)
            ref.ref_hashtable_populate(self.h, a, len(a))

        def num_items(self):
            return ref.ref_hashtable_num_items(self.h)

        def unique(self):
            out = np.empty(max(self.num_items(), 1), np.uint32)
            ref.ref_hashtable_map_nodes(self.h, out, self.num_items())
            return out[: self.num_items()]

        def map_edges(self, src, dst):
            src, dst = u32(src), u32(dst)
            ns, nd = np.empty(max(len(src), 1), np.uint32), np.empty(max(len(src), 1), np.uint32)
            ref.ref_hashtable_map_edges(self.h, src, dst, len(src), ns, nd)
            return ns[: len(src)], nd[: len(src)]


# the engine with the non-uniform samplers: replay the Philox words layer by layer (salt = kSampleSalt + layer)
def minibatch_ext(g, seeds, fan, sample_type, seed, batch_key, rw=None):
    import philox
    O = sys.modules[__name__]
    L = len(fan)
    table = O.Table(200000)
    table.populate(seeds)
    cur = seeds.copy()
    out = {"row": [None] * L, "col": [None] * L, "data": [None] * L, "num_src": [0] * L, "num_dst": [0] * L}
    for l in range(L - 1, -1, -1):
        salt = philox.SALT + l
        n = len(cur)
        if sample_type == 3:
            Lw, p, W = rw
            src, dst, cnt = O.random_walk_topk(g["indptr"], g["indices"], cur, Lw, p, W, fan[l],
                                               philox.words_walks(seed, batch_key, n, W, Lw, salt=salt))
            out["data"][l] = cnt
        elif sample_type == 6:
            src, dst = O.sample_weighted_hash_dedup(g["indptr"], g["indices"], cur, fan[l],
                                                    philox.words_slots(seed, batch_key, n, 16 * fan[l], c3=1, salt=salt),
                                                    g["prob"], g["alias"])
        else:
            src, dst = O.sample_with_replacement(sample_type, g["indptr"], g["indices"], cur, fan[l],
                                                 philox.words_slots(seed, batch_key, n, fan[l], salt=salt),
                                                 prob=g["prob"], alias=g["alias"], prefix=g["prefix"])
        table.populate(dst)
        out["col"][l], out["row"][l] = table.map_edges(src, dst)
        out["num_dst"][l], out["num_src"][l] = n, table.num_items()
        cur = table.unique()
    out["input_nodes"] = cur
    return out


def ics22_in_degree_rank(indices, num_node):
    """ICS22SongDistGraph ctor (cuda/ics22_song_dist_feature.cc:196-213) + the sort at :48-56: weight[v] = number of
    times v occurs in `indices`, all nodes sorted by (weight, id) with std::greater (both descending)."""
    w = np.bincount(u32(indices), minlength=num_node).astype(np.uint64)
    words = np.sort((w << np.uint64(32)) | np.arange(num_node, dtype=np.uint64))[::-1]
    return (words & np.uint64(0xFFFFFFFF)).astype(np.uint32), (words >> np.uint64(32)).astype(np.uint32)


def ics22_placement(weight, num_cached, num_dev, clique_size, alpha):
    """ICS22SongPlacementSolver (cuda/ics22_song_dist_feature.cc:22-167) restated for ALL devices at once, as the
    reference runs it: returns (ranking u32[V], device_map[num_dev][V], new_idx_map[num_dev][V],
    cached_nodes[num_dev][num_cached], total_cached).  Pure-Python loop: small cases only.
    Parity unpinned by the reference's own tests (it has none for this solver); this follows the source line by line
    in meaning: the per-device maps, the group-of-(c-1) device order refresh, the unsigned underflow stop."""
    weight = np.asarray(weight, dtype=np.float64)
    V = len(weight)
    EMPTY = 0xFFFFFFFF
    ranked = sorted(range(V), key=lambda v: (weight[v], v), reverse=True)  # std::greater<pair<double, IdType>>
    dev_map = np.full((num_dev, V), EMPTY, np.uint32)
    idx_map = np.full((num_dev, V), EMPTY, np.uint32)
    cached = np.zeros((num_dev, num_cached), np.uint32)
    for d in range(num_dev):
        for i in range(num_cached):
            idx_map[d, ranked[i]] = i
            cached[d, i] = ranked[i]
            dev_map[d, ranked[i]] = d
    total = num_cached
    if clique_size == 1:
        return np.array(ranked, np.uint32), dev_map, idx_map, cached, total
    p_accum = [0.0] * clique_size

    def device_order():
        return [d for _, d in sorted((p_accum[d], d) for d in range(clique_size))]

    order = device_order()
    num_cliques = num_dev // clique_size
    i = 0
    while i < V - num_cached:
        if i % (clique_size - 1) == 0:
            order = device_order()
        cand = ranked[num_cached + i]
        slot = num_cached - 1 - i // (clique_size - 1)
        if slot < 0:  # IdType underflow -> ">= num_cached_node" -> break (:112-114)
            break
        victim = ranked[slot]
        if not (weight[cand] >= alpha * weight[victim]):
            break
        cur = order[i % (clique_size - 1)]
        p_accum[cur] += weight[cand]
        for j in range(num_cliques):
            for k in range(clique_size):
                dev_map[j * clique_size + k, cand] = j * clique_size + cur
                idx_map[j * clique_size + k, cand] = slot
            dev_map[j * clique_size + cur, victim] = j * clique_size + order[-1]
            cached[j * clique_size + cur, slot] = cand
        i += 1
    total = num_cached + i
    return np.array(ranked, np.uint32), dev_map, idx_map, cached, total


def ics22_compact_table(ranking, dev_map, idx_map, total_cached, device, bitwidth=3):
    """GPUCacheManager's compact hashtable for one worker (cuda/cuda_cache_manager_host.cc:203-222):
    table[node] = (device_map[node] << (32 - b)) | idx_map[node] for node in ranking[:total_cached]."""
    shift = 32 - bitwidth
    table = np.full(dev_map.shape[1], 0xFFFFFFFF, np.uint32)
    nodes = u32(ranking[:total_cached])
    table[nodes] = ((dev_map[device, nodes].astype(np.uint64) << np.uint64(shift)) | idx_map[device, nodes]).astype(np.uint32)
    return table


def extract_ics22(parts, table, host_feat, nodes, bitwidth=3):
    """DeviceICS22SongDistFeature::Get (cuda/ics22_song_dist_feature.h:28-47) for cached nodes, host rows otherwise"""
    shift = 32 - bitwidth
    out = np.empty((len(nodes), host_feat.shape[1]), host_feat.dtype)
    for i, v in enumerate(u32(nodes)):
        t = int(table[v])
        out[i] = host_feat[v] if t == 0xFFFFFFFF else parts[t >> shift][t & ((1 << shift) - 1)]
    return out


# ---- the reference's own code compiled in place (oracle/_ref/libxgnn_ref_cuda.so) ---------------------------------
if refcu is not None:
    _sig(refcu, "refcu_available", c_int)
    _sig(refcu, "refcu_graph_create", c_void_p, c_int, u32p, u32p, c_uint32, c_size_t, c_void_p, c_void_p, c_void_p)
    _sig(refcu, "refcu_graph_destroy", None, c_void_p)
    _sig(refcu, "refcu_reset_states", None, c_void_p)
    _sig(refcu, "refcu_table_create", c_void_p, c_int, c_size_t)
    _sig(refcu, "refcu_table_destroy", None, c_void_p)
    _sig(refcu, "refcu_table_reset", None, c_void_p)
    _sig(refcu, "refcu_table_num_items", c_size_t, c_void_p)
    _sig(refcu, "refcu_table_fill", c_size_t, c_void_p, c_int, u32p, c_size_t, u32p, c_size_t)
    _sig(refcu, "refcu_table_map_edges", None, c_void_p, u32p, u32p, c_size_t, u32p, u32p)
    _sig(refcu, "refcu_cache_create", c_void_p, c_int, c_void_p, c_size_t, c_size_t, c_int, c_size_t, u32p, c_double)
    _sig(refcu, "refcu_cache_destroy", None, c_void_p)
    _sig(refcu, "refcu_cache_extract", None, c_void_p, u32p, c_size_t, c_void_p, u32p, u32p, POINTER(c_size_t), u32p, u32p,
         POINTER(c_size_t))
    _sig(refcu, "refcu_graph_set_parts", None, c_void_p, c_uint32, c_uint32, u32p, c_void_p, u32p, c_void_p)
    _sig(refcu, "refcu_sample", c_size_t, c_void_p, c_int, u32p, c_size_t, c_size_t, c_size_t, c_double, c_size_t, u32p,
         u32p, u32p)
    _sig(refcu, "ref_ics22_solve", c_uint32, c_void_p, c_uint32, c_uint32, c_uint32, c_double, c_uint32, u32p, u32p, u32p,
         u32p)
    _sig(refcu, "ref_presample_rank", None, c_uint32, c_int, c_size_t, u32p, c_void_p, u32p, u32p)
    _sig(refcu, "refcu_dist_feature_get", None, c_int, f32p, c_void_p, c_uint32, c_uint32, c_uint32, c_uint32, c_void_p,
         c_void_p, c_uint32, u32p, c_size_t, f32p)


def ref_dist_feature_get(parts, num_cache_node, ranks, compact_bitwidth=0, device_map=None, new_idx_map=None, device=0):
    """rows of `ranks` through the REFERENCE's device accessors: DeviceDistFeature::Get (cuda/dist_graph.h:193-207) over
    the float32 parts, or DeviceICS22SongDistFeature::Get (cuda/ics22_song_dist_feature.h:27-47) when compact_bitwidth
    or the (device_map, new_idx_map) pair is given"""
    dim = parts[0].shape[1]
    flat = [np.ascontiguousarray(p, np.float32).ravel() for p in parts]
    off = np.zeros(len(parts) + 1, np.uint64)
    np.cumsum([len(f) for f in flat], out=off[1:])
    cat = np.concatenate(flat + [np.zeros(1, np.float32)])
    ranks = u32(ranks)
    out = np.empty((len(ranks), dim), np.float32)
    dm = u32(device_map) if device_map is not None else None
    im = u32(new_idx_map) if new_idx_map is not None else None
    refcu.refcu_dist_feature_get(device, cat, off.ctypes.data, len(parts), num_cache_node, dim, compact_bitwidth,
                                 dm.ctypes.data if dm is not None else None, im.ctypes.data if im is not None else None,
                                 len(dm) if dm is not None else 0, ranks, len(ranks), out.reshape(-1))
    return out


def ref_ics22_solve(weight, num_cached, num_dev, clique_size, alpha):
    """The reference's ICS22SongPlacementSolver itself (cuda/ics22_song_dist_feature.cc:22-167); same return
    tuple as ics22_placement() above."""
    w = np.ascontiguousarray(weight, np.float64)
    V = len(w)
    ranking = np.empty(V, np.uint32)
    dev_map, idx_map = np.empty((num_dev, V), np.uint32), np.empty((num_dev, V), np.uint32)
    cached = np.empty((num_dev, max(num_cached, 1)), np.uint32)
    total = refcu.ref_ics22_solve(w.ctypes.data, V, num_cached, num_dev, alpha, clique_size, ranking, dev_map.reshape(-1),
                                  idx_map.reshape(-1), cached.reshape(-1)[:max(num_dev * num_cached, 1)])
    cached = cached.reshape(-1)[:num_dev * num_cached].reshape(num_dev, num_cached)
    return ranking, dev_map, idx_map, cached, int(total)


def ref_presample_rank(num_node, input_node_lists, epochs=1):
    """The reference's dist::PreSampler (dist/pre_sampler.cc:40-165) fed with one input-node list per batch."""
    lists = [u32(a) for a in input_node_lists]
    assert len(lists) % epochs == 0
    nodes = np.concatenate(lists + [np.zeros(1, np.uint32)])
    off = np.cumsum([0] + [len(a) for a in lists]).astype(np.uint64)
    rank_nodes, rank_freq = np.empty(num_node, np.uint32), np.empty(num_node, np.uint32)
    refcu.ref_presample_rank(num_node, epochs, len(lists) // epochs, nodes, off.ctypes.data, rank_nodes, rank_freq)
    return rank_nodes, rank_freq


class RefCudaGraph:
    """A CSR (+ optional weight tables) resident on the GPU for the reference's own CUDA samplers."""

    def __init__(self, indptr, indices, prob=None, alias=None, prefix=None, device=0):
        self._keep = [u32(indptr), u32(indices),
                      None if prob is None else np.ascontiguousarray(prob, np.float32),
                      None if alias is None else u32(alias),
                      None if prefix is None else np.ascontiguousarray(prefix, np.float32)]
        k = self._keep
        self.h = refcu.refcu_graph_create(device, k[0], k[1], len(k[0]) - 1, len(k[1]), _p(k[2]), _p(k[3]), _p(k[4]))

    def __del__(self):
        if getattr(self, "h", None):
            refcu.refcu_graph_destroy(self.h)
            self.h = None

    def set_parts(self, parts, num_cache_node):
        """GGMS partitioning (list of (indptr, indices) per part, as partition_csr() returns them): KHop0 / KHop3 /
        random walk then go through the reference's DeviceDistGraph (cuda/dist_graph.h:114-151)"""
        ip = [u32(a) for a, _ in parts]
        ix = [u32(b) if len(b) else np.zeros(1, np.uint32) for _, b in parts]
        ipo = np.cumsum([0] + [len(a) for a in ip]).astype(np.uint64)
        ixo = np.cumsum([0] + [len(b) for b in ix]).astype(np.uint64)
        refcu.refcu_graph_set_parts(self.h, len(parts), num_cache_node, np.concatenate(ip), ipo.ctypes.data,
                                    np.concatenate(ix), ixo.ctypes.data)

    def reset_states(self):
        """drop the curand states: the next call seeds fresh ones from the wall clock (cuda_random_states.cu:106-108)"""
        refcu.refcu_reset_states(self.h)

    def sample(self, sample_type, inp, fanout, walk_length=0, restart_prob=0.0, num_walk=0):
        """-> (src, dst, data) as the reference's GPUSample* wrote them (data only for random walk)"""
        inp = u32(inp)
        cap = max(len(inp) * fanout, 1)
        src, dst, data = np.empty(cap, np.uint32), np.empty(cap, np.uint32), np.empty(cap, np.uint32)
        n = refcu.refcu_sample(self.h, sample_type, inp, len(inp), fanout, walk_length, restart_prob, num_walk, src,
                               dst, data)
        return src[:n].copy(), dst[:n].copy(), data[:n].copy()


class RefCudaCache:
    """The reference's GPUCacheManager (single-GPU feature cache) built from (feat, rank list, cache percentage)"""

    def __init__(self, feat, rank_nodes, cache_percentage, device=0):
        feat = np.ascontiguousarray(feat)
        self.feat, self.ranks = feat, u32(rank_nodes)
        code = REF_DTYPE[feat.dtype] if ref is not None else {np.dtype(np.float32): 0, np.dtype(np.float16): 2}[feat.dtype]
        self.h = refcu.refcu_cache_create(device, feat.ctypes.data, feat.shape[0], feat.shape[1], code, feat.itemsize,
                                          self.ranks, cache_percentage)

    def __del__(self):
        if getattr(self, "h", None):
            refcu.refcu_cache_destroy(self.h)
            self.h = None

    def extract(self, nodes):
        """-> (rows, miss_src, miss_dst, cache_src, cache_dst) of GetMissCacheIndex + GPUExtractMissData + CombineCacheData"""
        nodes = u32(nodes)
        n = len(nodes)
        out = np.empty((n, self.feat.shape[1]), self.feat.dtype)
        idx = [np.empty(max(n, 1), np.uint32) for _ in range(4)]
        nm, nc = c_size_t(0), c_size_t(0)
        refcu.refcu_cache_extract(self.h, nodes, n, out.ctypes.data, idx[0], idx[1], ctypes.byref(nm), idx[2], idx[3],
                                  ctypes.byref(nc))
        return out, idx[0][:nm.value].copy(), idx[1][:nm.value].copy(), idx[2][:nc.value].copy(), idx[3][:nc.value].copy()


class RefCudaTable:
    """The reference's GPU OrderedHashTable (cuda/cuda_hashtable.cu) + GPUMapEdges (cuda/cuda_mapping.cu)"""

    def __init__(self, size, device=0):
        self.size = size
        self.h = refcu.refcu_table_create(device, size)

    def __del__(self):
        if getattr(self, "h", None):
            refcu.refcu_table_destroy(self.h)
            self.h = None

    def reset(self):
        refcu.refcu_table_reset(self.h)

    def fill(self, ids, unique_input=False):
        """FillWithDuplicates (or FillWithUnique) -> the whole unique list afterwards"""
        ids = u32(ids)
        out = np.empty(self.size, np.uint32)
        n = refcu.refcu_table_fill(self.h, 1 if unique_input else 0, ids, len(ids), out, self.size)
        return out[:n].copy()

    def map_edges(self, src, dst):
        src, dst = u32(src), u32(dst)
        ns, nd = np.empty(max(len(src), 1), np.uint32), np.empty(max(len(src), 1), np.uint32)
        refcu.refcu_table_map_edges(self.h, src, dst, len(src), ns, nd)
        return ns[:len(src)].copy(), nd[:len(src)].copy()
