"""Operator-level Python mirror of the reference's CUDA operators, over the `xg_*` C-ABI.

torch is used for device memory and streams only.  ids are int32 tensors viewing uint32 (as the
reference exposes them, samgraph/torch/adapter.cc:86-120).  Names follow the reference:
OrderedHashTable (cuda/cuda_hashtable.h:103), GPUSampleKHop0/3 (cuda/cuda_function.h),
GPUMapEdges (cuda/cuda_mapping.cu:68), GPUExtract (cuda/cuda_extraction.cu:74),
GetMissCacheIndex / CombineCacheData / GPUExtractMissData (cuda/cuda_cache_manager.h).
"""
from __future__ import annotations

import ctypes
from ctypes import byref, c_size_t, c_uint32, c_void_p

import torch

from . import _clib
from ._clib import XG_EMPTY_KEY, XG_MAX_PARTS, XgFeatureStore, XgGraph, check, lib

kKHop0, kKHop1, kWeightedKHop, kRandomWalk, kWeightedKHopPrefix, kKHop2, kWeightedKHopHashDedup, kKHop3 = range(8)
_DTYPE_CODE = {torch.float32: 0, torch.float64: 1, torch.float16: 2, torch.uint8: 3, torch.int32: 4,
               torch.int8: 5, torch.int64: 6, torch.bfloat16: 2}


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ids(t: torch.Tensor) -> torch.Tensor:
    assert t.dtype == torch.int32 and t.is_contiguous(), "ids must be contiguous int32 (uint32 view)"
    return t


class OrderedHashTable:
    """cuda::OrderedHashTable (cuda/cuda_hashtable.h:103-153) with first-occurrence numbering."""

    def __init__(self, size: int, device: int | torch.device = 0):
        self.device = torch.device("cuda", device) if isinstance(device, int) else device
        h = c_void_p()
        check(lib.xg_hashtable_create(byref(h), size, self.device.index or 0))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            lib.xg_hashtable_destroy(self._h)
            self._h = None

    def Reset(self):
        check(lib.xg_hashtable_reset(self._h, _stream()))

    def FillWithUnique(self, input: torch.Tensor):
        check(lib.xg_hashtable_fill_with_unique(self._h, _ptr(_ids(input)), input.numel(), _stream()))

    def FillWithDupRevised(self, input: torch.Tensor):
        check(lib.xg_hashtable_fill_with_dup_revised(self._h, _ptr(_ids(input)), input.numel(), _stream()))

    def FillWithDupMutable(self, input: torch.Tensor):
        check(lib.xg_hashtable_fill_with_dup_mutable(self._h, _ptr(_ids(input)), input.numel(), _stream()))

    def FillWithDuplicates(self, input: torch.Tensor, unique: torch.Tensor | None = None) -> torch.Tensor:
        """Returns the unique list (old prefix + new ids)."""
        n_old = self.NumItems()
        if unique is None:
            unique = torch.empty(n_old + input.numel(), dtype=torch.int32, device=self.device)
        n = c_uint32(0)
        check(lib.xg_hashtable_fill_with_duplicates(self._h, _ptr(_ids(input)), input.numel(), _ptr(unique),
                                                    byref(n), _stream()))
        return unique[: n.value]

    def NumItems(self) -> int:
        return int(lib.xg_hashtable_num_items(self._h))

    def RefUnique(self) -> torch.Tensor:
        """Copy of the table-resident unique list (the reference hands out the raw pointer)."""
        p = c_void_p()
        n = c_uint32(0)
        check(lib.xg_hashtable_ref_unique(self._h, byref(p), byref(n)))
        out = torch.empty(n.value, dtype=torch.int32, device=self.device)
        check(lib.xg_hashtable_copy_unique(self._h, _ptr(out), _stream()))
        return out

    def CopyUnique(self, unique: torch.Tensor):
        check(lib.xg_hashtable_copy_unique(self._h, _ptr(unique), _stream()))

    def SearchO2N(self, ids: torch.Tensor) -> torch.Tensor:
        out = torch.empty_like(ids)
        check(lib.xg_hashtable_lookup(self._h, _ptr(_ids(ids)), ids.numel(), _ptr(out), _stream()))
        return out

    @property
    def handle(self):
        return self._h


def GPUMapEdges(global_src: torch.Tensor, global_dst: torch.Tensor, table: OrderedHashTable):
    """cuda::GPUMapEdges (cuda/cuda_mapping.cu:68-80) -> (new_src, new_dst)."""
    new_src = torch.empty_like(global_src)
    new_dst = torch.empty_like(global_dst)
    check(lib.xg_map_edges(table.handle, _ptr(_ids(global_src)), _ptr(new_src), _ptr(_ids(global_dst)),
                           _ptr(new_dst), global_src.numel(), _stream()))
    return new_src, new_dst


def make_graph(indptr: torch.Tensor | None = None, indices: torch.Tensor | None = None, *,
               part_indptr=None, part_indices=None, num_cache_node: int = 0, num_node: int | None = None,
               host_indptr: torch.Tensor | int | None = None,
               host_indices: torch.Tensor | int | None = None) -> XgGraph:
    """Flat (DeviceNormalGraph) or sharded (DeviceDistGraph) topology handle, cuda/dist_graph.h:114-180.
    Partition entries may be tensors or raw device pointers (ints, e.g. IPC-mapped peer memory)."""
    g = XgGraph()

    def p(x):
        return x if isinstance(x, int) or x is None else x.data_ptr()

    if part_indptr is None:
        g.part_indptr[0] = p(indptr)
        g.part_indices[0] = p(indices)
        g.num_part = 0
        g.num_cache_node = 0
        g.num_node = (indptr.numel() - 1) if num_node is None else num_node
    else:
        P = len(part_indptr)
        assert P <= XG_MAX_PARTS
        for i in range(P):
            g.part_indptr[i] = p(part_indptr[i])
            g.part_indices[i] = p(part_indices[i])
        g.part_indptr[P] = p(host_indptr)
        g.part_indices[P] = p(host_indices)
        g.num_part = P
        g.num_cache_node = num_cache_node
        g.num_node = num_node or 0
    # the handle holds raw pointers: keep the tensors behind them alive as long as the handle lives
    g._keep = [indptr, indices, part_indptr, part_indices, host_indptr, host_indices]
    return g


def GPUSampleKHop(graph: XgGraph, input: torch.Tensor, fanout: int, *, seed: int = 0, batch_key: int = 0,
                  sample_type: int = kKHop0):
    """GPUSampleKHop0/2/3<GraphType> (cuda/cuda_sampling_khop0.cu:243-335, khop3.cu:234-318).
    Returns (out_src, out_dst) already trimmed to num_out (this wrapper reads the count back)."""
    n = input.numel()
    out_src = torch.empty(max(n * fanout, 1), dtype=torch.int32, device=input.device)
    out_dst = torch.empty(max(n * fanout, 1), dtype=torch.int32, device=input.device)
    num_out = torch.zeros(1, dtype=torch.int64, device=input.device)
    check(lib.xg_sample_khop(byref(graph), sample_type, _ptr(_ids(input)), n, fanout, _ptr(out_src), _ptr(out_dst),
                             _ptr(num_out), seed, batch_key, _stream()))
    e = int(num_out.item())
    return out_src[:e], out_dst[:e]


def GPUSampleKHop1(graph: XgGraph, input: torch.Tensor, fanout: int, *, seed: int = 0, batch_key: int = 0):
    """GPUSampleKHop1 (cuda/cuda_sampling_khop1.cu:130-236) -> (out_src, out_dst), trimmed."""
    n = input.numel()
    out_src = torch.empty(max(n * fanout, 1), dtype=torch.int32, device=input.device)
    out_dst = torch.empty(max(n * fanout, 1), dtype=torch.int32, device=input.device)
    num_out = torch.zeros(1, dtype=torch.int64, device=input.device)
    check(lib.xg_sample_khop1(byref(graph), _ptr(_ids(input)), n, fanout, _ptr(out_src), _ptr(out_dst), _ptr(num_out),
                              seed, batch_key, _stream()))
    e = int(num_out.item())
    return out_src[:e], out_dst[:e]


def GPUSampleWeightedKHop(graph: XgGraph, input: torch.Tensor, fanout: int, *, prob_table=None, alias_table=None,
                          prob_prefix_table=None, sample_type: int = kWeightedKHop, seed: int = 0, batch_key: int = 0):
    """GPUSampleWeightedKHop / ...Prefix / ...HashDedup (cuda/cuda_sampling_weighted_khop*.cu)."""
    n = input.numel()
    out_src = torch.empty(max(n * fanout, 1), dtype=torch.int32, device=input.device)
    out_dst = torch.empty(max(n * fanout, 1), dtype=torch.int32, device=input.device)
    num_out = torch.zeros(1, dtype=torch.int64, device=input.device)
    check(lib.xg_sample_weighted_khop(byref(graph), sample_type, _ptr(prob_table), _ptr(alias_table),
                                      _ptr(prob_prefix_table), _ptr(_ids(input)), n, fanout, _ptr(out_src),
                                      _ptr(out_dst), _ptr(num_out), seed, batch_key, _stream()))
    e = int(num_out.item())
    return out_src[:e], out_dst[:e]


def GPUSampleRandomWalk(graph: XgGraph, input: torch.Tensor, random_walk_length: int, random_walk_restart_prob: float,
                        num_random_walk: int, K: int, *, seed: int = 0, batch_key: int = 0):
    """GPUSampleRandomWalk + FrequencyHashmap::GetTopK (cuda/cuda_sampling_random_walk.cu:116-165)
    -> (out_src, out_dst, out_data = visit counts), trimmed."""
    n = input.numel()
    outs = [torch.empty(max(n * K, 1), dtype=torch.int32, device=input.device) for _ in range(3)]
    num_out = torch.zeros(1, dtype=torch.int64, device=input.device)
    check(lib.xg_sample_random_walk(byref(graph), _ptr(_ids(input)), n, random_walk_length, random_walk_restart_prob,
                                    num_random_walk, K, _ptr(outs[0]), _ptr(outs[1]), _ptr(outs[2]), _ptr(num_out),
                                    seed, batch_key, _stream()))
    e = int(num_out.item())
    return outs[0][:e], outs[1][:e], outs[2][:e]


def GPUExtract(src: torch.Tensor | int, index: torch.Tensor, dim: int | None = None,
               dtype: torch.dtype | None = None) -> torch.Tensor:
    """cuda::GPUExtract (cuda/cuda_extraction.cu:74-116). `src` is a [num_node, dim] tensor on the device or in
    pinned host memory (read through UVA), or a raw pointer with dim/dtype given."""
    if isinstance(src, torch.Tensor):
        dtype = src.dtype
        dim = src.shape[1] if src.dim() == 2 else 1
        sp = src.data_ptr()
    else:
        sp = src
    shape = (index.numel(), dim) if dim != 1 or (isinstance(src, torch.Tensor) and src.dim() == 2) else (index.numel(),)
    out = torch.empty(shape, dtype=dtype, device=index.device)
    check(lib.xg_extract(_ptr(out), sp, _ptr(_ids(index)), index.numel(), dim, _DTYPE_CODE[dtype], _stream()))
    return out


def GPUMockExtract(src: torch.Tensor, index: torch.Tensor, empty_feat: int) -> torch.Tensor:
    """cuda::GPUMockExtract (cuda/cuda_extraction.cu:118-160): out[i] = src[index[i] & (2^empty_feat - 1)]"""
    dim = src.shape[1] if src.dim() == 2 else 1
    out = torch.empty((index.numel(),) + tuple(src.shape[1:]), dtype=src.dtype, device=index.device)
    check(lib.xg_mock_extract(_ptr(out), src.data_ptr(), _ptr(_ids(index)), index.numel(), dim * src.element_size(),
                              empty_feat, _stream()))
    return out


def make_feature_store(row_bytes: int, host: torch.Tensor | int | None, cache_table: torch.Tensor | None = None,
                       parts=None, local_parts=(), replica=None, num_replicated: int = 0,
                       compact_bitwidth: int = 0, replica_words: torch.Tensor | None = None,
                       empty_feat: int = 0) -> XgFeatureStore:
    """DeviceDistFeature + cache table + host table in one handle (cuda/dist_graph.h:182-213).
    empty_feat = k > 0: SAMGRAPH_EMPTY_FEAT mock features - `host` holds 2^k rows and host-resident node v reads row
    v & (2^k - 1) (cuda/cuda_extraction.cu:51-72)."""
    st = XgFeatureStore()
    st.host_row_mask = ((1 << empty_feat) - 1) if empty_feat else 0

    def p(x):
        return x if isinstance(x, int) or x is None else x.data_ptr()

    st.host = p(host)
    st.cache_table = p(cache_table)
    st.row_bytes = row_bytes
    parts = parts or []
    assert len(parts) <= XG_MAX_PARTS
    for i, t in enumerate(parts):
        st.part[i] = p(t)
    st.num_part = len(parts)
    mask = 0
    for i in local_parts:
        mask |= 1 << i
    st.local_part_mask = mask
    st.replica = p(replica) if num_replicated else None
    st.num_replicated = num_replicated if replica is not None else 0
    assert not (compact_bitwidth and st.num_replicated), "the compact (ICS'22) map has no replicated tier of its own"
    st.compact_bitwidth = compact_bitwidth
    st.replica_words = p(replica_words) if (replica_words is not None and st.num_replicated) else None
    st.replica_words_cover_all = int(bool(st.replica_words) and getattr(replica_words, "cover_all", False))
    st._keep = [host, cache_table, parts, replica, replica_words]  # raw pointers inside: keep the tensors alive
    return st


def ExtractGGMS(store: XgFeatureStore, nodes: torch.Tensor, dim: int, dtype: torch.dtype,
                stats: torch.Tensor | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
    """GetMissCacheIndex + GPUExtractMissData + CombineCacheData fused (one launch)."""
    if out is None:
        out = torch.empty((nodes.numel(), dim), dtype=dtype, device=nodes.device)
    check(lib.xg_extract_ggms(byref(store), _ptr(_ids(nodes)), nodes.numel(), _ptr(out), _ptr(stats), _stream()))
    return out


def GetMissCacheIndex(cache_table: torch.Tensor, nodes: torch.Tensor):
    """GPUCacheManager::GetMissCacheIndex (cuda/cuda_cache_manager_device.cu:355-441)
    -> (miss_src, miss_dst, cache_src, cache_dst)."""
    n = nodes.numel()
    bufs = [torch.empty(max(n, 1), dtype=torch.int32, device=nodes.device) for _ in range(4)]
    nm, nc = c_size_t(0), c_size_t(0)
    check(lib.xg_get_miss_cache_index(_ptr(cache_table), _ptr(_ids(nodes)), n, _ptr(bufs[0]), _ptr(bufs[1]), byref(nm),
                                      _ptr(bufs[2]), _ptr(bufs[3]), byref(nc), _stream()))
    return bufs[0][: nm.value], bufs[1][: nm.value], bufs[2][: nc.value], bufs[3][: nc.value]


def GPUExtractMissData(out: torch.Tensor, host_feat: torch.Tensor | int, miss_src, miss_dst, row_bytes: int):
    sp = host_feat if isinstance(host_feat, int) else host_feat.data_ptr()
    check(lib.xg_extract_miss_data(_ptr(out), sp, _ptr(miss_src), _ptr(miss_dst), miss_src.numel(), row_bytes, _stream()))


def CombineCacheData(out: torch.Tensor, store: XgFeatureStore, cache_src, cache_dst):
    check(lib.xg_combine_cache_data(_ptr(out), byref(store), _ptr(cache_src), _ptr(cache_dst), cache_src.numel(), _stream()))


def launch_count() -> int:
    return int(lib.xg_launch_count())


# ------------------------------------------------------------------------------------------------
# Minibatch engine: DoGPUSample + extract stage functions as one replayed CUDA graph (pipeline.cu)
# ------------------------------------------------------------------------------------------------
_TORCH_OF_CODE = {0: torch.float32, 1: torch.float64, 2: torch.float16, 3: torch.uint8, 4: torch.int32,
                  5: torch.int8, 6: torch.int64}


class _DevMem:
    """__cuda_array_interface__ view of memory owned by a batch slot (keeps the batch alive)."""

    def __init__(self, ptr: int, shape, typestr: str, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False),
                                         "version": 3, "strides": None}
        self._owner = owner


_TYPESTR = {torch.float32: "<f4", torch.float64: "<f8", torch.float16: "<f2", torch.uint8: "|u1",
            torch.int32: "<i4", torch.int8: "|i1", torch.int64: "<i8"}


class Minibatch:
    """One finished Task (common.h:246-285): zero-copy torch views of the batch slot's buffers, as
    samgraph/torch/adapter.cc:62-134 hands them out (ids as int32). The slot is released when this
    object dies (or on release())."""

    def __init__(self, sampler: "MinibatchSampler", handle: c_void_p, view: _clib.XgBatchView):
        self._sampler, self._h, self.view = sampler, handle, view
        self.key = view.key
        self.num_layer = view.num_layer
        self.num_src = [int(view.num_src[l]) for l in range(self.num_layer)]
        self.num_dst = [int(view.num_dst[l]) for l in range(self.num_layer)]
        self.num_edge = [int(view.num_edge[l]) for l in range(self.num_layer)]
        self.num_input, self.num_output = int(view.num_input), int(view.num_output)

    def _t(self, ptr, shape, dtype):
        n = 1
        for s in shape:
            n *= s
        if n == 0 or not ptr:
            return torch.empty(shape, dtype=dtype, device=self._sampler.device)
        return torch.as_tensor(_DevMem(ptr, shape, _TYPESTR[dtype], self), device=self._sampler.device)

    def row(self, l):
        return self._t(self.view.row[l], (self.num_edge[l],), torch.int32)

    def col(self, l):
        return self._t(self.view.col[l], (self.num_edge[l],), torch.int32)

    def data(self, l):
        """per-edge payload (random-walk visit counts), samgraph_torch_get_graph_data"""
        return self._t(self.view.data[l], (self.num_edge[l],), torch.int32) if self.view.data[l] else None

    def csc_indptr(self, l):
        """[num_dst + 1] edge offsets of the layer's destination nodes (KHop0/2/3), else None: with row(l) as the
        indices array this is the block in CSC form (no COO -> CSC conversion on the caller side)."""
        p = self.view.csc_indptr[l]
        return self._t(p, (self.num_dst[l] + 1,), torch.int32) if p else None

    def timeline(self):
        """XG_TIMELINE=1: ms since sampler creation of (sample begin, sample end, extract begin, extract end)."""
        ms = (ctypes.c_float * 4)()
        check(lib.xg_batch_timeline(self._h, ms))
        return tuple(ms)

    @property
    def input_nodes(self):
        return self._t(self.view.input_nodes, (self.num_input,), torch.int32)

    @property
    def output_nodes(self):
        return self._t(self.view.output_nodes, (self.num_output,), torch.int32)

    @property
    def feat(self):
        s = self._sampler
        return self._t(self.view.input_feat, (self.num_input, s.feat_dim), s.feat_dtype) if self.view.input_feat else None

    @property
    def label(self):
        return self._t(self.view.output_label, (self.num_output,), torch.int64) if self.view.output_label else None

    def release(self):
        if self._h is not None:
            lib.xg_batch_release(self._h)
            self._h = None

    def __del__(self):
        self.release()


class MinibatchSampler:
    """dist::DoGPUSample + DoGPULabelExtract + DoArch6GetCacheMissIndex + DoArch6GPUCacheFeatureCopy
    (dist/dist_loops.cc) for one sampler GPU."""

    def __init__(self, graph: XgGraph, fanout, batch_size: int, *, store: XgFeatureStore | None = None,
                 feat_dim: int = 0, feat_dtype: torch.dtype = torch.float32, label: torch.Tensor | int | None = None,
                 sample_type: int = kKHop0, seed: int = 0, num_slots: int = 2, max_input_nodes: int = 0,
                 device: int = 0, use_graph: bool = True, stream: torch.cuda.Stream | None = None,
                 stage_timing: bool = False, single_stream: bool = False, random_walk_length: int = 0, num_random_walk: int = 0,
                 random_walk_restart_prob: float = 0.0, prob_table: torch.Tensor | None = None,
                 alias_table: torch.Tensor | None = None, prob_prefix_table: torch.Tensor | None = None):
        cfg = _clib.XgSamplerConfig()
        cfg.num_layer = len(fanout)
        for i, f in enumerate(fanout):
            cfg.fanout[i] = int(f)
        cfg.batch_size = batch_size
        cfg.sample_type = sample_type
        cfg.num_slots = num_slots
        cfg.flags = (0 if use_graph else 1) | (2 if stage_timing else 0) | (4 if single_stream else 0)
        cfg.feat_dim = feat_dim
        cfg.feat_dtype = _DTYPE_CODE[feat_dtype]
        cfg.seed = seed
        cfg.max_input_nodes = max_input_nodes
        cfg.stream = stream.cuda_stream if stream is not None else None
        cfg.random_walk_length, cfg.num_random_walk = random_walk_length, num_random_walk
        cfg.random_walk_restart_prob = random_walk_restart_prob
        cfg.prob_table, cfg.alias_table, cfg.prob_prefix_table = _ptr(prob_table), _ptr(alias_table), _ptr(prob_prefix_table)
        self.device = torch.device("cuda", device)
        self.feat_dim, self.feat_dtype = feat_dim, feat_dtype
        self._keep = [graph, store, label, stream, prob_table, alias_table, prob_prefix_table]
        h = c_void_p()
        lp = label if isinstance(label, int) or label is None else label.data_ptr()
        check(lib.xg_sampler_create(byref(h), byref(graph), byref(store) if store is not None else None, lp,
                                    byref(cfg), device))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # `lib` is gone when the interpreter is shutting down
            lib.xg_sampler_destroy(self._h)
            self._h = None

    def submit(self, seeds, batch_key: int = 0) -> c_void_p:
        """Enqueue one minibatch; `seeds` is a host uint32 numpy array / CPU int32 tensor, or a CUDA int32
        tensor. Returns the in-flight handle for wait()."""
        b = c_void_p()
        if isinstance(seeds, torch.Tensor):
            on_dev = 1 if seeds.is_cuda else 0
            ptr, n = seeds.data_ptr(), seeds.numel()
            if on_dev and n:
                # the engine copies the seeds on ITS stream: order that copy after whatever produced the tensor on
                # torch's current stream, and keep the caching allocator from reusing the memory before the copy ran
                mine = getattr(self, "_sample_stream", None)
                if mine is None:
                    mine = self._sample_stream = torch.cuda.ExternalStream(self.stream_ptr, device=seeds.device)
                cur = torch.cuda.current_stream(seeds.device)
                if cur.cuda_stream != mine.cuda_stream:
                    mine.wait_stream(cur)
                    seeds.record_stream(mine)
        else:
            on_dev, ptr, n = 0, seeds.ctypes.data, len(seeds)
        check(lib.xg_sampler_submit(self._h, ptr, n, on_dev, batch_key, byref(b)))
        return b

    def wait(self, handle: c_void_p) -> Minibatch:
        view = _clib.XgBatchView()
        rc = lib.xg_batch_wait(handle, byref(view))
        if rc != 0:
            lib.xg_batch_release(handle)
            check(rc)
        return Minibatch(self, handle, view)

    def sample(self, seeds, batch_key: int = 0) -> Minibatch:
        return self.wait(self.submit(seeds, batch_key))

    @property
    def launches_per_batch(self) -> int:
        return int(lib.xg_sampler_launches_per_batch(self._h))

    @property
    def stream_ptr(self) -> int:
        return int(lib.xg_sampler_stream(self._h) or 0)

    @property
    def extract_stream(self) -> torch.cuda.Stream:
        """Stream the extract phase (and therefore the end of every batch) is ordered on."""
        return torch.cuda.ExternalStream(int(lib.xg_sampler_extract_stream(self._h) or 0), device=self.device)


def presample_rank(sampler: "MinibatchSampler", num_node: int, batches, key0: int = 0):
    """dist::PreSampler::DoPreSample + GetRankNode/GetFreq (dist/pre_sampler.cc:76-152): run the given seed batches
    through `sampler`, count how many batches need each node as an input node, return (rank_nodes, rank_freq)
    as int32 CUDA tensors (frequency descending, ties by node id descending)."""
    freq = torch.zeros(num_node, dtype=torch.int32, device=sampler.device)
    for i, seeds in enumerate(batches):
        mb = sampler.sample(seeds, batch_key=key0 + i)
        nodes = mb.input_nodes
        check(lib.xg_presample_accumulate(_ptr(freq), num_node, _ptr(nodes), nodes.numel(), None, _stream()))
        torch.cuda.current_stream().synchronize()  # the slot is recycled on release
        mb.release()
    rank_nodes = torch.empty(num_node, dtype=torch.int32, device=sampler.device)
    rank_freq = torch.empty(num_node, dtype=torch.int32, device=sampler.device)
    check(lib.xg_presample_rank(_ptr(freq), num_node, _ptr(rank_nodes), _ptr(rank_freq), _stream()))
    return rank_nodes, rank_freq


def build_replica_words(cache_table: torch.Tensor, num_replicated: int) -> torch.Tensor:
    """rank/select words of the replicated tier (xg_build_replica_words): int32[ceil(V / 32), 2] = {bits, prefix}.
    Raises when the replica is not laid out by node id."""
    V = cache_table.numel()
    words = torch.empty(((V + 31) // 32, 2), dtype=torch.int32, device=cache_table.device)
    cover = ctypes.c_uint32(0)
    check(lib.xg_build_replica_words(_ptr(cache_table), V, num_replicated, _ptr(words), ctypes.addressof(cover), _stream()))
    words.cover_all = bool(cover.value)  # every node cached + sharded tier by node id: the table is never read
    return words


def sanity_check_ids(ids: torch.Tensor, num_node: int, epoch_map: torch.Tensor | None = None) -> int:
    """GPUBatchSanityCheck / GPUSanityCheckList (cuda/cuda_sanity_check.cu:28-88) as a count instead of a device assert:
    ids equal to kEmptyKey, ids >= num_node and - with an `epoch_map` (int32[num_node], zero at the start of an epoch,
    updated in place) - ids already seen this epoch."""
    bad = torch.zeros(1, dtype=torch.int64, device=ids.device)
    check(lib.xg_sanity_check_ids(_ptr(ids), ids.numel(), num_node, _ptr(epoch_map) if epoch_map is not None else None,
                                  _ptr(bad), _stream()))
    return int(bad.item())


def sanity_check_edges(row: torch.Tensor, col: torch.Tensor, num_src: int, num_dst: int) -> int:
    """number of edges whose local endpoints fall outside their block (row >= num_src or col >= num_dst)"""
    assert row.numel() == col.numel()
    bad = torch.zeros(1, dtype=torch.int64, device=row.device)
    check(lib.xg_sanity_check_edges(_ptr(row), _ptr(col), row.numel(), num_src, num_dst, _ptr(bad), _stream()))
    return int(bad.item())


def predict_num_nodes(batch_size: int, fanout) -> int:
    arr = (c_uint32 * len(fanout))(*[int(f) for f in fanout])
    return int(lib.xg_predict_num_nodes(batch_size, arr, len(fanout)))
