"""Synthetic GGMS stores of the BASELINE.json shapes, built directly in HBM.

Host-side mirror of what DistGraph::GraphLoad / FeatureLoad and the GPUCacheManager constructor leave
behind (cuda/dist_graph.cu:309-547, cuda/cuda_cache_manager_host.cc:133-254): per-GPU topology
partitions (node v -> part v % P, row v / P), feature-cache partitions (cache rank r -> part r % P,
row r / P), and the direct-address cache table u32[V].  The data itself comes from the device-side
generator (csrc/synth.cu); torch only owns the memory.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch

from . import ops
from ._clib import check, lib

# name -> (num_node, target num_edge, feat dim, feat dtype, num_class, train fraction)
SHAPES = {
    "products": (2_449_029, 61_859_140, 100, torch.float32, 47, 0.08),
    "papers100M": (111_059_956, 1_615_685_872, 128, torch.float32, 172, 0.0109),
    "twitter": (41_652_230, 1_468_365_182, 256, torch.float32, 150, 0.01),
    "friendster": (65_608_366, 3_612_134_270, 128, torch.float32, 100, 0.01),
    "mag240M-homo": (244_160_499, 1_728_364_232, 768, torch.float16, 153, 0.0045),
}
SYNTH_SEED = 0x5A4D47524150


def degree_constant(num_node: int, num_edge: int) -> float:
    """c of degree(v) = floor(c * sqrt(V / (v + 1))): sum over v ~= 2 c V - V/2 (floor loses ~1/2 each)."""
    return (num_edge + 0.5 * num_node) / (2.0 * num_node)


def train_set(num_node: int, count: int) -> np.ndarray:
    """`count` distinct node ids spread over the whole id range (multiplicative walk, gcd(K, V) = 1)."""
    K = 2654435761  # prime
    assert num_node % K != 0
    i = np.arange(count, dtype=np.uint64)
    return ((i * np.uint64(K) + np.uint64(12345)) % np.uint64(num_node)).astype(np.uint32)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def gen_degrees(begin: int, stride: int, n: int, num_node: int, c: float, device, max_deg: int = 0xFFFFFFFF):
    deg = torch.empty(n, dtype=torch.int32, device=device)
    check(lib.xg_synth_degrees(deg.data_ptr(), begin, stride, n, num_node, c, max_deg, _stream()))
    return deg


def gen_csr(begin: int, stride: int, n: int, num_node: int, c: float, seed: int, device, max_deg: int = 0xFFFFFFFF):
    """(indptr u32[n+1], indices u32[E]) of nodes begin, begin+stride, ... as int32 views."""
    deg = gen_degrees(begin, stride, n, num_node, c, device, max_deg)
    indptr64 = torch.zeros(n + 1, dtype=torch.int64, device=device)
    torch.cumsum(deg, 0, out=indptr64[1:])
    num_edge = int(indptr64[-1].item())
    assert num_edge < (1 << 32), "edge offsets are 32-bit (IdType)"
    indptr = indptr64.to(torch.int32)  # wraps: the bits are the uint32 offsets
    del indptr64, deg
    indices = torch.empty(max(num_edge, 1), dtype=torch.int32, device=device)
    check(lib.xg_synth_indices(indices.data_ptr(), indptr.data_ptr(), begin, stride, n, num_node, seed, _stream()))
    return indptr, indices, num_edge


def gen_features(n_rows: int, dim: int, dtype, device, *, row_ids=None, begin: int = 0, stride: int = 1, out=None):
    if out is None:
        out = torch.empty((n_rows, dim), dtype=dtype, device=device)
    code = 0 if dtype == torch.float32 else 2
    check(lib.xg_synth_features(out.data_ptr(), None if row_ids is None else row_ids.data_ptr(), begin, stride,
                                n_rows, dim, code, _stream()))
    return out


@dataclass
class GGMSStore:
    """What one GPU holds after sample_init + train_init (dist_engine.cc:198-244, 589-700)."""
    num_node: int
    num_edge: int
    dim: int
    dtype: torch.dtype
    num_class: int
    rank: int
    world: int
    graph: object = None           # XgGraph handle (by-value pointers: local + peer parts)
    store: object = None           # XgFeatureStore handle
    label: torch.Tensor = None     # int64[V] on the device
    train: np.ndarray = None       # whole train set (host, uint32)
    local: dict = field(default_factory=dict)  # tensors this GPU owns

    @property
    def row_bytes(self) -> int:
        return self.dim * torch.empty((), dtype=self.dtype).element_size()


def cache_rank_nodes(num_node: int, num_cached: int, num_replicated: int = 0) -> np.ndarray:
    """The cached set in cache-rank order: first `num_cached` entries of the degree ranking (for the
    synthetic shapes degree is non-increasing in the id, so the stable ranking is the identity),
    shuffled as GPUCacheManager does (std::shuffle with mt19937(num_cached),
    cuda_cache_manager_host.cc:162-171) so that rank r -> part r % P is load balanced."""
    nodes = np.arange(num_cached, dtype=np.uint32)
    # two-tier store: the replicated tier keeps the policy order (it is the HOTTEST prefix and needs no
    # balancing); only the partitioned remainder is shuffled, by the same rule applied to what is left
    rest = nodes[num_replicated:]
    if len(rest):
        check(lib.xg_shuffle_cache_ranks(rest.ctypes.data, len(rest)))
    return nodes


def build_store(shape: str, device: int, rank: int = 0, world: int = 1, *, scale: int = 1, cache_pct: float = 1.0,
                replicate_pct: float = 0.0, with_features: bool = True, seed: int = SYNTH_SEED,
                flat_when_full: bool = True, replica_words: bool = True, replicate_topology: bool = False,
                empty_feat: int = 20, exportable: bool = False) -> GGMSStore:
    """Generate THIS rank's shards (topology part `rank`, feature part `rank`) of the named shape,
    scaled down by `scale`. Peer parts are attached afterwards (attach_peers) from IPC handles.
    replicate_topology: every GPU keeps the whole CSR (B200 option: papers100M's 6.9 GB fits everywhere, so the sample
    phase needs no NVLink; the reference's `v % P` sharding stays the default and is what config 3 asks for).
    cache_pct < 1: the uncached rows live in pinned host memory; the host tier is a SAMGRAPH_EMPTY_FEAT-style mock table
    of 2^empty_feat rows (node v reads row v & mask, cuda/cuda_extraction.cu:51-72), as the reference's evaluation
    scripts run shapes whose feature file does not fit - the PCIe traffic is real, the host footprint small.
    exportable: the feature shard is generated straight into an xg_device_malloc block (its own cudaMalloc, 2 MiB granular:
    what a CUDA IPC handle can name, dist_graph.cu:274-307) instead of torch's pool, so that sharing it needs no second copy
    of a shard that may fill half the GPU; the raw pointer is kept in st.local["raw"]."""
    V, E, dim, dtype, num_class, train_frac = SHAPES[shape]
    V, E = V // scale, E // scale
    dev = torch.device("cuda", device)
    c = degree_constant(V, E)
    P = world
    st = GGMSStore(num_node=V, num_edge=0, dim=dim, dtype=dtype, num_class=num_class, rank=rank, world=world)
    with torch.cuda.device(dev):
        # topology: use_dist_graph = 1.0 -> num_cache_node = V, every node lives in some GPU part
        if replicate_topology:
            indptr, indices, ne = gen_csr(0, 1, V, V, c, seed, dev)
        else:
            n_part = (V - rank + P - 1) // P
            indptr, indices, ne = gen_csr(rank, P, n_part, V, c, seed, dev)
        st.local.update(indptr=indptr, indices=indices, num_local_edge=ne, replicate_topology=replicate_topology)
        st.label = (torch.arange(V, dtype=torch.int64, device=dev) % num_class)
        if with_features and P == 1 and cache_pct >= 1.0 and flat_when_full:
            # one part holding every row: WHERE a cached row is stored is ours to choose ("GGMS routing must
            # match in which row is read, not in where it is stored"), so the single part is laid out by node id and
            # the direct-address table (4 B x V, one 128-byte DRAM line per random lookup) is not needed at all
            st.local["feat_part"] = gen_features(V, dim, dtype, dev, begin=0)
            st.local["num_cached"] = V
            st.local["flat"] = True
        elif with_features:
            n_cached = int(V * cache_pct)
            n_rep = min(n_cached, int(V * replicate_pct)) if (P > 1 or not flat_when_full) else 0
            ranks = cache_rank_nodes(V, n_cached, n_rep)
            ranks[:n_rep].sort()  # the replicated tier is stored by node id: its rows resolve by rank/select (replica_words)
            if n_rep and replica_words and n_cached == V:
                # everything cached: the sharded tier by node id as well (round-robin over the parts in id order is as
                # balanced as the shuffle) - then rank/select covers every node and the DRAM table is never read
                ranks[n_rep:].sort()
            table = np.full(V, 0xFFFFFFFF, np.uint32)
            table[ranks] = np.arange(n_cached, dtype=np.uint32)
            st.local["cache_table"] = torch.from_numpy(table.view(np.int32)).to(dev)
            mine = torch.from_numpy(np.ascontiguousarray(ranks[n_rep:][rank::P]).view(np.int32)).to(dev)
            part_out = None
            if exportable:
                import ctypes
                raw = ctypes.c_void_p()
                rows = max(mine.numel(), 1)
                check(lib.xg_device_malloc(ctypes.byref(raw), rows * st.row_bytes))
                st.local.setdefault("raw", {})["feat_part"] = raw.value
                part_out = torch.as_tensor(ops._DevMem(raw.value, (rows, dim), ops._TYPESTR[dtype], None), device=dev)
            st.local["feat_part"] = gen_features(max(mine.numel(), 1), dim, dtype, dev, row_ids=mine if mine.numel() else None,
                                                 out=part_out)
            st.local["num_cached"] = n_cached
            st.local["num_replicated"] = n_rep
            if n_cached < V:  # cold tier: pinned host memory, read over PCIe through UVA
                rows = 1 << empty_feat
                host = torch.empty((rows, dim), dtype=dtype).pin_memory()
                step = max(1, (256 << 20) // st.row_bytes)
                for o in range(0, rows, step):
                    n = min(step, rows - o)
                    host[o:o + n].copy_(gen_features(n, dim, dtype, dev, begin=o))
                st.local["host_feat"] = host
                st.local["empty_feat"] = empty_feat
            if n_rep:  # ranks [0, n_rep) are nodes 0..n_rep-1 in policy order: every GPU generates its own copy
                hot = torch.from_numpy(np.ascontiguousarray(ranks[:n_rep]).view(np.int32)).to(dev)
                st.local["replica"] = gen_features(n_rep, dim, dtype, dev, row_ids=hot)
                del hot
                if replica_words:
                    st.local["replica_words"] = ops.build_replica_words(st.local["cache_table"], n_rep)
            del mine
        torch.cuda.synchronize()
    st.train = train_set(V, max(1, int(V * train_frac)))
    return st


def attach(st: GGMSStore, part_indptr=None, part_indices=None, feat_parts=None, host_feat=None) -> GGMSStore:
    """Fill the by-value handles. With world == 1 the local shards are the whole store; otherwise the
    caller passes every part's pointer (local tensor or IPC-mapped peer address) in part order."""
    P = st.world
    if st.local.get("replicate_topology") or part_indptr is None:
        assert P == 1 or st.local.get("replicate_topology")
        part_indptr, part_indices = [st.local["indptr"]], [st.local["indices"]]
    if feat_parts is None:
        assert P == 1
        feat_parts = [st.local.get("feat_part")]
    if host_feat is None:
        host_feat = st.local.get("host_feat")
    st.graph = ops.make_graph(part_indptr=part_indptr, part_indices=part_indices, num_cache_node=st.num_node,
                              num_node=st.num_node, host_indptr=None, host_indices=None)
    if st.local.get("flat"):
        st.store = ops.make_feature_store(st.row_bytes, st.local["feat_part"], None)
    elif st.local.get("cache_table") is not None:
        st.store = ops.make_feature_store(st.row_bytes, host_feat, st.local["cache_table"], parts=feat_parts,
                                          local_parts=(st.rank,), replica=st.local.get("replica"),
                                          num_replicated=st.local.get("num_replicated", 0),
                                          replica_words=st.local.get("replica_words"),
                                          empty_feat=st.local.get("empty_feat", 0) if host_feat is not None else 0)
    return st
