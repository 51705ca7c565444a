"""Size-independent properties of one minibatch, checked on the device against the generator's closed forms.

Used by the full-size GPU tests (tests/test_gpu_fullsize.py) and by bench.py's untimed per-rank self-check - at sizes
where the CPU oracle cannot replay a batch in seconds, and at N > 1 where the topology and the feature rows are read
from NVLink peers:
  * unique/remap: input_nodes duplicate-free, seeds are its prefix in order (prefix property), every
    layer's destination list is the previous layer's source list, all local ids in range;
  * sampling (uniform samplers): per-node edge count == min(degree, fanout) for EVERY node of every layer, edges
    node-major in input order, CSC offsets == the scan of those counts, and for every node with degree <= fanout the
    remapped neighbours equal the CSR entries at the same offsets (sampler + unique + remap at full scale);
    a sample of the remaining edges is checked for membership in the CSR list of its source;
  * features: every row equals the generator's closed form for its node id - 0 ULP, byte compare - whichever tier
    (local HBM, NVLink peer, pinned host) served it; labels likewise.
No oracle and no CPU path is involved: the graph, the features and the labels are pure functions of the node id
(csrc/synth.cu), so the expected values are recomputed on the GPU.
"""
from __future__ import annotations

import numpy as np
import torch


def _u32(t):  # int32 view -> int64 values of the uint32 bits
    return t.to(torch.int64) & 0xFFFFFFFF


class Topology:
    """degree / neighbour lookups over a flat CSR or over v % P parts (local tensors or IPC-mapped peer memory)"""

    def __init__(self, part_indptr, part_indices):
        self.ip = [_u32(t) for t in part_indptr]
        self.ix = part_indices
        self.P = len(part_indptr)

    @classmethod
    def of_store(cls, st, ptrs=None):
        from . import ops
        if ptrs is None or st.local.get("replicate_topology") or st.world == 1:
            return cls([st.local["indptr"]], [st.local["indices"]])
        P, V = st.world, st.num_node
        ips, ixs = [], []
        for p in range(P):
            rows = (V - p + P - 1) // P
            ip = ptrs["indptr"][p]
            t_ip = ip if isinstance(ip, torch.Tensor) else torch.as_tensor(ops._DevMem(ip, (rows + 1,), "<i4", st), device="cuda")
            ne = int(_u32(t_ip[-1:]).item())
            ix = ptrs["indices"][p]
            t_ix = ix if isinstance(ix, torch.Tensor) else torch.as_tensor(ops._DevMem(ix, (max(ne, 1),), "<i4", st), device="cuda")
            ips.append(t_ip)
            ixs.append(t_ix)
        return cls(ips, ixs)

    def begin_end(self, nodes):
        """CSR range of every node, as (part, begin, end) int64 tensors"""
        part = nodes % self.P
        row = nodes // self.P
        b = torch.empty_like(nodes)
        e = torch.empty_like(nodes)
        for p in range(self.P):
            m = part == p
            if bool(m.any()):
                r = row[m]
                b[m] = self.ip[p][r]
                e[m] = self.ip[p][r + 1]
        return part, b, e

    def entries(self, part, pos):
        """indices[part][pos] for per-element (part, pos)"""
        out = torch.empty_like(pos)
        for p in range(self.P):
            m = part == p
            if bool(m.any()):
                out[m] = _u32(self.ix[p][pos[m]])
        return out


def expected_feature_rows(st, input_nodes):
    """node id whose generator row each gathered row must equal: the node itself, or - for host-resident nodes of a
    SAMGRAPH_EMPTY_FEAT-style mock host tier - node & (2^k - 1)"""
    k = st.local.get("empty_feat", 0) if st.local.get("host_feat") is not None else 0
    if not k:
        return input_nodes
    table = st.local["cache_table"]
    ids = _u32(input_nodes)
    cold = _u32(table[ids]) == 0xFFFFFFFF
    return torch.where(cold, ids & ((1 << k) - 1), ids).to(torch.int32)


def check_minibatch_properties(mb, st, D, seeds_np, fanout, uniform=True, topo: Topology | None = None,
                               membership_samples: int = 400):
    dev = mb.input_nodes.device
    L = mb.num_layer
    topo = topo or Topology.of_store(st)
    inp = _u32(mb.input_nodes)
    seeds = torch.from_numpy(np.asarray(seeds_np).astype(np.int64)).to(dev)
    # ---- unique / remap ----
    assert inp.numel() == mb.num_input
    assert torch.unique(inp).numel() == inp.numel(), "input_nodes must be duplicate-free"
    assert torch.equal(inp[: len(seeds_np)], seeds), "seeds are the prefix of input_nodes, in order"
    assert int(inp.max()) < st.num_node
    assert mb.num_dst[L - 1] == len(seeds_np) and mb.num_src[0] == mb.num_input
    for l in range(L - 1):
        assert mb.num_dst[l] == mb.num_src[l + 1]  # layer l expands exactly the nodes layer l+1 discovered
    for l in range(L):
        row, col = _u32(mb.row(l)), _u32(mb.col(l))
        assert row.numel() == mb.num_edge[l] == col.numel()
        if row.numel() == 0:
            continue
        assert int(row.max()) < mb.num_src[l] and int(col.max()) < mb.num_dst[l]
        nodes = inp[: mb.num_dst[l]]
        part, b, e = topo.begin_end(nodes)
        deg = e - b
        if uniform:
            # node-major, input order; count == min(deg, fanout) for every expanded node
            assert bool((col[1:] >= col[:-1]).all())
            want = torch.clamp(deg, max=fanout[l])
            got = torch.bincount(col, minlength=mb.num_dst[l])
            assert torch.equal(got, want), "per-node edge count != min(degree, fanout)"
            ip = _u32(mb.csc_indptr(l))  # CSC offsets == exclusive scan of the per-node counts
            assert int(ip[0]) == 0 and torch.equal(ip[1:] - ip[:-1], want)
            # nodes with degree <= fanout keep their whole list in CSR order (the bit-exact branch): EVERY such edge's
            # remapped neighbour must be the CSR entry at the same offset - checks sampler + unique + remap at scale
            whole = deg[col] <= fanout[l]
            off = torch.arange(col.numel(), device=dev) - ip[col]
            cw = col[whole]
            want_dst = topo.entries(part[cw], b[cw] + off[whole])
            assert torch.equal(inp[row[whole]], want_dst), "remapped neighbour differs from the CSR entry"
        # membership of a sample of edges in the CSR list of their source node (k-hop samplers only: a random walk's
        # "edge" joins a seed to a node it VISITED, which need not be a neighbour - pass membership_samples = 0)
        if membership_samples <= 0:
            continue
        g = torch.Generator(device="cpu").manual_seed(l)
        pick = torch.randint(0, row.numel(), (membership_samples,), generator=g).to(dev)
        cp = col[pick]
        dn = inp[row[pick]].cpu().tolist()
        pp, bb, ee = part[cp].cpu().tolist(), b[cp].cpu().tolist(), e[cp].cpu().tolist()
        for k in range(len(dn)):
            nbrs = _u32(topo.ix[pp[k]][bb[k]:ee[k]])
            assert bool((nbrs == dn[k]).any()), "sampled edge is not in the graph"
    # ---- features / labels: closed form of the generator ----
    if mb.feat is not None:
        want = D.gen_features(mb.num_input, st.dim, st.dtype, dev, row_ids=expected_feature_rows(st, mb.input_nodes))
        view = torch.int32 if st.dtype == torch.float32 else torch.int16
        assert torch.equal(mb.feat.view(view), want.view(view)), "feature rows differ from the generator (0 ULP bar)"
    assert torch.equal(mb.label, seeds % st.num_class)
