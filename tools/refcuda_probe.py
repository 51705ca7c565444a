"""Debug probe (GPU): the reference's curand states cold (fresh wall-clock seeds) vs warmed up (n calls discarded)."""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np
import oracle_lib as O
import test_ref_cuda as T
import synth
from stats import one_sample_chi2, two_sample_chi2

deg = np.concatenate([np.repeat(T.DEGS, T.M), np.zeros(max(T.DEGS), np.int64)])
indptr = np.zeros(T.V + 1, np.uint32); np.cumsum(deg, out=indptr[1:])
indices = np.concatenate([T.T0 + np.arange(d) for d in np.repeat(T.DEGS, T.M)]).astype(np.uint32)
pos = indices.astype(np.int64) - T.T0
w = 0.2 + ((pos * 7) % 5).astype(np.float64) + (pos % 3 == 0) * 3.0
prob, alias = synth.alias_tables(indptr, indices, w)
prefix = synth.prefix_table(indptr, w)
rg = O.RefCudaGraph(indptr, indices, prob, alias, prefix)
hubs = np.random.default_rng(3).permutation(T.T0).astype(np.uint32)
for warm in (0, 1, 2, 4, 8, 16, 64):
    rg.reset_states()
    for _ in range(warm):
        rg.sample(4, hubs, 1); rg.sample(1, hubs, 6)
    t = np.zeros((len(T.DEGS), 40), np.int64)
    h = np.zeros((len(T.DEGS), 7), np.int64)
    for r in range(20):
        s, d, _ = rg.sample(4, hubs, 1)
        t += T._draw_table(s, d)
        s, d, _ = rg.sample(1, hubs, 6)
        h += T._edges_per_node_hist(s, 6)
    c = 5; dd = 16; b = int(indptr[c * T.M]); p = w[b:b + dd] / w[b:b + dd].sum()
    print("warm", warm, "prefix fanout1 deg16 p vs exact %.2e" % one_sample_chi2(t[c, :dd], p * t[c, :dd].sum())[2],
          "khop1 f6 deg2 hist", h[2].tolist(), "deg16", h[5].tolist())
