"""Golden fixtures for the two host-side cache-rank routines, generated from the REFERENCE's own code
(cuda/ics22_song_dist_feature.cc:22-167 and dist/pre_sampler.cc:40-165, compiled in place into
oracle/_ref/libxgnn_ref_cuda.so by `make -C oracle refcuda`).  Run only where /root/reference exists:

    python tests/golden/make_golden_rank.py   ->  tests/golden/ref_rank_small.npz
"""
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))
import oracle_lib as O  # noqa: E402

assert O.refcu is not None, "build oracle/_ref first (make -C oracle refcuda)"

ICS22_CASES = [(200, 40, 4, 4, 0.0), (200, 40, 4, 2, 0.5), (300, 25, 8, 4, 0.9), (120, 30, 2, 2, 0.3), (64, 64, 4, 4, 0.0),
               (150, 10, 6, 3, 0.0), (100, 20, 3, 1, 0.5), (97, 5, 8, 8, 0.05)]


def ics22_weight(V, num_dev):
    rng = np.random.default_rng(V * 31 + num_dev)
    indices = np.minimum((rng.pareto(1.2, V * 12) * 3).astype(np.int64), V - 1).astype(np.uint32)  # skewed, many ties
    return indices, np.bincount(indices, minlength=V).astype(np.float64)


def presample_lists(V, num_lists, seed):
    """input-node lists as a sampler produces them: unique inside one batch, skewed across batches"""
    rng = np.random.default_rng(seed)
    p = 1.0 / (1.0 + np.arange(V)) ** 0.7
    p /= p.sum()
    return [rng.choice(V, size=int(rng.integers(0, V // 2)), replace=False, p=p).astype(np.uint32)
            for _ in range(num_lists)]


if __name__ == "__main__":
    out = {}
    for i, (V, n_cached, num_dev, clique, alpha) in enumerate(ICS22_CASES):
        _, w = ics22_weight(V, num_dev)
        ranking, dev_map, idx_map, cached, total = O.ref_ics22_solve(w, n_cached, num_dev, clique, alpha)
        out[f"ics22_{i}_ranking"], out[f"ics22_{i}_dev_map"], out[f"ics22_{i}_idx_map"] = ranking, dev_map, idx_map
        out[f"ics22_{i}_cached"], out[f"ics22_{i}_total"] = cached, np.int64(total)
    for i, (V, n, epochs) in enumerate([(500, 6, 1), (1000, 8, 2), (64, 3, 3)]):
        nodes, freq = O.ref_presample_rank(V, presample_lists(V, n * epochs, 77 + i), epochs)
        out[f"presample_{i}_nodes"], out[f"presample_{i}_freq"] = nodes, freq
    np.savez_compressed(HERE / "ref_rank_small.npz", **out)
    print("wrote", HERE / "ref_rank_small.npz")
