"""Seeded synthetic graphs of the BASELINE.json shapes (numpy, for tests and CPU-side checks).

Degree sequence: discrete power law clipped to [0, max_deg] and rescaled so that sum(deg) ~= E;
neighbour lists are sorted, ids uniform over V.  Features are a counter hash
of (row, col) so any row can be re-derived without a second copy; labels = node % num_class.
"""
from __future__ import annotations

import numpy as np

SEED = 0x5A4D47524150


def power_law_degrees(rng, V, E, alpha=2.1, max_deg=None):
    max_deg = max_deg or max(16, int(V ** 0.5) * 4)
    u = rng.random(V)
    d = np.floor((1.0 - u) ** (-1.0 / (alpha - 1.0))).astype(np.float64)
    d = np.minimum(d, max_deg)
    d *= E / d.sum()
    deg = np.floor(d).astype(np.int64)
    # a share of isolated / low-degree nodes keeps the deg <= fanout branch exercised
    deg[rng.random(V) < 0.02] = 0
    return np.minimum(deg, max_deg)


def make_graph(V, E, seed=SEED, alpha=2.1, max_deg=None):
    rng = np.random.default_rng(seed)
    deg = power_law_degrees(rng, V, E, alpha, max_deg)
    indptr = np.zeros(V + 1, np.uint32)
    np.cumsum(deg, out=indptr[1:])
    total = int(indptr[-1])
    indices = rng.integers(0, V, size=total, dtype=np.uint32)
    # sort inside each list
    order = np.lexsort((indices, np.repeat(np.arange(V, dtype=np.uint32), deg)))
    indices = indices[order]
    return indptr, indices


def feature_rows(rows: np.ndarray, dim: int, dtype=np.float32) -> np.ndarray:
    """counter hash of (row, col) -> exact values representable in fp16/fp32."""
    r = rows.astype(np.uint64)[:, None]
    c = np.arange(dim, dtype=np.uint64)[None, :]
    h = (r * np.uint64(0x9E3779B97F4A7C15) + c * np.uint64(0xD1B54A32D192ED03)) >> np.uint64(40)
    v = (h % np.uint64(2048)).astype(np.float32) / np.float32(8.0) - np.float32(128.0)
    return v.astype(dtype)


def make_features(V, dim, dtype=np.float32):
    return feature_rows(np.arange(V), dim, dtype)


def make_labels(V, num_class=47):
    return (np.arange(V, dtype=np.int64) % num_class).astype(np.int64)


def train_set(V, frac, seed=SEED):
    rng = np.random.default_rng(seed ^ 0xABCDEF)
    perm = rng.permutation(V).astype(np.uint32)
    return perm[: max(1, int(V * frac))]


def degree_rank(indptr):
    """nodes by out-degree, descending, stable (utility/data-process/toolkit/cache/cache_by_degree.cc:29-63)"""
    deg = np.diff(indptr.astype(np.int64))
    return np.argsort(-deg, kind="stable").astype(np.uint32)


def edge_weights(indptr, indices, seed=11):
    """positive per-edge weights (seeded); a few zero-probability edges exercise the alias fallback"""
    rng = np.random.default_rng(seed)
    w = rng.random(int(indptr[-1])).astype(np.float64) + 0.05
    w[rng.random(len(w)) < 0.05] *= 20.0
    return w


def alias_tables(indptr, indices, weights):
    """Vose alias tables per node, laid out like prob_table.bin / alias_table.bin
    (utility/data-process/toolkit/weight/create_alias_table.cc): prob f32[E], alias u32[E] holding NODE IDS."""
    E = int(indptr[-1])
    prob = np.ones(E, np.float32)
    alias = indices.copy().astype(np.uint32)
    for v in range(len(indptr) - 1):
        b, e = int(indptr[v]), int(indptr[v + 1])
        n = e - b
        if n == 0:
            continue
        p = weights[b:e] / weights[b:e].sum() * n
        small = [i for i in range(n) if p[i] < 1.0]
        large = [i for i in range(n) if p[i] >= 1.0]
        while small and large:
            s, l = small.pop(), large.pop()
            prob[b + s] = p[s]
            alias[b + s] = indices[b + l]
            p[l] = p[l] + p[s] - 1.0
            (small if p[l] < 1.0 else large).append(l)
    return prob, alias


def prefix_table(indptr, weights):
    """per-node running sums of normalised weights (create_prob_prefix_table.cc:94-131), f32[E]"""
    out = np.zeros(int(indptr[-1]), np.float32)
    for v in range(len(indptr) - 1):
        b, e = int(indptr[v]), int(indptr[v + 1])
        if e > b:
            out[b:e] = np.cumsum(weights[b:e] / weights[b:e].sum()).astype(np.float32)
    return out


def write_dataset(path, indptr, indices, feat, label, train, num_class=47, valid=None, test=None, weights=None,
                  feat_dtype_name=None):
    """the reference's on-disk format (samgraph/common/engine.cc:109-443, constant.cc:23-51, datagen/README.md)"""
    import os
    os.makedirs(path, exist_ok=True)
    valid = np.zeros(0, np.uint32) if valid is None else valid
    test = np.zeros(0, np.uint32) if test is None else test
    meta = {"NUM_NODE": len(indptr) - 1, "NUM_EDGE": int(indptr[-1]), "FEAT_DIM": feat.shape[1], "NUM_CLASS": num_class,
            "NUM_TRAIN_SET": len(train), "NUM_VALID_SET": len(valid), "NUM_TEST_SET": len(test)}
    with open(os.path.join(path, "meta.txt"), "w") as f:
        for k, v in meta.items():
            f.write(f"{k} {v}\n")
        if feat_dtype_name:
            f.write(f"FEAT_DATA_TYPE {feat_dtype_name}\n")
    indptr.astype(np.uint32).tofile(os.path.join(path, "indptr.bin"))
    indices.astype(np.uint32).tofile(os.path.join(path, "indices.bin"))
    np.ascontiguousarray(feat).tofile(os.path.join(path, "feat.bin"))
    label.astype(np.uint64).tofile(os.path.join(path, "label.bin"))
    train.astype(np.uint32).tofile(os.path.join(path, "train_set.bin"))
    valid.astype(np.uint32).tofile(os.path.join(path, "valid_set.bin"))
    test.astype(np.uint32).tofile(os.path.join(path, "test_set.bin"))
    degree_rank(indptr).tofile(os.path.join(path, "cache_by_degree.bin"))
    if weights is not None:
        prob, alias = alias_tables(indptr, indices, weights)
        prob.tofile(os.path.join(path, "prob_table.bin"))
        alias.tofile(os.path.join(path, "alias_table.bin"))
        prefix_table(indptr, weights).tofile(os.path.join(path, "prob_prefix_table.bin"))


def minstd_shuffle_epoch(data: np.ndarray, epoch: int) -> np.ndarray:
    """Pure-Python restatement of DistAlignedShuffler::ReShuffle (dist/dist_shuffler_aligned.cc:86-109) on top of
    libstdc++'s std::default_random_engine (= minstd_rand0: x <- 16807 x mod 2^31-1, seed 0 -> 1) and its
    std::uniform_int_distribution downscaling (bits/uniform_int_dist.h: reject above uerange * scaling, divide)."""
    M = 2147483647
    x = epoch % M or 1
    v = data.copy()
    n = len(v)
    urng_range = (M - 1) - 1  # max() - min() = 2147483646 - 1
    for i in range(n - 1):
        uerange = (n - 1 - i) + 1
        assert uerange - 1 < urng_range
        scaling = urng_range // uerange
        past = uerange * scaling
        while True:
            x = (x * 16807) % M
            ret = x - 1
            if ret < past:
                break
        j = i + ret // scaling
        v[i], v[j] = v[j], v[i]
    return v
