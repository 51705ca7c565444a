"""numpy Philox4x32-10 (Salmon et al., SC'11) -- an independent twin of csrc/common.cuh's generator, used
to replay a kernel's random words through the CPU oracle.  TEST INFRASTRUCTURE."""
import numpy as np

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32(key_lo, key_hi, c0, c1, c2, c3):
    """All arguments broadcast; returns uint32 array [..., 4]."""
    c0, c1, c2, c3 = [np.asarray(x, dtype=np.uint64) & MASK for x in np.broadcast_arrays(c0, c1, c2, c3)]
    k0 = np.uint64(key_lo & 0xFFFFFFFF)
    k1 = np.uint64(key_hi & 0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(M0) * c0
        p1 = np.uint64(M1) * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & MASK, lo1, (hi0 ^ c3 ^ k1) & MASK, lo0
        k0 = (k0 + np.uint64(W0)) & MASK
        k1 = (k1 + np.uint64(W1)) & MASK
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def batch_key(seed: int, key: int):
    """philox_key(seed, batch_key) of csrc/sampling.cuh -> (lo, hi)."""
    k64 = (seed ^ ((key * 0x9E3779B97F4A7C15 + 0x7F4A7C15F39CC060) & 0xFFFFFFFFFFFFFFFF)) & 0xFFFFFFFFFFFFFFFF
    return k64 & 0xFFFFFFFF, k64 >> 32


SALT = 0x5A4D4752  # kSampleSalt


def words_slots(seed, key, num_input, fanout, c3=0, salt=SALT):
    """[num_input * fanout, 4]: counter (input position, slot, salt, c3) -- KHop1 / weighted / hash-dedup tries"""
    lo, hi = batch_key(seed, key)
    i = np.repeat(np.arange(num_input, dtype=np.uint64), fanout)
    s = np.tile(np.arange(fanout, dtype=np.uint64), num_input)
    return np.ascontiguousarray(philox4x32(lo, hi, i, s, salt, c3).reshape(-1, 4))


def words_walks(seed, key, num_input, num_walk, walk_len, salt=SALT):
    """[((i * W + w) * L + t), 4]: counter (input position, walk, step, salt)"""
    lo, hi = batch_key(seed, key)
    i, w, t = np.meshgrid(np.arange(num_input, dtype=np.uint64), np.arange(num_walk, dtype=np.uint64),
                          np.arange(walk_len, dtype=np.uint64), indexing="ij")
    return np.ascontiguousarray(philox4x32(lo, hi, i.ravel(), w.ravel(), t.ravel(), salt).reshape(-1, 4))
