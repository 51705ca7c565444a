// Device-side view of the ordered hash table (unique + remap).
//
// Replaces DeviceOrderedHashTable / MutableDeviceOrderedHashTable
// (reference: samgraph/common/cuda/cuda_hashtable.h:35-101, cuda_hashtable.cu:36-135).
//
// Layout in HBM (all per table, sized once for the worst-case batch):
//   slots : uint2 {key, val}[capacity]   capacity = 1.25 x max_items by default (XG_TABLE_SLOTS_PCT): the
//                                        worst-case batch of the headline config (8.45 M ids) gives an
//                                        84 MB table that stays inside B200's 126 MB L2 while a batch
//                                        is being sampled; 0xFFFFFFFF'FFFFFFFF = free
//   n2o   : uint32[max_items]            local id -> global id  (the "unique" list, append only)
//   items : uint32[kMaxFills + 1]        items[k] = number of unique ids after the k-th fill since
//                                        Reset (items[0] = 0); lives on the device so that a whole
//                                        minibatch can be enqueued without a host round trip.
//
// Determinism: while a fill is in flight `val` of a NEW key holds 0x80000000 | (smallest input
// position that carries the key) -- maintained with atomicMin -- so the occurrence that numbers
// the key is the FIRST in input order, not whichever CAS lands first as in the reference
// (cuda_hashtable.cu:54-79).  Keys from earlier fills hold their local id (< 2^31), which atomicMin
// against a tagged position can never change.
#pragma once
#include "common.cuh"

namespace xg {

constexpr int kMaxFills = 62;
constexpr uint32_t kPosTag = 0x80000000u;

struct TableView {
  uint2 *slots;
  uint32_t *n2o;
  uint32_t *n2o_slot;  // [max_items] slot position of every numbered key: lets Reset clear only those
  uint32_t *items;     // [kMaxFills + 1]
  uint32_t *overflow;  // set to 1 if capacity was exceeded
  uint32_t cap;        // number of slots (any value, multiple of 2; not a power of two: sized to sit in L2)
  uint32_t n2o_cap;
};

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t table_hash(uint32_t k) {
  // murmur3 finaliser: full avalanche so that consecutive node ids do not form probe clusters
  k ^= k >> 16; k *= 0x85EBCA6Bu; k ^= k >> 13; k *= 0xC2B2AE35u; k ^= k >> 16;
  return k;
}

// Claim (or find) the slot of `key`; fold `tagged_pos` into its val with atomicMin.
// Returns the slot position, or kEmpty if the table is full.
__device__ __forceinline__ uint32_t table_insert(const TableView &t, uint32_t key, uint32_t tagged_pos) {
  uint32_t p = __umulhi(table_hash(key), t.cap);  // multiply-shift: any capacity, uses the hash's high bits
  for (uint32_t probes = 0; probes < t.cap; ++probes) {
    uint32_t *kp = &t.slots[p].x;
    uint32_t k = *reinterpret_cast<volatile uint32_t *>(kp);
    if (k == kEmpty) k = atomicCAS(kp, kEmpty, key);
    if (k == kEmpty || k == key) {
      atomicMin(&t.slots[p].y, tagged_pos);
      return p;
    }
    p = (p + 1 == t.cap) ? 0u : p + 1;
  }
  *t.overflow = 1;
  return kEmpty;
}

// The same insert split in two so that a thread can have several home-slot operations in flight before it
// resolves any of them (sample_uniform_kernel pipelines 4 edges per lane this way on the small, latency-bound
// layers).  Here the home slot is claimed with ONE 64-bit CAS on the whole {key, val} word: a free slot takes the
// key and the position in a single L2 round trip; only a key that is already there needs a second operation, a
// fire-and-forget RED.min, skipped when the val that came back is already smaller (val only decreases during a
// fill).  Measured on B200 (same box, papers100M shape): layers of 8 k / 128 k nodes 0.045 -> 0.037 ms and
// 0.080 -> 0.075 ms; on the big layer (0.9 M nodes, bound by random-access throughput rather than latency) the
// CAS-on-every-probe form is SLOWER than load-then-CAS (0.254 vs 0.229 ms), so table_insert above keeps that.
constexpr unsigned long long kFreeSlot = ~0ull;
__device__ __forceinline__ unsigned long long slot_word(uint32_t key, uint32_t val) {
  return ((unsigned long long)val << 32) | key;  // uint2 {x = key, y = val}, little endian
}
__device__ __forceinline__ uint32_t table_insert_begin(const TableView &t, uint32_t key, uint32_t tagged_pos,
                                                       unsigned long long *old_out) {
  const uint32_t p = __umulhi(table_hash(key), t.cap);
  *old_out = atomicCAS(reinterpret_cast<unsigned long long *>(&t.slots[p]), kFreeSlot, slot_word(key, tagged_pos));
  return p;
}
__device__ __forceinline__ uint32_t table_insert_finish(const TableView &t, uint32_t key, uint32_t tagged_pos,
                                                        uint32_t p, unsigned long long old) {
  if (old == kFreeSlot) return p;
  if ((uint32_t)old == key) {
    if ((uint32_t)(old >> 32) > tagged_pos) atomicMin(&t.slots[p].y, tagged_pos);
    return p;
  }
  // the home slot belongs to another key: linear probing from the next slot, looking before claiming
  p = (p + 1 == t.cap) ? 0u : p + 1;
  for (uint32_t probes = 1; probes < t.cap; ++probes) {
    uint32_t *kp = &t.slots[p].x;
    uint32_t k = *reinterpret_cast<volatile uint32_t *>(kp);
    if (k == kEmpty) k = atomicCAS(kp, kEmpty, key);
    if (k == kEmpty || k == key) {
      atomicMin(&t.slots[p].y, tagged_pos);
      return p;
    }
    p = (p + 1 == t.cap) ? 0u : p + 1;
  }
  *t.overflow = 1;
  return kEmpty;
}

// SearchO2N: slot position of `key`, kEmpty when absent.
__device__ __forceinline__ uint32_t table_find(const TableView &t, uint32_t key) {
  uint32_t p = __umulhi(table_hash(key), t.cap);
  for (uint32_t probes = 0; probes < t.cap; ++probes) {
    uint32_t k = __ldg(&t.slots[p].x);
    if (k == key) return p;
    if (k == kEmpty) return kEmpty;
    p = (p + 1 == t.cap) ? 0u : p + 1;
  }
  return kEmpty;
}
#endif

// host-side launchers shared by hashtable.cu and the minibatch pipeline
struct ScanSpace {      // zeroed before every launch that uses it
  uint32_t *ticket;     // 1 word (padded to 8 B)
  uint64_t *state;      // one word per tile
  size_t max_tiles;
};

constexpr int kAssignBlock = 256;
constexpr int kAssignItems = 4;
constexpr int kAssignTile = kAssignBlock * kAssignItems;

// insert keys[0..n) with positions 0..n-1; pos_out[i] = slot position. n from *d_n if d_n != NULL.
int launch_table_insert(const TableView &t, const uint32_t *keys, uint32_t *pos_out, size_t n_max,
                        const uint64_t *d_n, cudaStream_t s);
// number the first occurrences: items[fill] = items[fill-1] + #new; n2o appended; slot.val = local
// local_out != NULL: also write local_out[i] = local id of the key at position i (the remap, fused: no second pass)
int launch_table_assign(const TableView &t, const uint32_t *pos, size_t n_max, const uint64_t *d_n,
                        int fill, const ScanSpace &scan, cudaStream_t s, uint32_t *local_out = nullptr);
// Reset in O(items): slots[n2o_slot[i]] = free for i < items[fill], then items[] = 0 (stream-ordered)
int launch_table_clear(const TableView &t, int fill, size_t max_items, cudaStream_t s);
// out[i] = slots[pos[i]].val
int launch_table_gather_local(const TableView &t, const uint32_t *pos, uint32_t *out, size_t n_max,
                              const uint64_t *d_n, cudaStream_t s);

}  // namespace xg
