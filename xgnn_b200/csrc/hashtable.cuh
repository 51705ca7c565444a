// Device-side view of the ordered hash table (unique + remap).
//
// Replaces DeviceOrderedHashTable / MutableDeviceOrderedHashTable
// (reference: samgraph/common/cuda/cuda_hashtable.h:35-101, cuda_hashtable.cu:36-135).
//
// Layout in HBM (all per table, sized once for the worst-case batch):
//   slots : uint32[capacity]   ONE 4-byte word per slot.  capacity = 1.5 x max_items by default
//                              (XG_TABLE_SLOTS_PCT): the worst-case batch of the headline config (8.45 M ids) gives a
//                              51 MB table.  The r01 table kept {key, val} pairs (8 bytes, 84 MB): measured on B200, a
//                              randomly probed table stays L2-resident up to ~64 MB (fill cost per key doubles from
//                              64 to 84 MB and triples at 256 MB, tools/table_size_probe.py) and every probe that
//                              misses L2 is a DRAM row activation that the feature gather on the other stream pays
//                              for.  0xFFFFFFFF = free.
//   n2o   : uint32[max_items]  local id -> global id  (the "unique" list, append only)
//   items : uint32[kMaxFills + 1]  items[k] = number of unique ids after the k-th fill since Reset (items[0] = 0);
//                              lives on the device so that a whole minibatch can be enqueued without a host round trip.
//
// What a slot word means:
//   0x80000000 | pos   (during a fill) the smallest input position of the CURRENT fill that carries the slot's key,
//                      maintained with atomicMin - so the occurrence that numbers the key is the FIRST in input
//                      order, not whichever CAS lands first as in the reference (cuda_hashtable.cu:54-79);
//   local id < 2^31    (after the numbering pass, and for keys of earlier fills) - atomicMin against a tagged
//                      position can never change it.
// The KEY is not stored: it is keys[pos] of the current fill or n2o[local], both already in memory (and L2-resident:
// 17 MB + 20 MB for the headline batch).  Only a probe that lands on an occupied slot has to look it up.
#pragma once
#include "common.cuh"

namespace xg {

constexpr int kMaxFills = 62;
constexpr uint32_t kPosTag = 0x80000000u;

struct TableView {
  uint32_t *slots;
  uint32_t *n2o;
  uint32_t *n2o_slot;  // [max_items] slot position of every numbered key: lets Reset clear only those
  uint32_t *items;     // [kMaxFills + 1]
  uint32_t *overflow;  // set to 1 if capacity was exceeded
  uint32_t cap;        // number of slots (any value, multiple of 4; not a power of two: sized to sit in L2)
  uint32_t n2o_cap;
};

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t table_hash(uint32_t k) {
  // murmur3 finaliser: full avalanche so that consecutive node ids do not form probe clusters
  k ^= k >> 16; k *= 0x85EBCA6Bu; k ^= k >> 13; k *= 0xC2B2AE35u; k ^= k >> 16;
  return k;
}
__device__ __forceinline__ uint32_t table_home(const TableView &t, uint32_t key) {
  return __umulhi(table_hash(key), t.cap);  // multiply-shift: any capacity, uses the hash's high bits
}

// key of an occupied slot.  `keys` = input array of the fill in flight (NULL outside a fill: every word is a local id).
// Both arrays are read-only while the kernels that call this run, so plain loads are safe.
__device__ __forceinline__ uint32_t table_key_of(const TableView &t, const uint32_t *__restrict__ keys, uint32_t word) {
  if (word & kPosTag) return keys[word & ~kPosTag];
  return word < t.n2o_cap ? t.n2o[word] : kEmpty;  // >= n2o_cap: the "numbered out of range" mark of an overflowed batch
}

// Claim (or find) the slot of keys[pos] and fold 0x80000000 | pos into it.  Returns the slot position, or kEmpty if the
// table is full.  A free home slot - the common case - costs ONE atomic and nothing else.
__device__ __forceinline__ uint32_t table_insert_resolve(const TableView &t, const uint32_t *__restrict__ keys,
                                                         uint32_t key, uint32_t tagged_pos, uint32_t p, uint32_t old) {
  for (uint32_t probes = 0; probes < t.cap; ++probes) {
    if (probes) old = atomicCAS(&t.slots[p], kEmpty, tagged_pos);
    if (old == kEmpty) return p;
    if (table_key_of(t, keys, old) == key) {
      if (old > tagged_pos) atomicMin(&t.slots[p], tagged_pos);  // only tagged words can be larger; fire and forget
      return p;
    }
    p = (p + 1 == t.cap) ? 0u : p + 1;
  }
  *t.overflow = 1;
  return kEmpty;
}
__device__ __forceinline__ uint32_t table_insert(const TableView &t, const uint32_t *__restrict__ keys, uint32_t key,
                                                 uint32_t tagged_pos) {
  const uint32_t p = table_home(t, key);
  const uint32_t old = atomicCAS(&t.slots[p], kEmpty, tagged_pos);
  return table_insert_resolve(t, keys, key, tagged_pos, p, old);
}

// SearchO2N: slot position of `key`, kEmpty when absent.  Outside a fill every occupied word is a local id.
__device__ __forceinline__ uint32_t table_find(const TableView &t, uint32_t key) {
  uint32_t p = table_home(t, key);
  for (uint32_t probes = 0; probes < t.cap; ++probes) {
    const uint32_t w = __ldg(&t.slots[p]);
    if (w == kEmpty) return kEmpty;
    if (w < t.n2o_cap && t.n2o[w] == key) return p;
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

// insert keys[0..n) with positions 0..n-1; pos_out[i] = slot position (must not alias keys). n from *d_n if d_n != NULL.
int launch_table_insert(const TableView &t, const uint32_t *keys, uint32_t *pos_out, size_t n_max,
                        const uint64_t *d_n, cudaStream_t s);
// number the first occurrences: items[fill] = items[fill-1] + #new; n2o appended; slot word = local id.  `keys` = the
// array the positions were inserted from.
// local_out != NULL (16-byte aligned): also write local_out[i] = local id of the key at position i (the remap, fused)
// ids_into_table = false: the last fill before Reset - the ids are only handed out through local_out / n2o, the
// slots keep their tagged positions (lookups after this fill are not possible)
int launch_table_assign(const TableView &t, const uint32_t *keys, const uint32_t *pos, size_t n_max, const uint64_t *d_n,
                        int fill, const ScanSpace &scan, cudaStream_t s, uint32_t *local_out = nullptr,
                        bool ids_into_table = true);
// Reset in O(items): slots[n2o_slot[i]] = free for i < items[fill], then items[] = 0 (stream-ordered)
int launch_table_clear(const TableView &t, int fill, size_t max_items, cudaStream_t s);
// out[i] = slots[pos[i]] (the local id, after the numbering pass)
int launch_table_gather_local(const TableView &t, const uint32_t *pos, uint32_t *out, size_t n_max,
                              const uint64_t *d_n, cudaStream_t s);

}  // namespace xg
