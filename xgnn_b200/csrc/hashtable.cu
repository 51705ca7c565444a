// Ordered hash table: unique + remap.
//
// Reference behaviour being replaced: cuda::OrderedHashTable (samgraph/common/cuda/cuda_hashtable.cu
// :699-1064) and GPUMapEdges (cuda/cuda_mapping.cu:50-80).  The reference runs three probing passes
// per fill over 16-byte versioned buckets with a host sync after every kernel; here a fill is
//   insert  : one probe per input item, remembers the slot position (4 B/item, streaming)
//   assign  : single-pass chained scan over "is first occurrence" flags -> local ids, n2o append
//   (gather : out[i] = local id, one 8-byte load at the remembered position)
// on 8-byte slots, all stream-ordered with device-resident counts (no host round trip).
#include "hashtable.cuh"

#include <cstdlib>
#include <mutex>
#include <vector>

namespace xg {

namespace {

__device__ __forceinline__ uint64_t resolve_n(size_t n_max, const uint64_t *d_n) {
  uint64_t n = d_n ? *d_n : (uint64_t)n_max;
  return n < n_max ? n : n_max;
}

// kInsertItems keys per thread: all home-slot CASes are issued before any result is consumed, so a thread has that
// many independent L2 atomics in flight (the kernel is a stream of coalesced key loads + one random atomic per key).
constexpr int kInsertItems = 4;
__global__ void __launch_bounds__(256)
table_insert_kernel(TableView t, const uint32_t *__restrict__ keys, uint32_t *__restrict__ pos_out,
                    size_t n_max, const uint64_t *d_n) {
  const uint64_t n = resolve_n(n_max, d_n);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i0 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i0 < n; i0 += stride * kInsertItems) {
    uint32_t key[kInsertItems], p[kInsertItems], old[kInsertItems];
#pragma unroll
    for (int k = 0; k < kInsertItems; ++k) {
      const uint64_t i = i0 + (uint64_t)k * stride;
      key[k] = i < n ? keys[i] : 0u;
    }
#pragma unroll
    for (int k = 0; k < kInsertItems; ++k) {
      const uint64_t i = i0 + (uint64_t)k * stride;
      if (i < n) {
        p[k] = table_home(t, key[k]);
        old[k] = atomicCAS(&t.slots[p[k]], kEmpty, kPosTag | (uint32_t)i);
      }
    }
#pragma unroll
    for (int k = 0; k < kInsertItems; ++k) {
      const uint64_t i = i0 + (uint64_t)k * stride;
      if (i < n) pos_out[i] = table_insert_resolve(t, keys, key[k], kPosTag | (uint32_t)i, p[k], old[k]);
    }
  }
}

// FillWithUnique: the caller vouches for uniqueness, so item i simply becomes local id base + i - once every key sits
// in its slot (a probe must be able to read the key of any occupied slot, and that key is keys[pos] until n2o has it).
__global__ void __launch_bounds__(256)
table_number_unique_kernel(TableView t, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ pos, size_t n,
                           int fill) {
  const uint32_t base = t.items[fill - 1];
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t local = base + (uint32_t)i;
    const uint32_t p = pos[i];
    if (p != kEmpty && local < t.n2o_cap) {
      t.slots[p] = local;
      t.n2o[local] = keys[i];
      t.n2o_slot[local] = p;
    } else {
      *t.overflow = 1;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) t.items[fill] = base + (uint32_t)n;
}

// One tile = kAssignTile consecutive input positions, striped over the CTA for coalesced loads.
// Tiles are handed out by a ticket so that the chained scan's ordering requirement holds for any
// grid size.
__global__ void __launch_bounds__(kAssignBlock)
table_assign_kernel(TableView t, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ pos,
                    uint32_t *__restrict__ local_out, size_t n_max, const uint64_t *d_n,
                    int fill, bool ids_into_table, uint32_t *ticket, uint64_t *scan_state) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_cnt[kAssignItems * (kAssignBlock / 32)];
  __shared__ uint64_t s_tile_prefix;

  const uint64_t n = resolve_n(n_max, d_n);
  const uint32_t num_tiles = (uint32_t)((n + kAssignTile - 1) / kAssignTile);
  const uint32_t base_items = t.items[fill - 1];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int kWarps = kAssignBlock / 32;

  if (num_tiles == 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) t.items[fill] = base_items;
    return;
  }

  while (true) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= num_tiles) break;
    const uint64_t tile_base = (uint64_t)tile * kAssignTile;

    uint32_t slot_pos[kAssignItems];
    uint32_t word[kAssignItems];
    bool first[kAssignItems];
    uint32_t before[kAssignItems];  // flags before me inside my (row, warp)
#pragma unroll
    for (int k = 0; k < kAssignItems; ++k) {
      const uint64_t e = tile_base + (uint64_t)k * kAssignBlock + threadIdx.x;
      slot_pos[k] = (e < n) ? pos[e] : kEmpty;
    }
#pragma unroll
    for (int k = 0; k < kAssignItems; ++k) {
      const uint64_t e = tile_base + (uint64_t)k * kAssignBlock + threadIdx.x;
      word[k] = kEmpty;
      if (slot_pos[k] != kEmpty) word[k] = t.slots[slot_pos[k]];
      first[k] = slot_pos[k] != kEmpty && word[k] == (kPosTag | (uint32_t)e);
      const unsigned b = __ballot_sync(0xffffffffu, first[k]);
      before[k] = __popc(b & ((1u << lane) - 1));
      if (lane == 0) s_cnt[k * kWarps + warp] = __popc(b);
    }
    __syncthreads();
    // exclusive scan of the kAssignItems * kWarps (= 32) per-(row, warp) counts by warp 0
    if (warp == 0) {
      static_assert(kAssignItems * kWarps == 32, "scan below assumes 32 partial counts");
      const uint32_t c = s_cnt[lane];
      uint32_t incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += v;
      }
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      s_cnt[lane] = incl - c;
      const uint64_t excl = lookback_exclusive(scan_state, tile, total);
      if (lane == 0) {
        s_tile_prefix = excl;
        if (tile == num_tiles - 1) t.items[fill] = base_items + (uint32_t)(excl + total);
      }
    }
    __syncthreads();
    const uint32_t tile_first_local = base_items + (uint32_t)s_tile_prefix;
#pragma unroll
    for (int k = 0; k < kAssignItems; ++k) {
      if (first[k]) {
        const uint64_t e = tile_base + (uint64_t)k * kAssignBlock + threadIdx.x;
        const uint32_t local = tile_first_local + s_cnt[k * kWarps + warp] + before[k];
        if (local < t.n2o_cap) {
          t.n2o[local] = keys[e];  // nothing probes while this pass runs: the next insert kernel starts after it
          t.n2o_slot[local] = slot_pos[k];
          // the id goes into the slot for the fills that follow; the LAST fill of a batch has no follower (the next
          // thing that touches the table is Reset, which finds the slot through n2o_slot), so its ~4 M scattered
          // stores - a fifth of the layer's random accesses - are skipped
          if (ids_into_table) t.slots[slot_pos[k]] = local;
        } else {
          *t.overflow = 1;
          t.slots[slot_pos[k]] = kPosTag - 1;  // numbered "out of range"; the batch is discarded
        }
        word[k] = local;
      }
    }
    __syncthreads();  // s_tile / s_cnt are reused by the next tile; this tile's local ids are visible CTA-wide
    if (local_out != nullptr) {
      // Fused remap (GPUMapEdges for the dst side): every position gets the local id of its key in the same pass.
      // A later occurrence whose slot still held a tagged position when this tile loaded it (its FIRST occurrence sits
      // in a tile that is being numbered right now) cannot know the id yet: it stores the tagged word as a "pending"
      // mark and table_remap_pending_kernel resolves it after this pass.  (r01 spun on the slot here; with ~1200 tiles
      // in flight that parked whole CTAs behind their slowest lane: stalled_barrier 96 per issue.)
#pragma unroll
      for (int k = 0; k < kAssignItems; ++k) {
        const uint64_t e = tile_base + (uint64_t)k * kAssignBlock + threadIdx.x;
        if (e < n) local_out[e] = word[k];  // local id | pending (>= kPosTag) | kEmpty (table overflow)
      }
    }
  }
}

// second half of the fused remap: a position left "pending" by the numbering pass carries 0x80000000 | e', the
// position of its key's FIRST occurrence - whose own entry of local_out is final (a first occurrence is never
// pending), so the id is read from there: a 17 MB array instead of the table.  Four positions per thread.
__global__ void __launch_bounds__(256)
table_remap_pending_kernel(uint32_t *__restrict__ local_out, size_t n_max, const uint64_t *d_n) {
  const uint64_t n = resolve_n(n_max, d_n);
  const uint64_t n4 = n / 4;
  uint4 *out4 = reinterpret_cast<uint4 *>(local_out);
  auto fix = [&](uint32_t v) { return (v >= kPosTag && v != kEmpty) ? local_out[v & ~kPosTag] : v; };
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n4; i += (uint64_t)gridDim.x * blockDim.x) {
    uint4 v = out4[i];
    if ((v.x | v.y | v.z | v.w) & kPosTag) {
      v.x = fix(v.x); v.y = fix(v.y); v.z = fix(v.z); v.w = fix(v.w);
      out4[i] = v;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const uint64_t i = n4 * 4 + threadIdx.x;
    local_out[i] = fix(local_out[i]);
  }
}

// Reset.  The reference bumps a version word (cuda_hashtable.cu:739-742); 4-byte slots have no room for
// one.  Two ways, chosen on the device from the number of ids the finished batch numbered:
//   few ids  -> free exactly those slots (one scattered 4-byte store per id, n2o_slot remembers where);
//   many ids -> stream 0xFF over the whole table with 16-byte stores: cap * 4 bytes at copy bandwidth
//               beats one random sector per id, and it leaves the (L2-sized) table resident in L2 for
//               the inserts that follow.
__global__ void __launch_bounds__(256)
table_clear_kernel(TableView t, int fill) {
  const uint32_t n = min(t.items[fill], t.n2o_cap);
  const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const uint64_t nthr = (uint64_t)gridDim.x * blockDim.x;
  if ((uint64_t)n * 8 >= t.cap) {  // ~32 bytes of L2 traffic per scattered 4-byte store vs 4 streamed bytes per slot
    uint4 *p = reinterpret_cast<uint4 *>(t.slots);
    const uint64_t n4 = t.cap / 4;  // cap is a multiple of 4
    const uint4 free4 = make_uint4(kEmpty, kEmpty, kEmpty, kEmpty);
    for (uint64_t i = tid; i < n4; i += nthr) p[i] = free4;
  } else {
    for (uint64_t i = tid; i < n; i += nthr) t.slots[t.n2o_slot[i]] = kEmpty;
  }
}

__global__ void __launch_bounds__(256)
table_gather_local_kernel(TableView t, const uint32_t *__restrict__ pos, uint32_t *__restrict__ out,
                          size_t n_max, const uint64_t *d_n) {
  const uint64_t n = resolve_n(n_max, d_n);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t p = pos[i];
    out[i] = (p == kEmpty) ? kEmpty : t.slots[p];
  }
}

// SearchO2N for arrays; gridDim.y selects the (in, out) pair like map_edge_ids (cuda_mapping.cu:50)
__global__ void __launch_bounds__(256)
table_lookup_kernel(TableView t, const uint32_t *__restrict__ in0, uint32_t *__restrict__ out0,
                    const uint32_t *__restrict__ in1, uint32_t *__restrict__ out1, size_t n) {
  const uint32_t *in = blockIdx.y == 0 ? in0 : in1;
  uint32_t *out = blockIdx.y == 0 ? out0 : out1;
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t p = table_find(t, in[i]);
    out[i] = (p == kEmpty) ? kEmpty : t.slots[p];
  }
}

}  // namespace

int launch_table_insert(const TableView &t, const uint32_t *keys, uint32_t *pos_out, size_t n_max,
                        const uint64_t *d_n, cudaStream_t s) {
  if (n_max == 0) return XG_OK;
  table_insert_kernel<<<grid_for((n_max + 256 * kInsertItems - 1) / (256 * kInsertItems), 8), 256, 0, s>>>(t, keys, pos_out, n_max,
                                                                                                           d_n);
  XG_LAUNCHED();
  return XG_OK;
}

int launch_table_assign(const TableView &t, const uint32_t *keys, const uint32_t *pos, size_t n_max, const uint64_t *d_n,
                        int fill, const ScanSpace &scan, cudaStream_t s, uint32_t *local_out, bool ids_into_table) {
  size_t tiles = (n_max + kAssignTile - 1) / kAssignTile;
  XG_REQUIRE(tiles <= scan.max_tiles, "scan workspace too small");
  XG_REQUIRE(ids_into_table || local_out != nullptr, "a fill that leaves no ids in the table must hand them out");
  table_assign_kernel<<<grid_for(tiles, 8), kAssignBlock, 0, s>>>(t, keys, pos, local_out, n_max, d_n, fill, ids_into_table,
                                                                  scan.ticket, scan.state);
  XG_LAUNCHED();
  if (local_out != nullptr && n_max > 0) {
    XG_REQUIRE((reinterpret_cast<uintptr_t>(local_out) & 15) == 0, "local_out must be 16-byte aligned");
    table_remap_pending_kernel<<<grid_for((n_max + 1023) / 1024, 8), 256, 0, s>>>(local_out, n_max, d_n);
    XG_LAUNCHED();
  }
  return XG_OK;
}

int launch_table_clear(const TableView &t, int fill, size_t max_items, cudaStream_t s) {
  if (fill > 0) {
    table_clear_kernel<<<grid_for((max_items + 255) / 256, 8), 256, 0, s>>>(t, fill);  // 8 x 256 thr = full SM
    XG_LAUNCHED();
  }
  XG_CUDA(cudaMemsetAsync(t.items, 0, (kMaxFills + 2) * sizeof(uint32_t), s));
  return XG_OK;
}

int launch_table_gather_local(const TableView &t, const uint32_t *pos, uint32_t *out, size_t n_max,
                              const uint64_t *d_n, cudaStream_t s) {
  if (n_max == 0) return XG_OK;
  table_gather_local_kernel<<<grid_for((n_max + 255) / 256, 8), 256, 0, s>>>(t, pos, out, n_max, d_n);
  XG_LAUNCHED();
  return XG_OK;
}

}  // namespace xg

// =================================================================================================
// C-ABI object
// =================================================================================================
using namespace xg;

struct xg_hashtable {
  int device = 0;
  size_t max_items = 0;
  size_t capacity = 0;
  TableView view{};
  int fill = 0;             // fills since Reset (index into view.items)
  uint32_t *scratch = nullptr;  // slot positions of the current fill (standalone API only)
  size_t scratch_len = 0;
  void *scan_mem = nullptr;  // ticket + state words
  size_t scan_tiles = 0;
  cudaStream_t last_stream = nullptr;
  bool dirty_full = false;  // an overflow was seen: the next Reset wipes the whole table
};

namespace {

struct DeviceGuard {
  int prev = 0;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    target = dev;
  }
  ~DeviceGuard() {
    if (prev != target) cudaSetDevice(prev);
  }
  int target;
};

int ensure_scratch(xg_hashtable *t, size_t n) {
  if (t->scratch_len < n) {
    if (t->scratch) XG_CUDA(cudaFree(t->scratch));
    t->scratch = nullptr;
    size_t len = n + n / 4 + 1024;
    XG_CUDA(cudaMalloc(&t->scratch, len * sizeof(uint32_t)));
    t->scratch_len = len;
  }
  size_t tiles = (n + kAssignTile - 1) / kAssignTile + 1;
  if (t->scan_tiles < tiles) {
    if (t->scan_mem) XG_CUDA(cudaFree(t->scan_mem));
    t->scan_mem = nullptr;
    size_t want = tiles + tiles / 4 + 64;
    XG_CUDA(cudaMalloc(&t->scan_mem, (want + 1) * sizeof(uint64_t)));
    t->scan_tiles = want;
  }
  return XG_OK;
}

ScanSpace scan_space(xg_hashtable *t) {
  ScanSpace s;
  s.ticket = reinterpret_cast<uint32_t *>(t->scan_mem);
  s.state = reinterpret_cast<uint64_t *>(t->scan_mem) + 1;
  s.max_tiles = t->scan_tiles;
  return s;
}

// When a caller issues more fills than the device-side counter array holds, fold the history:
// items[0] <- items[fill], fill <- 0.  Needs one small D2D copy, no sync.
int fold_fills(xg_hashtable *t, cudaStream_t s) {
  if (t->fill < kMaxFills) return XG_OK;
  XG_CUDA(cudaMemcpyAsync(t->view.items, t->view.items + t->fill, sizeof(uint32_t),
                          cudaMemcpyDeviceToDevice, s));
  t->fill = 0;
  return XG_OK;
}

int fill_dup(xg_hashtable *t, const uint32_t *d_input, uint32_t *pos, size_t n, cudaStream_t s) {
  int rc = fold_fills(t, s);
  if (rc) return rc;
  t->fill += 1;
  t->last_stream = s;
  size_t tiles = (n + kAssignTile - 1) / kAssignTile;
  XG_CUDA(cudaMemsetAsync(t->scan_mem, 0, (tiles + 2) * sizeof(uint64_t), s));
  rc = launch_table_insert(t->view, d_input, pos, n, nullptr, s);
  if (rc) return rc;
  return launch_table_assign(t->view, d_input, pos, n, nullptr, t->fill, scan_space(t), s);
}

int check_overflow(xg_hashtable *t, cudaStream_t s) {
  uint32_t ovf = 0;
  XG_CUDA(cudaMemcpyAsync(&ovf, t->view.overflow, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  XG_CUDA(cudaStreamSynchronize(s));
  if (ovf) {
    t->dirty_full = true;
    set_error("hash table capacity exceeded (max_items=" + std::to_string(t->max_items) + ")");
    return XG_ERR_CAPACITY;
  }
  return XG_OK;
}

}  // namespace

namespace xg {
// accessors used by the minibatch pipeline (pipeline.cu)
TableView table_view(xg_hashtable *t) { return t->view; }
int table_begin_fill(xg_hashtable *t, cudaStream_t s) {
  int rc = fold_fills(t, s);
  if (rc) return rc;
  t->fill += 1;
  t->last_stream = s;
  return t->fill;
}
}  // namespace xg

extern "C" {

int xg_hashtable_create(xg_hashtable **out, size_t max_items, int device) {
  XG_REQUIRE(out != nullptr, "out is NULL");
  XG_REQUIRE(max_items > 0 && max_items < (1ull << 30), "max_items must be in (0, 2^30)");
  DeviceGuard guard(device);
  auto *t = new xg_hashtable();
  t->device = device;
  t->max_items = max_items;
  // slots per item: 1.5 by default.  The reference takes 2-4 (cuda_hashtable.cu:146-149) in 16-byte buckets; here the
  // worst-case batch of the headline config (8.45 M ids) gives 51 MB of 4-byte slots, inside the part of the 126 MB
  // L2 a randomly probed table can really use (~64 MB, tools/table_size_probe.py): every probe is a random access
  // and an L2 hit costs a third of a DRAM miss.  A typical batch fills 39 % of that.  XG_TABLE_SLOTS_PCT overrides
  // (>= 110).
  size_t pct = 150;
  if (const char *e = std::getenv("XG_TABLE_SLOTS_PCT")) pct = std::strtoul(e, nullptr, 10);
  if (pct < 110) pct = 110;
  t->capacity = (max_items * pct / 100 + 1024) & ~(size_t)3;
  XG_REQUIRE(t->capacity < (1ull << 32), "table too large");
  XG_CUDA(cudaMalloc(&t->view.slots, t->capacity * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&t->view.n2o, max_items * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&t->view.n2o_slot, max_items * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&t->view.items, (kMaxFills + 2) * sizeof(uint32_t)));
  t->view.overflow = t->view.items + kMaxFills + 1;
  t->view.cap = static_cast<uint32_t>(t->capacity);
  t->view.n2o_cap = static_cast<uint32_t>(max_items);
  XG_CUDA(cudaMemset(t->view.slots, 0xFF, t->capacity * sizeof(uint32_t)));
  XG_CUDA(cudaMemset(t->view.items, 0, (kMaxFills + 2) * sizeof(uint32_t)));
  *out = t;
  return XG_OK;
}

int xg_hashtable_destroy(xg_hashtable *t) {
  if (!t) return XG_OK;
  DeviceGuard guard(t->device);
  cudaFree(t->view.slots);
  cudaFree(t->view.n2o);
  cudaFree(t->view.n2o_slot);
  cudaFree(t->view.items);
  cudaFree(t->scratch);
  cudaFree(t->scan_mem);
  delete t;
  return XG_OK;
}

int xg_hashtable_reset(xg_hashtable *t, void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  DeviceGuard guard(t->device);
  cudaStream_t s = as_stream(stream);
  if (t->dirty_full) {
    XG_CUDA(cudaMemsetAsync(t->view.slots, 0xFF, t->capacity * sizeof(uint32_t), s));
    XG_CUDA(cudaMemsetAsync(t->view.items, 0, (kMaxFills + 2) * sizeof(uint32_t), s));
    t->dirty_full = false;
  } else {
    int rc = launch_table_clear(t->view, t->fill, t->max_items, s);
    if (rc) return rc;
  }
  t->fill = 0;
  t->last_stream = s;
  return XG_OK;
}

int xg_hashtable_fill_with_unique(xg_hashtable *t, const uint32_t *d_input, size_t num_input,
                                  void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  if (num_input == 0) return XG_OK;
  DeviceGuard guard(t->device);
  cudaStream_t s = as_stream(stream);
  int rc = ensure_scratch(t, num_input);
  if (rc) return rc;
  if ((rc = fold_fills(t, s))) return rc;
  t->fill += 1;
  t->last_stream = s;
  // every key into its slot first (positions tagged), then number them: local id = base + position
  if ((rc = launch_table_insert(t->view, d_input, t->scratch, num_input, nullptr, s))) return rc;
  table_number_unique_kernel<<<grid_for((num_input + 255) / 256, 8), 256, 0, s>>>(t->view, d_input, t->scratch,
                                                                                num_input, t->fill);
  XG_LAUNCHED();
  return XG_OK;
}

int xg_hashtable_fill_with_dup_revised(xg_hashtable *t, const uint32_t *d_input, size_t num_input,
                                       void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  if (num_input == 0) return XG_OK;
  DeviceGuard guard(t->device);
  int rc = ensure_scratch(t, num_input);
  if (rc) return rc;
  return fill_dup(t, d_input, t->scratch, num_input, as_stream(stream));
}

int xg_hashtable_fill_with_dup_mutable(xg_hashtable *t, uint32_t *d_input, size_t num_input,
                                       void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  if (num_input == 0) return XG_OK;
  DeviceGuard guard(t->device);
  int rc = ensure_scratch(t, num_input);
  if (rc) return rc;
  // (the r01 table let the slot positions overwrite the input in place; keys are now read back through their
  // positions while the fill is in flight, so the input stays intact and the positions go to the scratch array)
  return fill_dup(t, d_input, t->scratch, num_input, as_stream(stream));
}

int xg_hashtable_fill_with_duplicates(xg_hashtable *t, const uint32_t *d_input, size_t num_input,
                                      uint32_t *d_unique, uint32_t *h_num_unique, void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  XG_REQUIRE(h_num_unique != nullptr, "h_num_unique is NULL");
  DeviceGuard guard(t->device);
  cudaStream_t s = as_stream(stream);
  if (num_input > 0) {
    int rc = ensure_scratch(t, num_input);
    if (rc) return rc;
    rc = fill_dup(t, d_input, t->scratch, num_input, s);
    if (rc) return rc;
  }
  uint32_t n = 0;
  XG_CUDA(cudaMemcpyAsync(&n, t->view.items + t->fill, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  int rc = check_overflow(t, s);  // synchronises
  if (rc) return rc;
  if (d_unique && n)
    XG_CUDA(cudaMemcpyAsync(d_unique, t->view.n2o, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
  *h_num_unique = n;
  return XG_OK;
}

size_t xg_hashtable_num_items(xg_hashtable *t) {
  if (!t) return 0;
  DeviceGuard guard(t->device);
  uint32_t n = 0;
  if (cudaMemcpyAsync(&n, t->view.items + t->fill, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                      t->last_stream) != cudaSuccess)
    return 0;
  cudaStreamSynchronize(t->last_stream);
  return n;
}

int xg_hashtable_ref_unique(xg_hashtable *t, const uint32_t **d_unique, uint32_t *h_num_unique) {
  XG_REQUIRE(t && d_unique && h_num_unique, "NULL argument");
  *d_unique = t->view.n2o;
  *h_num_unique = static_cast<uint32_t>(xg_hashtable_num_items(t));
  return check_overflow(t, t->last_stream);
}

int xg_hashtable_copy_unique(xg_hashtable *t, uint32_t *d_unique, void *stream) {
  XG_REQUIRE(t && d_unique, "NULL argument");
  DeviceGuard guard(t->device);
  size_t n = xg_hashtable_num_items(t);
  if (n)
    XG_CUDA(cudaMemcpyAsync(d_unique, t->view.n2o, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice,
                            as_stream(stream)));
  return XG_OK;
}

int xg_hashtable_lookup(xg_hashtable *t, const uint32_t *d_ids, size_t n, uint32_t *d_local,
                        void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  if (n == 0) return XG_OK;
  DeviceGuard guard(t->device);
  dim3 grid(grid_for((n + 255) / 256, 8), 1);
  table_lookup_kernel<<<grid, 256, 0, as_stream(stream)>>>(t->view, d_ids, d_local, nullptr, nullptr, n);
  XG_LAUNCHED();
  return XG_OK;
}

int xg_map_edges(xg_hashtable *t, const uint32_t *d_src, uint32_t *d_new_src, const uint32_t *d_dst,
                 uint32_t *d_new_dst, size_t num_edges, void *stream) {
  XG_REQUIRE(t != nullptr, "table is NULL");
  if (num_edges == 0) return XG_OK;
  DeviceGuard guard(t->device);
  dim3 grid(grid_for((num_edges + 255) / 256, 8), 2);
  table_lookup_kernel<<<grid, 256, 0, as_stream(stream)>>>(t->view, d_src, d_new_src, d_dst, d_new_dst,
                                                          num_edges);
  XG_LAUNCHED();
  return XG_OK;
}

}  // extern "C"
