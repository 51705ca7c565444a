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

__global__ void __launch_bounds__(256)
table_insert_kernel(TableView t, const uint32_t *keys, uint32_t *pos_out,  // may alias (DupMutable)
                    size_t n_max, const uint64_t *d_n) {
  const uint64_t n = resolve_n(n_max, d_n);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    pos_out[i] = table_insert(t, keys[i], kPosTag | (uint32_t)i);
  }
}

// FillWithUnique: the caller vouches for uniqueness, so item i simply becomes local id base + i.
__global__ void __launch_bounds__(256)
table_fill_unique_kernel(TableView t, const uint32_t *__restrict__ keys, size_t n, int fill) {
  const uint32_t base = t.items[fill - 1];
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t key = keys[i];
    const uint32_t local = base + (uint32_t)i;
    uint32_t p = table_insert(t, key, local);  // local < 2^31 < any tag: atomicMin keeps it
    if (p != kEmpty && local < t.n2o_cap) {
      t.n2o[local] = key;
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
table_assign_kernel(TableView t, const uint32_t *__restrict__ pos, uint32_t *__restrict__ local_out, size_t n_max,
                    const uint64_t *d_n,
                    int fill, uint32_t *ticket, uint64_t *scan_state) {
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
    uint2 slot[kAssignItems];
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
      slot[k] = make_uint2(kEmpty, 0);
      if (slot_pos[k] != kEmpty) slot[k] = t.slots[slot_pos[k]];
      first[k] = slot_pos[k] != kEmpty && slot[k].y == (kPosTag | (uint32_t)e);
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
        const uint32_t local = tile_first_local + s_cnt[k * kWarps + warp] + before[k];
        if (local < t.n2o_cap) {
          t.slots[slot_pos[k]].y = local;
          t.n2o[local] = slot[k].x;
          t.n2o_slot[local] = slot_pos[k];
        } else {
          *t.overflow = 1;
          t.slots[slot_pos[k]].y = kPosTag - 1;  // numbered "out of range": releases the waiters below; the batch is discarded
        }
        slot[k].y = local;
      }
    }
    __syncthreads();  // s_tile / s_cnt are reused by the next tile; this tile's local ids are visible CTA-wide
    if (local_out != nullptr) {
      // Fused remap (GPUMapEdges for the dst side): every position gets the local id of its key in the same pass.
      // A later occurrence whose slot still holds a tagged position waits for the FIRST occurrence to be numbered:
      // that one sits at a smaller position, i.e. in this tile (already written, see the barrier above) or in a tile
      // with a smaller ticket, which is running or done and never waits on a later tile - so the wait terminates.
#pragma unroll
      for (int k = 0; k < kAssignItems; ++k) {
        const uint64_t e = tile_base + (uint64_t)k * kAssignBlock + threadIdx.x;
        if (e >= n) continue;
        uint32_t v = kEmpty;
        if (slot_pos[k] != kEmpty) {
          v = slot[k].y;
          if (v >= kPosTag) {
            const volatile uint32_t *vp = &t.slots[slot_pos[k]].y;
            while ((v = *vp) >= kPosTag) __nanosleep(20);
          }
        }
        local_out[e] = v;
      }
    }
  }
}

// Reset.  The reference bumps a version word (cuda_hashtable.cu:739-742); 8-byte slots have no room for
// one.  Two ways, chosen on the device from the number of ids the finished batch numbered:
//   few ids  -> free exactly those slots (one scattered 8-byte store per id, n2o_slot remembers where);
//   many ids -> stream 0xFF over the whole table with 16-byte stores: cap * 8 bytes at copy bandwidth
//               beats one random sector per id, and it leaves the (L2-sized) table resident in L2 for
//               the inserts that follow.
__global__ void __launch_bounds__(256)
table_clear_kernel(TableView t, int fill) {
  const uint32_t n = min(t.items[fill], t.n2o_cap);
  const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const uint64_t nthr = (uint64_t)gridDim.x * blockDim.x;
  if ((uint64_t)n * 8 >= t.cap) {
    uint4 *p = reinterpret_cast<uint4 *>(t.slots);
    const uint64_t n4 = t.cap / 2;  // cap is even
    const uint4 free4 = make_uint4(kEmpty, kEmpty, kEmpty, kEmpty);
    for (uint64_t i = tid; i < n4; i += nthr) p[i] = free4;
  } else {
    for (uint64_t i = tid; i < n; i += nthr) t.slots[t.n2o_slot[i]] = make_uint2(kEmpty, kEmpty);
  }
}

__global__ void __launch_bounds__(256)
table_gather_local_kernel(TableView t, const uint32_t *__restrict__ pos, uint32_t *__restrict__ out,
                          size_t n_max, const uint64_t *d_n) {
  const uint64_t n = resolve_n(n_max, d_n);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t p = pos[i];
    out[i] = (p == kEmpty) ? kEmpty : t.slots[p].y;
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
    out[i] = (p == kEmpty) ? kEmpty : t.slots[p].y;
  }
}

}  // namespace

int launch_table_insert(const TableView &t, const uint32_t *keys, uint32_t *pos_out, size_t n_max,
                        const uint64_t *d_n, cudaStream_t s) {
  if (n_max == 0) return XG_OK;
  table_insert_kernel<<<grid_for((n_max + 255) / 256, 8), 256, 0, s>>>(t, keys, pos_out, n_max, d_n);
  XG_LAUNCHED();
  return XG_OK;
}

int launch_table_assign(const TableView &t, const uint32_t *pos, size_t n_max, const uint64_t *d_n,
                        int fill, const ScanSpace &scan, cudaStream_t s, uint32_t *local_out) {
  size_t tiles = (n_max + kAssignTile - 1) / kAssignTile;
  XG_REQUIRE(tiles <= scan.max_tiles, "scan workspace too small");
  table_assign_kernel<<<grid_for(tiles, 8), kAssignBlock, 0, s>>>(t, pos, local_out, n_max, d_n, fill, scan.ticket,
                                                                  scan.state);
  XG_LAUNCHED();
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
  return launch_table_assign(t->view, pos, n, nullptr, t->fill, scan_space(t), s);
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
  // slots per item: 1.25 by default.  The reference takes 2-4 (cuda_hashtable.cu:146-149) in 16-byte
  // buckets; here the worst-case batch of the headline config (8.45 M ids) must stay inside the 126 MB
  // L2 (84 MB at 1.25), because every probe is a random access and an L2 hit costs a third of a DRAM
  // miss.  A typical batch fills 45 % of that.  XG_TABLE_SLOTS_PCT overrides (>= 110).
  size_t pct = 125;
  if (const char *e = std::getenv("XG_TABLE_SLOTS_PCT")) pct = std::strtoul(e, nullptr, 10);
  if (pct < 110) pct = 110;
  t->capacity = (max_items * pct / 100 + 1024) & ~(size_t)1;
  XG_REQUIRE(t->capacity < (1ull << 32), "table too large");
  XG_CUDA(cudaMalloc(&t->view.slots, t->capacity * sizeof(uint2)));
  XG_CUDA(cudaMalloc(&t->view.n2o, max_items * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&t->view.n2o_slot, max_items * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&t->view.items, (kMaxFills + 2) * sizeof(uint32_t)));
  t->view.overflow = t->view.items + kMaxFills + 1;
  t->view.cap = static_cast<uint32_t>(t->capacity);
  t->view.n2o_cap = static_cast<uint32_t>(max_items);
  XG_CUDA(cudaMemset(t->view.slots, 0xFF, t->capacity * sizeof(uint2)));
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
    XG_CUDA(cudaMemsetAsync(t->view.slots, 0xFF, t->capacity * sizeof(uint2), s));
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
  int rc = fold_fills(t, s);
  if (rc) return rc;
  t->fill += 1;
  t->last_stream = s;
  table_fill_unique_kernel<<<grid_for((num_input + 255) / 256, 8), 256, 0, s>>>(t->view, d_input,
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
  // the slot positions overwrite the input in place: no scratch traffic
  return fill_dup(t, d_input, d_input, num_input, as_stream(stream));
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
