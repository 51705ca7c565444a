// Feature / label row gather over the GGMS feature store.
//
// Reference behaviour being replaced:
//   gpu_extract                        cuda/cuda_extraction.cu:31-116          (plain gather, 4-byte loads)
//   GetMissCacheIndex (4 kernels)      cuda/cuda_cache_manager_device.cu:41-169,355-441
//   extract_miss_data                  cuda/cuda_cache_manager_device.cu:234-253,573-625 (host UVA)
//   combine_cache_data_for_partition   cuda/cuda_cache_manager_device.cu:278-299,628-731 (local+peer)
// i.e. classification into index lists (with two scans and two D2H counts) followed by two
// element-wise gathers whose routing (`% P`, `/ P`, pointer-table load) is redone per ELEMENT.
//
// Here: ONE launch.  A CTA takes a tile of rows;
//   phase 1 (thread per row): node id -> cache_table -> (part, row) or host -> source row pointer,
//            staged in shared memory -- the three dependent loads are issued for a whole tile at once;
//   phase 2 (all threads): the tile is a flat list of 16-byte chunks (row-major), copied with
//            8 independent 128-bit loads in flight per thread before the matching stores, so lanes
//            are never idle on rows that are not a multiple of 512 B (e.g. 100 x fp32 = 25 chunks).
// Local HBM, NVLink-peer HBM and pinned host rows take the same path; only the pointer differs.
#include "extract.cuh"

namespace xg {

namespace {

constexpr int kGatherBlock = 256;
constexpr int kGatherUnroll = 8;
constexpr int kGatherChunksPerTile = kGatherBlock * kGatherUnroll;  // 2048 chunks (32 KB at 16 B)
constexpr int kGatherMaxRows = 256;                                  // rows per tile upper bound

template <int VW> struct Vec;
template <> struct Vec<16> { using T = int4; };
template <> struct Vec<8> { using T = int2; };
template <> struct Vec<4> { using T = int; };
template <> struct Vec<2> { using T = short; };
template <> struct Vec<1> { using T = char; };

template <typename T> __device__ __forceinline__ T ld_stream(const T *p) { return *p; }
template <> __device__ __forceinline__ int4 ld_stream<int4>(const int4 *p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
template <> __device__ __forceinline__ int2 ld_stream<int2>(const int2 *p) {
  int2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
template <typename T> __device__ __forceinline__ void st_stream(T *p, const T &v) { *p = v; }
template <> __device__ __forceinline__ void st_stream<int4>(int4 *p, const int4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ const char *route_row(const xg_feature_store &st, int mode, uint32_t idx,
                                                 int *klass) {
  uint32_t rank = kEmpty;
  if (mode == kCacheRank) rank = idx;
  else if (mode == kByNode && st.cache_table != nullptr) rank = __ldg(st.cache_table + idx);
  if (rank == kEmpty) {
    *klass = 2;  // host (or the flat table when nothing is cached)
    return static_cast<const char *>(st.host) + (size_t)idx * st.row_bytes;
  }
  const uint32_t part = rank % st.num_part;
  *klass = ((st.local_part_mask >> part) & 1u) ? 0 : 1;
  return static_cast<const char *>(st.part[part]) + (size_t)(rank / st.num_part) * st.row_bytes;
}

template <int VW>
__global__ void __launch_bounds__(kGatherBlock)
gather_rows_kernel(const xg_feature_store st, int mode, const uint32_t *__restrict__ src_idx,
                   const uint32_t *__restrict__ dst_idx, size_t num_rows_max,
                   const uint32_t *__restrict__ d_num_rows, uint32_t rows_per_tile,
                   uint32_t chunks_per_row, char *__restrict__ out, unsigned long long *stats) {
  using V = typename Vec<VW>::T;
  __shared__ const char *s_src[kGatherMaxRows];
  __shared__ char *s_dst[kGatherMaxRows];
  __shared__ unsigned s_class[3];

  size_t num_rows = num_rows_max;
  if (d_num_rows && *d_num_rows < num_rows_max) num_rows = *d_num_rows;  // count still on the device
  const size_t num_tiles = (num_rows + rows_per_tile - 1) / rows_per_tile;
  for (size_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const size_t row0 = tile * rows_per_tile;
    const uint32_t rows = (uint32_t)((num_rows - row0) < rows_per_tile ? (num_rows - row0) : rows_per_tile);
    if (stats && threadIdx.x < 3) s_class[threadIdx.x] = 0;
    if (stats) __syncthreads();
    // phase 1: resolve source / destination row pointers
    if (threadIdx.x < rows) {
      const size_t r = row0 + threadIdx.x;
      int klass;
      s_src[threadIdx.x] = route_row(st, mode, src_idx[r], &klass);
      const size_t drow = dst_idx ? (size_t)dst_idx[r] : r;
      s_dst[threadIdx.x] = out + drow * st.row_bytes;
      if (stats) atomicAdd(&s_class[klass], 1u);
    }
    __syncthreads();
    if (stats && threadIdx.x < 3 && s_class[threadIdx.x])
      atomicAdd(stats + threadIdx.x, (unsigned long long)s_class[threadIdx.x]);

    // phase 2: flat chunk copy, kGatherUnroll loads in flight per thread
    const uint32_t total = rows * chunks_per_row;
    for (uint32_t f0 = 0; f0 < total; f0 += kGatherChunksPerTile) {
      V buf[kGatherUnroll];
      uint32_t row[kGatherUnroll], col[kGatherUnroll];
#pragma unroll
      for (int u = 0; u < kGatherUnroll; ++u) {
        const uint32_t f = f0 + u * kGatherBlock + threadIdx.x;
        row[u] = f / chunks_per_row;
        col[u] = f - row[u] * chunks_per_row;
        if (f < total) buf[u] = ld_stream(reinterpret_cast<const V *>(s_src[row[u]]) + col[u]);
      }
#pragma unroll
      for (int u = 0; u < kGatherUnroll; ++u) {
        const uint32_t f = f0 + u * kGatherBlock + threadIdx.x;
        if (f < total) st_stream(reinterpret_cast<V *>(s_dst[row[u]]) + col[u], buf[u]);
      }
    }
    __syncthreads();  // pointer staging is reused by the next tile
  }
}

bool aligned_to(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

int pick_vector_width(const xg_feature_store &st, const void *out) {
  for (int vw = 16; vw > 1; vw >>= 1) {
    bool ok = (st.row_bytes % vw == 0) && aligned_to(out, vw) && (st.host == nullptr || aligned_to(st.host, vw));
    for (uint32_t p = 0; ok && p < st.num_part; ++p) ok = st.part[p] == nullptr || aligned_to(st.part[p], vw);
    if (ok) return vw;
  }
  return 1;
}

}  // namespace

int launch_gather(const xg_feature_store &st, int mode, const uint32_t *src_idx, const uint32_t *dst_idx,
                  size_t num_rows, const uint32_t *d_num_rows, void *out, uint64_t *stats, cudaStream_t s) {
  if (num_rows == 0) return XG_OK;
  XG_REQUIRE(st.row_bytes > 0, "row_bytes is 0");
  XG_REQUIRE(out != nullptr && src_idx != nullptr, "NULL buffer");
  XG_REQUIRE(st.num_part <= XG_MAX_PARTS, "too many partitions");
  XG_REQUIRE(mode != kCacheRank || st.num_part > 0, "cache-rank gather needs a partitioned cache");
  XG_REQUIRE(mode == kCacheRank || st.host != nullptr || st.cache_table != nullptr, "no source table");
  const int vw = pick_vector_width(st, out);
  const uint64_t cpr64 = st.row_bytes / vw;
  XG_REQUIRE(cpr64 < (1u << 24), "row too long");
  const uint32_t cpr = (uint32_t)cpr64;
  uint32_t rows_per_tile = kGatherChunksPerTile / cpr;
  if (rows_per_tile < 1) rows_per_tile = 1;
  if (rows_per_tile > kGatherMaxRows) rows_per_tile = kGatherMaxRows;
  const size_t tiles = (num_rows + rows_per_tile - 1) / rows_per_tile;
  const unsigned grid = grid_for(tiles, 8);
  auto *o = static_cast<char *>(out);
  auto *st64 = reinterpret_cast<unsigned long long *>(stats);
  switch (vw) {
    case 16: gather_rows_kernel<16><<<grid, kGatherBlock, 0, s>>>(st, mode, src_idx, dst_idx, num_rows, d_num_rows, rows_per_tile, cpr, o, st64); break;
    case 8: gather_rows_kernel<8><<<grid, kGatherBlock, 0, s>>>(st, mode, src_idx, dst_idx, num_rows, d_num_rows, rows_per_tile, cpr, o, st64); break;
    case 4: gather_rows_kernel<4><<<grid, kGatherBlock, 0, s>>>(st, mode, src_idx, dst_idx, num_rows, d_num_rows, rows_per_tile, cpr, o, st64); break;
    case 2: gather_rows_kernel<2><<<grid, kGatherBlock, 0, s>>>(st, mode, src_idx, dst_idx, num_rows, d_num_rows, rows_per_tile, cpr, o, st64); break;
    default: gather_rows_kernel<1><<<grid, kGatherBlock, 0, s>>>(st, mode, src_idx, dst_idx, num_rows, d_num_rows, rows_per_tile, cpr, o, st64); break;
  }
  XG_LAUNCHED();
  return XG_OK;
}

namespace {

// ---- GetMissCacheIndex: stable two-way split with one chained scan (both counts in one word) ----
constexpr int kSplitBlock = 256;
constexpr int kSplitItems = 4;
constexpr int kSplitTile = kSplitBlock * kSplitItems;

__global__ void __launch_bounds__(kSplitBlock)
split_miss_cache_kernel(const uint32_t *__restrict__ table, const uint32_t *__restrict__ nodes, size_t n,
                        uint32_t *__restrict__ miss_src, uint32_t *__restrict__ miss_dst,
                        uint32_t *__restrict__ cache_src, uint32_t *__restrict__ cache_dst,
                        unsigned long long *counts /*[2]: miss, cache*/, uint32_t *ticket,
                        uint64_t *scan_state) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_miss[kSplitItems * (kSplitBlock / 32)];
  __shared__ uint64_t s_prefix;
  constexpr int kWarps = kSplitBlock / 32;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t num_tiles = (uint32_t)((n + kSplitTile - 1) / kSplitTile);
  while (true) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= num_tiles) break;
    const size_t base = (size_t)tile * kSplitTile;
    uint32_t node[kSplitItems], rank[kSplitItems], before[kSplitItems];
    bool valid[kSplitItems];
#pragma unroll
    for (int k = 0; k < kSplitItems; ++k) {
      const size_t i = base + (size_t)k * kSplitBlock + threadIdx.x;
      valid[k] = i < n;
      node[k] = valid[k] ? nodes[i] : 0;
    }
#pragma unroll
    for (int k = 0; k < kSplitItems; ++k) {
      rank[k] = valid[k] ? __ldg(table + node[k]) : 0;
      const bool miss = valid[k] && rank[k] == kEmpty;
      const unsigned b = __ballot_sync(0xffffffffu, miss);
      before[k] = __popc(b & ((1u << lane) - 1));
      if (lane == 0) s_miss[k * kWarps + warp] = __popc(b);
    }
    __syncthreads();
    if (warp == 0) {
      static_assert(kSplitItems * kWarps == 32, "32 partial counts expected");
      const uint32_t c = s_miss[lane];
      uint32_t incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += v;
      }
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      s_miss[lane] = incl - c;
      const uint64_t excl = lookback_exclusive(scan_state, tile, total);
      if (lane == 0) {
        s_prefix = excl;
        if (tile == num_tiles - 1) {
          counts[0] = excl + total;
          counts[1] = n - (excl + total);
        }
      }
    }
    __syncthreads();
    const uint64_t miss_before_tile = s_prefix;
#pragma unroll
    for (int k = 0; k < kSplitItems; ++k) {
      if (!valid[k]) continue;
      const size_t i = base + (size_t)k * kSplitBlock + threadIdx.x;
      const uint64_t miss_before = miss_before_tile + s_miss[k * kWarps + warp] + before[k];
      if (rank[k] == kEmpty) {
        miss_src[miss_before] = node[k];
        miss_dst[miss_before] = (uint32_t)i;
      } else {
        const uint64_t c = i - miss_before;  // hits before me = position - misses before me
        cache_src[c] = rank[k];
        cache_dst[c] = (uint32_t)i;
      }
    }
    __syncthreads();
  }
}

struct SplitWorkspace {
  void *mem = nullptr;
  size_t tiles = 0;
};
SplitWorkspace g_split_ws[XG_MAX_PARTS * 4];

}  // namespace

}  // namespace xg

using namespace xg;

extern "C" {

int xg_extract(void *d_dst, const void *src, const uint32_t *d_index, size_t num_index, size_t dim,
               int dtype, void *stream) {
  size_t elem = 0;
  switch (dtype) {
    case XG_F32: case XG_I32: elem = 4; break;
    case XG_F64: case XG_I64: elem = 8; break;
    case XG_F16: elem = 2; break;
    case XG_U8: case XG_I8: elem = 1; break;
    default: XG_REQUIRE(false, "unknown dtype");
  }
  XG_REQUIRE(src != nullptr, "src is NULL");
  xg_feature_store st{};
  st.host = src;
  st.row_bytes = dim * elem;
  return launch_gather(st, kHostRow, d_index, nullptr, num_index, nullptr, d_dst, nullptr, as_stream(stream));
}

int xg_extract_ggms(const xg_feature_store *store, const uint32_t *d_nodes, size_t num_nodes, void *d_out,
                    uint64_t *d_stats, void *stream) {
  XG_REQUIRE(store != nullptr, "store is NULL");
  return launch_gather(*store, kByNode, d_nodes, nullptr, num_nodes, nullptr, d_out, d_stats, as_stream(stream));
}

int xg_extract_miss_data(void *d_out, const void *host_feat, const uint32_t *d_miss_src,
                         const uint32_t *d_miss_dst, size_t num_miss, size_t row_bytes, void *stream) {
  XG_REQUIRE(host_feat != nullptr, "host_feat is NULL");
  xg_feature_store st{};
  st.host = host_feat;
  st.row_bytes = row_bytes;
  return launch_gather(st, kHostRow, d_miss_src, d_miss_dst, num_miss, nullptr, d_out, nullptr, as_stream(stream));
}

int xg_combine_cache_data(void *d_out, const xg_feature_store *store, const uint32_t *d_cache_src,
                          const uint32_t *d_cache_dst, size_t num_cache, void *stream) {
  XG_REQUIRE(store != nullptr, "store is NULL");
  return launch_gather(*store, kCacheRank, d_cache_src, d_cache_dst, num_cache, nullptr, d_out, nullptr,
                       as_stream(stream));
}

int xg_get_miss_cache_index(const uint32_t *d_cache_table, const uint32_t *d_nodes, size_t num_nodes,
                            uint32_t *d_miss_src, uint32_t *d_miss_dst, size_t *h_num_miss,
                            uint32_t *d_cache_src, uint32_t *d_cache_dst, size_t *h_num_cache,
                            void *stream) {
  XG_REQUIRE(d_cache_table && h_num_miss && h_num_cache, "NULL argument");
  cudaStream_t s = as_stream(stream);
  *h_num_miss = 0;
  *h_num_cache = 0;
  if (num_nodes == 0) return XG_OK;
  int dev = 0;
  XG_CUDA(cudaGetDevice(&dev));
  SplitWorkspace &ws = g_split_ws[dev % (XG_MAX_PARTS * 4)];
  const size_t tiles = (num_nodes + kSplitTile - 1) / kSplitTile + 1;
  if (ws.tiles < tiles) {
    if (ws.mem) {
      XG_CUDA(cudaStreamSynchronize(s));
      XG_CUDA(cudaFree(ws.mem));
    }
    size_t want = tiles + tiles / 2 + 64;
    XG_CUDA(cudaMalloc(&ws.mem, (want + 3) * sizeof(uint64_t)));
    ws.tiles = want;
  }
  // layout: [0] ticket, [1..2] counts, [3..] scan state
  XG_CUDA(cudaMemsetAsync(ws.mem, 0, (tiles + 3) * sizeof(uint64_t), s));
  auto *words = reinterpret_cast<uint64_t *>(ws.mem);
  split_miss_cache_kernel<<<grid_for(tiles, 8), kSplitBlock, 0, s>>>(
      d_cache_table, d_nodes, num_nodes, d_miss_src, d_miss_dst, d_cache_src, d_cache_dst,
      reinterpret_cast<unsigned long long *>(words + 1), reinterpret_cast<uint32_t *>(words), words + 3);
  XG_LAUNCHED();
  uint64_t counts[2] = {0, 0};
  XG_CUDA(cudaMemcpyAsync(counts, words + 1, sizeof(counts), cudaMemcpyDeviceToHost, s));
  XG_CUDA(cudaStreamSynchronize(s));
  *h_num_miss = counts[0];
  *h_num_cache = counts[1];
  return XG_OK;
}

}  // extern "C"
