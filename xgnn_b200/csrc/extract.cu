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
// Here: ONE launch, warp-autonomous (see gather_rows_kernel): lane i of a warp resolves row i of a
// 32-row group (node id -> cache_table -> (part, row) or host -> source pointer), then the warp copies the
// group as a flat list of 16-byte chunks with 8 independent 128-bit loads in flight per lane, so lanes
// are never idle on rows that are not a multiple of 512 B (e.g. 100 x fp32 = 25 chunks).
// Local HBM, NVLink-peer HBM and pinned host rows take the same path; only the pointer differs.
#include "extract.cuh"

#include <cub/cub.cuh>

#include <cstdlib>
#include <mutex>

namespace xg {

namespace {

constexpr int kGatherBlock = 256;
constexpr int kGatherUnroll = 8;

template <int VW> struct Vec;
template <> struct Vec<16> { using T = int4; };
template <> struct Vec<8> { using T = int2; };
template <> struct Vec<4> { using T = int; };
template <> struct Vec<2> { using T = short; };
template <> struct Vec<1> { using T = char; };

// Feature rows are touched once per batch and the batch's rows (GBs) dwarf the 126 MB L2: mark them
// evict-first so the stream does not push the hash table and the hot CSR lines of the sample phase out of L2.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ int4 ld_stream_hint(const int4 *p, uint64_t pol) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.s32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ void st_stream_hint(int4 *p, const int4 &v, uint64_t pol) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.s32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}
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

// node id -> cache-table value.  With `replica_words` (one {membership bits, exclusive popcount} pair per 32 node ids, 8 B x V / 32:
// 28 MB for 111 M nodes, i.e. L2-resident, where the 4 B x V table is not and costs one 128-byte DRAM line per lookup)
// a node of the replicated tier gets its row by rank/select without touching the table; everything else falls through
// to the table.  The replica is laid out by node id, so select == table value (checked when the words are built).
__device__ __forceinline__ uint32_t lookup_rank(const xg_feature_store &st, uint32_t idx) {
  if (st.replica_words != nullptr) {
    // default L2 priority; the rows themselves stream through with evict_first, so the words stay resident
    const uint2 w = __ldg(reinterpret_cast<const uint2 *>(st.replica_words) + (idx >> 5));
    const uint32_t bit = idx & 31u;
    const uint32_t before = w.y + __popc(w.x & ((1u << bit) - 1u));  // replicated nodes with a smaller id
    if ((w.x >> bit) & 1u) return before;
    // every node cached and the sharded tier in node-id order too: its rank is the count of NON-replicated smaller ids
    if (st.replica_words_cover_all) return st.num_replicated + (idx - before);
  }
  return __ldg(st.cache_table + idx);
}

// cache-table value -> (part, row).  Default: DeviceDistFeature::Get (cuda/dist_graph.h:203-207), part r % P, row r / P.
// compact_bitwidth != 0: DeviceICS22SongDistFeature::_GetRealPartId (cuda/ics22_song_dist_feature.h:38-47), the value
// carries the owning GPU in its top bits and the row in the rest.
__device__ __forceinline__ void split_rank(const xg_feature_store &st, uint32_t rank, uint32_t *part, uint32_t *row) {
  if (st.compact_bitwidth) {
    const uint32_t shift = 32u - st.compact_bitwidth;
    *part = rank >> shift;
    *row = rank & ((1u << shift) - 1u);
  } else {
    *part = rank % st.num_part;
    *row = rank / st.num_part;
  }
}

// SAMGRAPH_EMPTY_FEAT mock features (cuda_extraction.cu:51-72, cuda_cache_manager_device.cu:589-592): the host table
// has only 2^k rows and every host-row index is masked to it; 0 = no mask
__device__ __forceinline__ uint32_t host_row(const xg_feature_store &st, uint32_t idx) {
  return st.host_row_mask ? (idx & st.host_row_mask) : idx;
}

__device__ __forceinline__ const char *route_row(const xg_feature_store &st, int mode, uint32_t idx,
                                                 int *klass) {
  uint32_t rank = kEmpty;
  if (mode == kCacheRank) rank = idx;
  else if (mode == kByNode && st.cache_table != nullptr) rank = lookup_rank(st, idx);
  if (rank == kEmpty) {
    *klass = 2;  // host (or the flat table when nothing is cached)
    return static_cast<const char *>(st.host) + (size_t)host_row(st, idx) * st.row_bytes;
  }
  if (rank < st.num_replicated) {
    *klass = 0;
    return static_cast<const char *>(st.replica) + (size_t)rank * st.row_bytes;
  }
  rank -= st.num_replicated;
  uint32_t part, row;
  split_rank(st, rank, &part, &row);
  *klass = ((st.local_part_mask >> part) & 1u) ? 0 : 1;
  return static_cast<const char *>(st.part[part]) + (size_t)row * st.row_bytes;
}

// One warp owns 32 consecutive rows at a time (a "group"); nothing in the kernel crosses a warp, so
// there is no shared memory and no __syncthreads (the first version staged a CTA tile's row pointers in
// shared memory and lost 40 % of its stall cycles on the barriers around that, ncu r01).
//   lane i resolves row i of the group: index -> cache_table -> (part, row) | host -> source pointer;
//   the 32 rows are then one flat list of 16-byte chunks (chunk f -> row f / cpr, col f % cpr), copied
//   with kGatherUnroll independent 128-bit loads in flight per lane before the matching stores; the
//   source / destination row pointers travel by warp shuffle.
// The index and cache-table loads of the NEXT groups are issued before the current group is copied
// (two-deep software pipeline), so the dependent index -> rank -> row chain never stalls the copy.
template <int VW, int U, int MINB>
__global__ void __launch_bounds__(kGatherBlock, MINB)
gather_rows_kernel(const xg_feature_store st, int mode, const uint32_t *__restrict__ src_idx,
                   const uint32_t *__restrict__ dst_idx, size_t num_rows_max,
                   const uint32_t *__restrict__ d_num_rows, uint32_t cpr /* chunks per row */,
                   char *__restrict__ out, unsigned long long *stats) {
  using V = typename Vec<VW>::T;
  uint32_t num_rows = (uint32_t)num_rows_max;  // launch_gather checks num_rows_max < 2^32 - 2^20
  if (d_num_rows && *d_num_rows < num_rows) num_rows = *d_num_rows;  // count still on the device
  const unsigned lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t num_warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t num_groups = (num_rows + 31) / 32;
  const uint32_t q = 32u / cpr, rem = 32u % cpr;  // chunk f + 32 -> (row + q, col + rem) with one carry
  const bool by_table = (mode == kByNode) && st.cache_table != nullptr;

  // pipeline registers: idx of group g+1 (arrived or in flight), rank of group g
  uint32_t g = warp;
  uint32_t idx_cur = 0, rank_cur = kEmpty, idx_nxt = 0;
  if (g < num_groups) {
    const uint32_t r = g * 32 + lane;
    if (r < num_rows) {
      idx_cur = src_idx[r];
      if (by_table) rank_cur = lookup_rank(st, idx_cur);
    }
    const uint64_t rn = (uint64_t)(g + num_warps) * 32 + lane;
    if (rn < num_rows) idx_nxt = src_idx[rn];
  }
  for (; g < num_groups; g += num_warps) {
    const uint32_t row0 = g * 32;
    const uint32_t rows = (num_rows - row0) < 32 ? (num_rows - row0) : 32;
    // ---- issue the look-ahead loads: rank of group g+1, index of group g+2 ----
    uint32_t rank_nxt = kEmpty, idx_nn = 0;
    {
      const uint64_t rn = (uint64_t)(g + num_warps) * 32 + lane;
      if (by_table && rn < num_rows) rank_nxt = lookup_rank(st, idx_nxt);
      const uint64_t rnn = ((uint64_t)g + 2ull * num_warps) * 32 + lane;
      if (rnn < num_rows) idx_nn = src_idx[rnn];
    }
    // ---- route my row ----
    const char *sp = nullptr;
    char *dp = nullptr;
    int klass = 2;
    if (lane < rows) {
      uint32_t rank = kEmpty;
      if (mode == kCacheRank) rank = idx_cur;
      else if (by_table) rank = rank_cur;
      if (rank == kEmpty) {
        sp = static_cast<const char *>(st.host) + (size_t)host_row(st, idx_cur) * st.row_bytes;
      } else if (rank < st.num_replicated) {
        klass = 0;
        sp = static_cast<const char *>(st.replica) + (size_t)rank * st.row_bytes;
      } else {
        rank -= st.num_replicated;
        uint32_t part, row;
        split_rank(st, rank, &part, &row);
        klass = ((st.local_part_mask >> part) & 1u) ? 0 : 1;
        sp = static_cast<const char *>(st.part[part]) + (size_t)row * st.row_bytes;
      }
      const size_t drow = dst_idx ? (size_t)dst_idx[row0 + lane] : row0 + lane;
      dp = out + drow * st.row_bytes;
    }
    if (stats) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const unsigned m = __ballot_sync(0xffffffffu, lane < rows && klass == c);
        if (lane == 0 && m) atomicAdd(stats + c, (unsigned long long)__popc(m));
      }
    }
    // ---- flat chunk copy ----
    const uint32_t total = rows * cpr;
    uint32_t row = lane / cpr, col = lane - row * cpr;  // chunk `lane`
    for (uint32_t f0 = 0; f0 < total; f0 += 32 * U) {
      V buf[U];
      uint32_t lrow = row, lcol = col;  // the load loop walks a copy; the store loop re-walks (fewer registers)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t f = f0 + u * 32 + lane;
        const unsigned long long s64 =
            __shfl_sync(0xffffffffu, (unsigned long long)reinterpret_cast<uintptr_t>(sp), lrow & 31);
        if (f < total) buf[u] = ld_stream(reinterpret_cast<const V *>(s64) + lcol);
        lrow += q;
        lcol += rem;
        if (lcol >= cpr) { lcol -= cpr; ++lrow; }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t f = f0 + u * 32 + lane;
        const unsigned long long d64 =
            __shfl_sync(0xffffffffu, (unsigned long long)reinterpret_cast<uintptr_t>(dp), row & 31);
        if (f < total) st_stream(reinterpret_cast<V *>(d64) + col, buf[u]);
        row += q;
        col += rem;
        if (col >= cpr) { col -= cpr; ++row; }
      }
    }
    idx_cur = idx_nxt;
    rank_cur = rank_nxt;
    idx_nxt = idx_nn;
  }
}

// ---- low-footprint variant: rows staged through a per-warp shared-memory ring with cp.async -------------
// Same routing and the same flat 16-byte-chunk walk as gather_rows_kernel, but the bytes in flight sit in
// shared memory (LDGSTS: global -> smem without a register round trip) instead of in registers:
//   * one CTA of 8 warps per SM keeps kRingStages x 2 KB per warp = 96 KB per SM in flight (Little: 6.5 TB/s x
//     ~800 ns / 148 SMs = 35 KB per SM), with ~40 registers per thread;
//   * so the SM's register file and warp slots stay free for the NEXT batch's sample phase, which the engine
//     runs concurrently on its own stream (pipeline.cu) - the register-staged kernel needs the whole register
//     file (4 CTAs x 256 threads x 64 registers) to cover the same latency and therefore runs alone.
// Two warp-uniform cursors walk the warp's groups: `issue` starts stage copies, `drain` (kRingStages - 1
// stages behind) moves finished stages smem -> global with 128-bit LDS/STG.  A stage is 128 chunks (4 per
// lane) of one 32-row group; the last stage of a group may be partial.
constexpr int kRingWarps = 8;
constexpr int kStageChunks = 128;  // 2 KB

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc, uint64_t pol) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "l"(pol) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

struct GroupCursor {   // warp-uniform position + this lane's walker for chunk (stage * 128 + u * 32 + lane)
  uint32_t g;          // group index
  uint32_t stage;      // stage within the group
  uint32_t rows;       // rows in the group (<= 32)
  uint32_t row, col;   // walker: row / column (in chunks) of this lane's next chunk
};

template <int STAGES>
__global__ void __launch_bounds__(kRingWarps * 32)
gather_rows_ring_kernel(const xg_feature_store st, int mode, const uint32_t *__restrict__ src_idx,
                        const uint32_t *__restrict__ dst_idx, size_t num_rows_max,
                        const uint32_t *__restrict__ d_num_rows, uint32_t cpr, char *__restrict__ out) {
  extern __shared__ int4 ring_all[];
  const uint64_t pol = l2_evict_first_policy();
  uint32_t num_rows = (uint32_t)num_rows_max;
  if (d_num_rows && *d_num_rows < num_rows) num_rows = *d_num_rows;
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  int4 *ring = ring_all + (size_t)wib * STAGES * kStageChunks;
  const uint32_t warp = blockIdx.x * kRingWarps + wib;
  const uint32_t num_warps = gridDim.x * kRingWarps;
  const uint32_t num_groups = (num_rows + 31) / 32;
  const uint32_t q = 32u / cpr, rem = 32u % cpr;
  const bool by_table = (mode == kByNode) && st.cache_table != nullptr;
  const uint32_t lane_row0 = lane / cpr, lane_col0 = lane - lane_row0 * cpr;

  auto group_rows = [&](uint32_t g) { return (num_rows - g * 32) < 32 ? (num_rows - g * 32) : 32u; };
  auto route = [&](uint32_t idx, uint32_t rank_in) -> const char * {
    uint32_t rank = kEmpty;
    if (mode == kCacheRank) rank = idx;
    else if (by_table) rank = rank_in;
    if (rank == kEmpty) return static_cast<const char *>(st.host) + (size_t)host_row(st, idx) * st.row_bytes;
    if (rank < st.num_replicated) return static_cast<const char *>(st.replica) + (size_t)rank * st.row_bytes;
    rank -= st.num_replicated;
    uint32_t part, row;
    split_rank(st, rank, &part, &row);
    return static_cast<const char *>(st.part[part]) + (size_t)row * st.row_bytes;
  };
  auto dst_of = [&](uint32_t g, uint32_t rows) -> char * {
    if (lane >= rows) return nullptr;
    const size_t drow = dst_idx ? (size_t)dst_idx[g * 32 + lane] : (size_t)g * 32 + lane;
    return out + drow * st.row_bytes;
  };

  GroupCursor is{warp, 0, 0, lane_row0, lane_col0}, dr = is;
  if (warp >= num_groups) return;
  is.rows = dr.rows = group_rows(warp);
  // look-ahead registers of the issue side: index of the next group, rank of the current one
  uint32_t idx_cur = 0, rank_cur = kEmpty, idx_nxt = 0, rank_nxt = kEmpty;
  {
    const uint32_t r = warp * 32 + lane;
    if (r < num_rows) {
      idx_cur = src_idx[r];
      if (by_table) rank_cur = lookup_rank(st, idx_cur);
    }
    const uint64_t rn = (uint64_t)(warp + num_warps) * 32 + lane;
    if (rn < num_rows) {
      idx_nxt = src_idx[rn];
      if (by_table) rank_nxt = lookup_rank(st, idx_nxt);
    }
  }
  const char *sp = lane < is.rows ? route(idx_cur, rank_cur) : nullptr;
  char *dp = dst_of(dr.g, dr.rows);
  bool issue_more = true;
  uint32_t inflight = 0, slot_i = 0, slot_d = 0;

  while (true) {
    // ---- issue stages until the ring is full ----
    while (issue_more && inflight < (uint32_t)STAGES) {
      const uint32_t total = is.rows * cpr;
      int4 *stage = ring + (size_t)slot_i * kStageChunks;
#pragma unroll
      for (int u = 0; u < kStageChunks / 32; ++u) {
        const uint32_t f = is.stage * kStageChunks + u * 32 + lane;
        const unsigned long long s64 =
            __shfl_sync(0xffffffffu, (unsigned long long)reinterpret_cast<uintptr_t>(sp), is.row & 31);
        if (f < total) cp_async16(stage + u * 32 + lane, reinterpret_cast<const int4 *>(s64) + is.col, pol);
        is.row += q;
        is.col += rem;
        if (is.col >= cpr) { is.col -= cpr; ++is.row; }
      }
      cp_async_commit();
      ++inflight;
      slot_i = slot_i + 1 == (uint32_t)STAGES ? 0 : slot_i + 1;
      if ((++is.stage) * kStageChunks >= total) {  // group finished: move to the next one
        is.g += num_warps;
        if (is.g >= num_groups) {
          issue_more = false;
        } else {
          is.stage = 0;
          is.rows = group_rows(is.g);
          is.row = lane_row0;
          is.col = lane_col0;
          sp = lane < is.rows ? route(idx_nxt, rank_nxt) : nullptr;
          const uint64_t rn = (uint64_t)(is.g + num_warps) * 32 + lane;  // look-ahead for the group after
          idx_nxt = 0;
          rank_nxt = kEmpty;
          if (rn < num_rows) {
            idx_nxt = src_idx[rn];
            if (by_table) rank_nxt = lookup_rank(st, idx_nxt);
          }
        }
      }
    }
    if (inflight == 0) break;
    // ---- drain the oldest stage ----
    if (inflight == (uint32_t)STAGES) cp_async_wait<STAGES - 1>();
    else cp_async_wait<0>();
    __syncwarp();
    {
      const uint32_t total = dr.rows * cpr;
      const int4 *stage = ring + (size_t)slot_d * kStageChunks;
      int4 v[kStageChunks / 32];
#pragma unroll
      for (int u = 0; u < kStageChunks / 32; ++u) v[u] = stage[u * 32 + lane];
#pragma unroll
      for (int u = 0; u < kStageChunks / 32; ++u) {
        const uint32_t f = dr.stage * kStageChunks + u * 32 + lane;
        const unsigned long long d64 =
            __shfl_sync(0xffffffffu, (unsigned long long)reinterpret_cast<uintptr_t>(dp), dr.row & 31);
        if (f < total) st_stream_hint(reinterpret_cast<int4 *>(d64) + dr.col, v[u], pol);
        dr.row += q;
        dr.col += rem;
        if (dr.col >= cpr) { dr.col -= cpr; ++dr.row; }
      }
      __syncwarp();  // every lane has read the stage before the issue side overwrites it
      --inflight;
      slot_d = slot_d + 1 == (uint32_t)STAGES ? 0 : slot_d + 1;
      if ((++dr.stage) * kStageChunks >= total) {
        dr.g += num_warps;
        if (dr.g < num_groups) {
          dr.stage = 0;
          dr.rows = group_rows(dr.g);
          dr.row = lane_row0;
          dr.col = lane_col0;
          dp = dst_of(dr.g, dr.rows);
        }
      }
    }
  }
}

// ---- bulk variant: the rows never touch a register (TMA engine in, TMA engine out) ---------------------------------
// For stores whose rows are mostly REMOTE (pure rank % P sharding: (P - 1) / P of the rows come over NVLink).  Measured
// with ncu byte counters (profiles/r02n_nvlink_requests.md): a row fetched by 32 lanes x 16 bytes costs four 128-byte
// read requests = 0.25 request bytes per row byte on the link; one cp.async.bulk per row costs 0.1875, and because every
// GPU of an all-pull gather both receives rows and carries the requests of the rows it serves, that is 650-664 GB/s of
// rows per direction instead of 612-624.  On local rows the same pipeline is ~8 % slower than the cp.async ring (2.72
// vs 2.96 TB/s of rows), so launch_gather picks it only when nearly all rows are remote (mostly_remote below).
//   one warp = one CTA; a stage = up to 32 rows (<= 16 KB); lane i resolves row i of the group (index -> rank ->
//   pointer, one step AHEAD of the copy), asks the engine for it (cp.async.bulk global -> shared, one mbarrier per
//   stage counts the bytes) and, when the stage has landed, hands it back (cp.async.bulk shared -> global).
constexpr int kBulkStages = 4;
constexpr int kBulkStageBytes = 16384;

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(32)
gather_rows_bulk_kernel(const xg_feature_store st, int mode, const uint32_t *__restrict__ src_idx,
                        const uint32_t *__restrict__ dst_idx, size_t num_rows_max, const uint32_t *__restrict__ d_num_rows,
                        uint32_t rows_per_stage, char *__restrict__ out) {
  extern __shared__ __align__(128) char bulk_smem[];
  uint64_t *bars = reinterpret_cast<uint64_t *>(bulk_smem + (size_t)kBulkStages * kBulkStageBytes);
  uint32_t num_rows = (uint32_t)num_rows_max;
  if (d_num_rows && *d_num_rows < num_rows) num_rows = *d_num_rows;
  const unsigned lane = threadIdx.x;
  const uint32_t rb = (uint32_t)st.row_bytes;
  const uint32_t num_groups = (num_rows + rows_per_stage - 1) / rows_per_stage;
  const uint32_t first = blockIdx.x, stride = gridDim.x;
  if (first >= num_groups) return;
  const uint32_t mine = (num_groups - first + stride - 1) / stride;
  if (lane == 0)
    for (int s = 0; s < kBulkStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bars + s)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();

  auto rows_of = [&](uint32_t g) { return min(rows_per_stage, num_rows - g * rows_per_stage); };
  auto resolve = [&](uint32_t k) -> const char * {  // source pointer of this lane's row of my k-th group
    if (k >= mine) return nullptr;
    const uint32_t g = first + k * stride;
    if (lane >= rows_of(g)) return nullptr;
    int klass;
    return route_row(st, mode, src_idx[g * rows_per_stage + lane], &klass);
  };
  auto issue = [&](uint32_t k, const char *src) {
    const uint32_t s = k % kBulkStages, g = first + k * stride;
    if (lane == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bars + s)), "r"(rows_of(g) * rb) : "memory");
    __syncwarp();
    if (src != nullptr)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_addr(bulk_smem + (size_t)s * kBulkStageBytes + (size_t)lane * rb)),
                   "l"(src), "r"(rb), "r"(smem_addr(bars + s))
                   : "memory");
  };
  for (uint32_t k = 0; k < (uint32_t)kBulkStages && k < mine; ++k) issue(k, resolve(k));
  const char *next_src = resolve(kBulkStages);  // pointer of the first refill, in flight during the first wait
  for (uint32_t k = 0; k < mine; ++k) {
    const uint32_t s = k % kBulkStages, parity = (k / kBulkStages) & 1u, g = first + k * stride;
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(smem_addr(bars + s)), "r"(parity) : "memory");
    if (lane < rows_of(g)) {
      const size_t r = (size_t)g * rows_per_stage + lane;
      const size_t drow = dst_idx ? (size_t)dst_idx[r] : r;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + drow * rb),
                   "r"(smem_addr(bulk_smem + (size_t)s * kBulkStageBytes + (size_t)lane * rb)), "r"(rb)
                   : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // the stage handed back one step ago has been read
    __syncwarp();
    if (k >= 1 && k - 1 + kBulkStages < mine) {  // refill that stage: group k - 1 + kBulkStages, resolved one step ago
      issue(k - 1 + kBulkStages, next_src);
      next_src = resolve(k + kBulkStages);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// mostly-remote store: sharded, no replicated tier, at most one part in eight local.  Measured (papers100M shape, pure
// rank % P sharding, profiles/r02t_bulk_ab_*.txt): 8 GPUs (87.5 % of the rows remote) bulk 3.332 ms vs ring 3.377 per
// gather, 3.79 vs 3.90 ms per minibatch (the bulk kernel keeps 96 threads per SM busy instead of 768, which the sample
// phase beside it gets); 2 GPUs (50 % remote) bulk 2.089 vs ring 1.986 ms - the local half pays the bulk pipeline's 8 %.
bool mostly_remote(const xg_feature_store &st, int mode) {
  if (mode == kHostRow || st.num_part < 2 || st.num_replicated != 0) return false;
  uint32_t local = 0;
  for (uint32_t p = 0; p < st.num_part; ++p) local += (st.local_part_mask >> p) & 1u;
  return 8 * local <= st.num_part;
}

bool aligned_to(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

int pick_vector_width(const xg_feature_store &st, const void *out) {
  for (int vw = 16; vw > 1; vw >>= 1) {
    bool ok = (st.row_bytes % vw == 0) && aligned_to(out, vw) && (st.host == nullptr || aligned_to(st.host, vw));
    for (uint32_t p = 0; ok && p < st.num_part; ++p) ok = st.part[p] == nullptr || aligned_to(st.part[p], vw);
    if (ok && st.num_replicated) ok = aligned_to(st.replica, vw);
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
  XG_REQUIRE(num_rows < 0xFFF00000ull, "too many rows for one launch");
  XG_REQUIRE(mode != kCacheRank || st.num_part > 0 || st.num_replicated > 0, "cache-rank gather needs a cache");
  XG_REQUIRE(st.num_replicated == 0 || st.replica != nullptr, "num_replicated > 0 but replica is NULL");
  XG_REQUIRE(mode == kCacheRank || st.host != nullptr || st.cache_table != nullptr, "no source table");
  const int vw = pick_vector_width(st, out);
  const uint64_t cpr64 = st.row_bytes / vw;
  XG_REQUIRE(cpr64 < (1u << 24), "row too long");
  const uint32_t cpr = (uint32_t)cpr64;
  // one warp per 32-row group, grid-stride; 8 CTAs of 256 threads per SM when registers allow
  const size_t groups = (num_rows + 31) / 32;
  const size_t ctas = (groups + kGatherBlock / 32 - 1) / (kGatherBlock / 32);
  auto *o = static_cast<char *>(out);
  auto *st64 = reinterpret_cast<unsigned long long *>(stats);
#define XG_GATHER(VW, U, MINB)                                                                       \
  gather_rows_kernel<VW, U, MINB><<<grid_for(ctas, MINB), kGatherBlock, 0, s>>>(st, mode, src_idx, dst_idx, \
                                                                               num_rows, d_num_rows, cpr, o, st64)
  static const int variant = [] {
    const char *e = std::getenv("XG_GATHER_VARIANT");
    return e ? std::atoi(e) : 2;  // 2 = shared-memory ring (default), 0 / 1 = register-staged <16,4,4> / <16,8,2>
  }();
  if (vw == 16 && stats == nullptr && st.row_bytes <= (size_t)kBulkStageBytes &&
      (variant == 3 || (variant == 2 && mostly_remote(st, mode)))) {
    // XG_GATHER_VARIANT=3 forces the bulk kernel for every store, =4 never uses it
    int dev = 0;
    XG_CUDA(cudaGetDevice(&dev));
    const size_t smem = (size_t)kBulkStages * kBulkStageBytes + kBulkStages * sizeof(uint64_t);
    static bool bulk_attr[64] = {};
    if (!bulk_attr[dev & 63]) {
      XG_REQUIRE(cudaFuncSetAttribute(gather_rows_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) ==
                     cudaSuccess, "cannot reserve shared memory for the bulk gather");
      bulk_attr[dev & 63] = true;
    }
    uint32_t rps = (uint32_t)(kBulkStageBytes / st.row_bytes);
    if (rps > 32) rps = 32;
    const size_t stage_groups = (num_rows + rps - 1) / rps;
    gather_rows_bulk_kernel<<<grid_for(stage_groups, 3), 32, smem, s>>>(st, mode, src_idx, dst_idx, num_rows, d_num_rows, rps, o);
    XG_LAUNCHED();
    return XG_OK;
  }
  if (vw == 16 && variant >= 2 && stats == nullptr) {
    // shared-memory ring: XG_GATHER_RING_CTAS CTAs per SM (default 3), XG_GATHER_RING_STAGES x 2 KB per warp
    // (default 4): 192 KB per SM in flight.  Measured on B200, papers100M-shape batch: alone the gather takes
    // 0.84-0.96 ms for every setting with >= 48 KB per SM in flight (profiles/r01d_sweep*.txt), but with the next
    // batch's sample phase running beside it - and, at N > 1, with a tenth to a fifth of the rows coming over NVLink
    // at ~3x the latency - the deeper budget wins: r02, 1 GPU 1.190 -> 1.155 ms/minibatch, 2 GPUs 8.4 -> 8.9-9.0 G
    // edges/s for (3 CTAs, 4 stages), (4, 3) and (2, 6) alike (all 192 KB) against (2, 3) = 96 KB.
    static const int ring_ctas = [] {
      const char *e = std::getenv("XG_GATHER_RING_CTAS");
      const int v = e ? std::atoi(e) : 3;
      return v < 1 ? 1 : (v > 4 ? 4 : v);
    }();
    static const int ring_stages = [] {
      const char *e = std::getenv("XG_GATHER_RING_STAGES");
      return e ? std::atoi(e) : 4;
    }();
    int ring_dev = 0;
    XG_CUDA(cudaGetDevice(&ring_dev));
    const size_t ring_ctas_needed = (groups + kRingWarps - 1) / kRingWarps;
    const unsigned grid = grid_for(ring_ctas_needed, ring_ctas);
#define XG_RING(S)                                                                                                \
  do {                                                                                                            \
    const size_t smem = (size_t)kRingWarps * S * kStageChunks * sizeof(int4);                                     \
    static bool attr_set[64] = {}; /* per DEVICE: a process may drive several GPUs */                            \
    if (!attr_set[ring_dev & 63]) {                                                                               \
      XG_REQUIRE(cudaFuncSetAttribute(gather_rows_ring_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                      (int)smem) == cudaSuccess, "cannot reserve shared memory for the ring gather"); \
      attr_set[ring_dev & 63] = true;                                                                             \
    }                                                                                                             \
    gather_rows_ring_kernel<S><<<grid, kRingWarps * 32, smem, s>>>(st, mode, src_idx, dst_idx, num_rows,          \
                                                                   d_num_rows, cpr, o);                           \
  } while (0)
    switch (ring_stages) {
      case 2: XG_RING(2); break;
      case 3: XG_RING(3); break;
      case 4: XG_RING(4); break;
      case 6: XG_RING(6); break;
      case 8: XG_RING(8); break;
      default: XG_RING(4); break;
    }
#undef XG_RING
    XG_LAUNCHED();
    return XG_OK;
  }
  switch (vw) {
    case 16:
      // measured on B200 (papers100M-shape batch, 4.9 M x 512 B rows): U=4/4 CTAs 0.925 ms, U=8/2 CTAs 0.920,
      // U=8/3 CTAs 0.933 (spills), U=6/4 0.951 (spills), U=4/6 0.983 (spills); r01 shared-memory version 1.050
      if (variant == 1) XG_GATHER(16, 8, 2);
      else XG_GATHER(16, 4, 4);
      break;
    case 8: XG_GATHER(8, 8, 3); break;
    case 4: XG_GATHER(4, 8, 3); break;
    case 2: XG_GATHER(2, 8, 3); break;
    default: XG_GATHER(1, 8, 3); break;
  }
#undef XG_GATHER
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

int xg_mock_extract(void *d_dst, const void *src, const uint32_t *d_index, size_t num_index, size_t row_bytes,
                    unsigned empty_feat_bits, void *stream) {
  XG_REQUIRE(src != nullptr || num_index == 0, "src is NULL");
  XG_REQUIRE(empty_feat_bits < 32, "empty_feat_bits must be below 32");
  xg_feature_store st{};
  st.host = src;
  st.row_bytes = row_bytes;
  st.host_row_mask = empty_feat_bits ? (uint32_t)((1ull << empty_feat_bits) - 1) : 0u;
  return launch_gather(st, kHostRow, d_index, nullptr, num_index, nullptr, d_dst, nullptr, as_stream(stream));
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

// ---- rank/select words of the replicated tier ------------------------------------------------------
namespace xg {
namespace {

__global__ void __launch_bounds__(256)
replica_bits_kernel(const uint32_t *__restrict__ table, uint32_t num_node, uint32_t num_replicated, uint32_t *bits,
                    uint32_t *counts, uint32_t num_words) {
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < num_words; w += (gridDim.x * blockDim.x) >> 5) {
    const uint32_t v = w * 32 + lane;
    const bool in = v < num_node && table[v] < num_replicated;
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, in);
    if (lane == 0) {
      bits[w] = b;
      counts[w] = __popc(b);
    }
  }
}

__global__ void __launch_bounds__(256)
replica_words_kernel(const uint32_t *__restrict__ table, uint32_t num_node, const uint32_t *__restrict__ bits,
                     const uint32_t *__restrict__ prefix, uint2 *words, uint32_t num_words, uint32_t num_replicated,
                     unsigned long long *mismatch) {
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < num_words; w += (gridDim.x * blockDim.x) >> 5) {
    const uint32_t b = bits[w], p = prefix[w], v = w * 32 + lane;
    if (lane == 0) words[w] = make_uint2(b, p);
    // select must reproduce the table: the replicated tier has to be laid out by node id
    const uint32_t before = p + __popc(b & ((1u << lane) - 1u));
    if ((b >> lane) & 1u) {
      if (table[v] != before) atomicAdd(mismatch, 1ull);
    } else if (v < num_node && table[v] != num_replicated + (v - before)) {
      atomicAdd(mismatch + 1, 1ull);  // not "all cached, sharded tier by node id": the table stays in use for these
    }
  }
}

}  // namespace
}  // namespace xg

extern "C" int xg_build_replica_words(const uint32_t *d_cache_table, uint32_t num_node, uint32_t num_replicated,
                                      void *d_words, uint32_t *h_cover_all, void *stream) {
  XG_REQUIRE(d_cache_table && d_words, "NULL argument");
  if (num_node == 0) return XG_OK;
  cudaStream_t s = as_stream(stream);
  const uint32_t num_words = (num_node + 31) / 32;
  uint32_t *bits = nullptr, *counts = nullptr, *prefix = nullptr;
  unsigned long long *mismatch = nullptr, h_mismatch[2] = {0, 0};
  void *temp = nullptr;
  size_t temp_bytes = 0;
  // three u32 arrays of num_words, then (8-byte aligned) the two 64-bit mismatch counters
  const size_t counter_word = ((size_t)num_words * 3 + 1) & ~(size_t)1;
  XG_CUDA(cudaMalloc(&bits, counter_word * sizeof(uint32_t) + sizeof(unsigned long long) * 2));
  counts = bits + num_words;
  prefix = counts + num_words;
  mismatch = reinterpret_cast<unsigned long long *>(bits + counter_word);
  cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, counts, prefix, (int)num_words, s);
  if (e == cudaSuccess) e = cudaMalloc(&temp, temp_bytes ? temp_bytes : 16);
  if (e == cudaSuccess) e = cudaMemsetAsync(mismatch, 0, 2 * sizeof(unsigned long long), s);
  if (e == cudaSuccess) {  // every exit below goes through the common cleanup: nothing returns with temp / bits allocated
    const unsigned grid = grid_for(((size_t)num_words * 32 + 255) / 256, 8);
    replica_bits_kernel<<<grid, 256, 0, s>>>(d_cache_table, num_node, num_replicated, bits, counts, num_words);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(temp, temp_bytes, counts, prefix, (int)num_words, s);
    if (e == cudaSuccess) {
      replica_words_kernel<<<grid, 256, 0, s>>>(d_cache_table, num_node, bits, prefix, static_cast<uint2 *>(d_words), num_words,
                                                num_replicated, mismatch);
      g_launches.fetch_add(1, std::memory_order_relaxed);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_mismatch, mismatch, sizeof(h_mismatch), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  }
  cudaFree(temp);
  cudaFree(bits);
  XG_CUDA(e);
  XG_REQUIRE(h_mismatch[0] == 0, "the replicated tier is not laid out by node id (cache_table[v] != rank of v among replicated nodes)");
  if (h_cover_all) *h_cover_all = h_mismatch[1] == 0 ? 1u : 0u;
  return XG_OK;
}

// ---- GetMissCacheIndex without any allocation ------------------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  The scan scratch of split_miss_cache_kernel - one ticket word plus one look-back word per tile - lives in a
// file-scope __device__ array: every GPU that loads the library holds XG_NOALLOC_MAX_NODES / 1024 + 4 words of 8 bytes
// (128 KB) until exit.  ALL calls on a device share it and it keeps the last call's contents, so the part a call uses
// is cleared on the caller's stream first; two calls that overlap on the GPU (different streams, no ordering) race.
namespace xg {
namespace {

constexpr size_t kStaticSplitTiles = (size_t)XG_NOALLOC_MAX_NODES / kSplitTile + 1;
static __device__ __align__(256) uint64_t g_static_split_scratch[kStaticSplitTiles + 3];  // [0] ticket, [1..2] unused, [3..] scan

constexpr int kStaticSplitDevices = 64;
std::mutex g_static_split_mu;
uint64_t *g_static_split_ptr[kStaticSplitDevices];  // device address of the array, per device (host code cannot use its name)

// once per device, with that device current: the symbol's device address
int static_split_scratch(uint64_t **out) {
  int dev = 0;
  XG_CUDA(cudaGetDevice(&dev));
  XG_REQUIRE(dev >= 0 && dev < kStaticSplitDevices, "device ordinal beyond the static scratch table");
  std::lock_guard<std::mutex> lock(g_static_split_mu);
  if (!g_static_split_ptr[dev]) {
    void *p = nullptr;
    XG_CUDA(cudaGetSymbolAddress(&p, g_static_split_scratch));
    XG_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned NULL for the static split scratch");
    g_static_split_ptr[dev] = static_cast<uint64_t *>(p);
  }
  *out = g_static_split_ptr[dev];
  return XG_OK;
}

}  // namespace
}  // namespace xg

extern "C" int xg_get_miss_cache_index_noalloc(const uint32_t *d_cache_table, const uint32_t *d_nodes, size_t num_nodes,
                                               uint32_t *d_miss_src, uint32_t *d_miss_dst, uint32_t *d_cache_src,
                                               uint32_t *d_cache_dst, uint64_t *d_counts, void *stream) {
  XG_REQUIRE(d_cache_table && d_counts, "NULL argument");
  if (num_nodes > (size_t)XG_NOALLOC_MAX_NODES) {
    set_error("invalid argument: xg_get_miss_cache_index_noalloc: num_nodes = " + std::to_string(num_nodes) +
              " exceeds XG_NOALLOC_MAX_NODES = " + std::to_string((size_t)XG_NOALLOC_MAX_NODES) +
              " (the static scratch is sized for that; use xg_get_miss_cache_index)");
    return XG_ERR_INVALID;
  }
  XG_REQUIRE(num_nodes == 0 || (d_nodes && d_miss_src && d_miss_dst && d_cache_src && d_cache_dst), "NULL argument");
  cudaStream_t s = as_stream(stream);
  if (num_nodes == 0) {
    XG_CUDA(cudaMemsetAsync(d_counts, 0, 2 * sizeof(uint64_t), s));
    return XG_OK;
  }
  uint64_t *words = nullptr;
  if (int rc = static_split_scratch(&words)) return rc;
  const size_t tiles = (num_nodes + kSplitTile - 1) / kSplitTile + 1;
  XG_CUDA(cudaMemsetAsync(words, 0, (tiles + 3) * sizeof(uint64_t), s));  // ticket + this call's scan words start at zero
  split_miss_cache_kernel<<<grid_for(tiles, 8), kSplitBlock, 0, s>>>(
      d_cache_table, d_nodes, num_nodes, d_miss_src, d_miss_dst, d_cache_src, d_cache_dst,
      reinterpret_cast<unsigned long long *>(d_counts), reinterpret_cast<uint32_t *>(words), words + 3);
  XG_LAUNCHED();
  return XG_OK;
}
