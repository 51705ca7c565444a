// Random-walk neighbourhood sampler with per-seed top-K by visit count (PinSAGE), SURVEY.md row a7.
//
// Reference behaviour being replaced: GPUSampleRandomWalk<GraphType> (samgraph/common/cuda/
// cuda_sampling_random_walk.cu:43-165) + FrequencyHashmap::GetTopK (cuda/cuda_frequency_hashmap.cu:643-841):
// a walk kernel writing N * W * Lw (seed, visited) pairs to global memory, then 8 more kernels, three
// device scans, a 64-bit radix sort of every unique (seed, node) pair and a 12-byte-per-slot global hash
// table with 128 slots per seed, each step followed by a host synchronisation.
//
// Contract kept: per seed, in input order, W walks of at most Lw steps; a walk
// ends at a node without neighbours or, after any step, with probability restart_prob; every visited node
// (the start node can be one) is counted per seed; the min(K, #distinct) most visited are emitted,
// count-descending, as (src = seed, dst = node, data = count).  Ties: the reference leaves them to a race
// (which thread's CAS created the bucket); here the node visited FIRST in the reference's own visit layout
// (step-major, walk-minor: cuda_sampling_random_walk.cu:66-80) wins, which is the order a sequential
// execution of the reference produces.
//
// How: W * Lw is a dozen to a few hundred visits, so the whole per-seed pipeline fits in one warp and
// shared memory: lanes walk, the warp counts duplicates and ranks the distinct nodes by (count desc, first
// visit asc) with O(M^2 / 32) register compares, a CTA stages 64 seeds' results and one chained scan places
// them.  One launch, no global scratch, no sort.
#include "sampling.cuh"

namespace xg {
namespace {

constexpr int kRwBlock = 256;
constexpr int kRwWarps = kRwBlock / 32;
constexpr int kRwTile = 64;  // seeds per tile (8 per warp)

__global__ void __launch_bounds__(kRwBlock)
random_walk_topk_kernel(const xg_graph g, const uint32_t *__restrict__ input, uint32_t n_max,
                        const uint32_t *__restrict__ d_n, uint32_t walk_len, double restart_prob, uint32_t num_walk,
                        uint32_t K, bool emit_local, const uint32_t *__restrict__ src_local,
                        uint32_t *__restrict__ out_src, uint32_t *__restrict__ out_dst,
                        uint32_t *__restrict__ out_data, uint64_t *__restrict__ d_num_out, uint2 rng_key,
                        const uint2 *__restrict__ d_rng_key, uint32_t salt, uint32_t *ticket, uint64_t *scan_state) {
  extern __shared__ uint32_t smem[];
  const uint32_t M = walk_len * num_walk;
  uint32_t *s_vis = smem;                         // [kRwWarps][M]  visited node per (step, walk)
  uint32_t *s_cnt = s_vis + kRwWarps * M;         // [kRwWarps][M]  count at first visits, 0 elsewhere
  uint32_t *s_dst = s_cnt + kRwWarps * M;         // [kRwTile][K]
  uint32_t *s_data = s_dst + kRwTile * K;         // [kRwTile][K]
  __shared__ uint32_t s_nout[kRwTile], s_off[kRwTile], s_warp[kRwBlock / 32];
  __shared__ uint32_t s_tile;
  __shared__ uint64_t s_base;

  uint32_t n = d_n ? *d_n : n_max;
  if (n > n_max) n = n_max;
  if (d_rng_key) rng_key = *d_rng_key;
  const uint32_t num_tiles = (n + kRwTile - 1) / kRwTile;
  if (num_tiles == 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *d_num_out = 0;
    return;
  }
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t *vis = s_vis + warp * M, *cnt = s_cnt + warp * M;

  while (true) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= num_tiles) break;

    for (uint32_t q = warp; q < kRwTile; q += kRwWarps) {
      const uint32_t i = tile * kRwTile + q;
      uint32_t distinct = 0;
      if (i < n) {
        const uint32_t start = input[i];
        // ---- walks: lane = walk ----
        for (uint32_t w = lane; w < num_walk; w += 32) {
          uint32_t node = start;
          for (uint32_t t = 0; t < walk_len; ++t) {
            uint32_t seen = kEmpty;
            if (node != kEmpty) {
              const NodeRoute r = route(g, node);
              const uint32_t off = r.indptr[r.row];
              const uint32_t len = r.indptr[r.row + 1] - off;
              if (len == 0) {
                node = kEmpty;
              } else {
                const uint4 rnd = Philox::rand4(rng_key, make_uint4(i, w, t, salt));
                seen = r.indices[off + rnd.x % len];
                node = (uniform_double(rnd.y) < restart_prob) ? kEmpty : seen;
              }
            }
            vis[t * num_walk + w] = seen;
          }
        }
        __syncwarp();
        // ---- visit counts; only the first visit of a node keeps its count ----
        for (uint32_t m0 = 0; m0 < M; m0 += 32) {
          const uint32_t m = m0 + lane;
          uint32_t c = 0;
          if (m < M) {
            const uint32_t v = vis[m];
            if (v != kEmpty) {
              bool first = true;
              for (uint32_t o = 0; o < M; ++o) {
                const bool same = vis[o] == v;
                c += same;
                first &= !(same && o < m);
              }
              if (!first) c = 0;
            }
            cnt[m] = c;
          }
          distinct += __popc(__ballot_sync(0xffffffffu, c != 0));
        }
        __syncwarp();
        // ---- rank the distinct nodes: count descending, first visit ascending ----
        for (uint32_t m0 = 0; m0 < M; m0 += 32) {
          const uint32_t m = m0 + lane;
          if (m < M) {
            const uint32_t c = cnt[m];
            if (c != 0) {
              uint32_t rank = 0;
              for (uint32_t o = 0; o < M; ++o) {
                const uint32_t co = cnt[o];
                rank += (co > c) || (co == c && o < m);
              }
              if (rank < K) {
                s_dst[q * K + rank] = vis[m];
                s_data[q * K + rank] = c;
              }
            }
          }
        }
        __syncwarp();
      }
      if (lane == 0) s_nout[q] = distinct < K ? distinct : K;
    }
    __syncthreads();
    // ---- place the tile's results ----
    uint32_t tile_total;
    const uint32_t mine = threadIdx.x < kRwTile ? s_nout[threadIdx.x] : 0u;
    const uint32_t excl = block_exclusive_scan<kRwBlock>(mine, s_warp, &tile_total);
    if (threadIdx.x < kRwTile) s_off[threadIdx.x] = excl;
    if (threadIdx.x < 32) {
      const uint64_t before = lookback_exclusive(scan_state, tile, tile_total);
      if (threadIdx.x == 0) {
        s_base = before;
        if (tile == num_tiles - 1) *d_num_out = before + tile_total;
      }
    }
    __syncthreads();
    const uint64_t base = s_base;
    for (uint32_t e = threadIdx.x; e < kRwTile * K; e += kRwBlock) {
      const uint32_t q = e / K, r = e - q * K;
      if (r < s_nout[q]) {
        const uint32_t i = tile * kRwTile + q;
        const uint64_t o = base + s_off[q] + r;
        out_src[o] = emit_local ? (src_local ? src_local[i] : i) : input[i];
        out_dst[o] = s_dst[e];
        out_data[o] = s_data[e];
      }
    }
    __syncthreads();
  }
}

}  // namespace

int launch_random_walk(const xg_graph &g, const uint32_t *d_input, size_t num_input_max, const uint32_t *d_num_input,
                       uint32_t walk_length, double restart_prob, uint32_t num_walk, uint32_t K, bool emit_local,
                       const uint32_t *src_local, uint32_t *d_out_src, uint32_t *d_out_dst, uint32_t *d_out_data,
                       uint64_t *d_num_out, uint2 rng_key, const uint2 *d_rng_key, uint32_t salt,
                       const ScanSpace &scan, cudaStream_t s) {
  XG_REQUIRE(walk_length >= 1 && num_walk >= 1 && K >= 1, "walk_length, num_walk and K must be positive");
  XG_REQUIRE((uint64_t)walk_length * num_walk <= 512, "random_walk_length * num_random_walk must be <= 512");
  XG_REQUIRE(K <= 64, "num_neighbor (K) must be <= 64");
  XG_REQUIRE(num_input_max < (1ull << 31) / K, "too many input nodes");
  if (num_input_max == 0) {
    XG_CUDA(cudaMemsetAsync(d_num_out, 0, sizeof(uint64_t), s));
    return XG_OK;
  }
  const size_t tiles = (num_input_max + kRwTile - 1) / kRwTile;
  XG_REQUIRE(tiles <= scan.max_tiles, "scan workspace too small");
  const size_t smem = ((size_t)2 * kRwWarps * walk_length * num_walk + (size_t)2 * kRwTile * K) * sizeof(uint32_t);
  if (smem > 48 * 1024) {
    // the attribute is per DEVICE: a process that drives several GPUs must raise it on each of them
    static bool raised[64] = {};
    int dev = 0;
    XG_CUDA(cudaGetDevice(&dev));
    if (!raised[dev & 63]) {
      XG_CUDA(cudaFuncSetAttribute(random_walk_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      raised[dev & 63] = true;
    }
  }
  random_walk_topk_kernel<<<grid_for(tiles, 4), kRwBlock, smem, s>>>(
      g, d_input, (uint32_t)num_input_max, d_num_input, walk_length, restart_prob, num_walk, K, emit_local, src_local,
      d_out_src, d_out_dst, d_out_data, d_num_out, rng_key, d_rng_key, salt, scan.ticket, scan.state);
  XG_LAUNCHED();
  return XG_OK;
}

}  // namespace xg

using namespace xg;

namespace {
struct RwWorkspace {
  void *scan_mem = nullptr;
  size_t scan_tiles = 0;
};
RwWorkspace g_rw_ws[XG_MAX_PARTS * 4];
}  // namespace

extern "C" int xg_sample_random_walk(const xg_graph *g, const uint32_t *d_input, size_t num_input,
                                     size_t random_walk_length, double random_walk_restart_prob,
                                     size_t num_random_walk, size_t K, uint32_t *d_out_src, uint32_t *d_out_dst,
                                     uint32_t *d_out_data, size_t *d_num_out, uint64_t seed, uint64_t batch_key,
                                     void *stream) {
  XG_REQUIRE(g && d_num_out, "NULL argument");
  XG_REQUIRE(g->num_part <= XG_MAX_PARTS, "too many partitions");
  int dev = 0;
  XG_CUDA(cudaGetDevice(&dev));
  RwWorkspace &ws = g_rw_ws[dev % (XG_MAX_PARTS * 4)];
  cudaStream_t s = as_stream(stream);
  const size_t tiles = (num_input + kRwTile - 1) / kRwTile + 1;
  if (ws.scan_tiles < tiles) {
    if (ws.scan_mem) {
      XG_CUDA(cudaStreamSynchronize(s));
      XG_CUDA(cudaFree(ws.scan_mem));
    }
    const size_t want = tiles + tiles / 2 + 64;
    XG_CUDA(cudaMalloc(&ws.scan_mem, (want + 1) * sizeof(uint64_t)));
    ws.scan_tiles = want;
  }
  XG_CUDA(cudaMemsetAsync(ws.scan_mem, 0, (tiles + 1) * sizeof(uint64_t), s));
  ScanSpace scan{reinterpret_cast<uint32_t *>(ws.scan_mem), reinterpret_cast<uint64_t *>(ws.scan_mem) + 1, ws.scan_tiles};
  return launch_random_walk(*g, d_input, num_input, nullptr, (uint32_t)random_walk_length, random_walk_restart_prob,
                            (uint32_t)num_random_walk, (uint32_t)K, false, nullptr, d_out_src, d_out_dst, d_out_data,
                            reinterpret_cast<uint64_t *>(d_num_out), philox_key(seed, batch_key), nullptr, kSampleSalt,
                            scan, s);
}
