// Uniform k-hop neighbour sampling over a flat or GGMS-sharded CSR.
//
// Reference behaviour being replaced: GPUSampleKHop0 (samgraph/common/cuda/cuda_sampling_khop0.cu
// :102-335), GPUSampleKHop3 (cuda_sampling_khop3.cu:41-318), GPUSampleKHop2 (cuda_sampling_khop2.cu
// :46-375) and the graph handles DeviceNormalGraph / DeviceDistGraph (cuda/dist_graph.h:114-180).
// The reference needs sample -> count_edge -> DeviceScan -> compact_edge (4 launches, 5 host syncs,
// a padded num_input*fanout staging buffer written and re-read) plus a curandState pool.
//
// Here ONE kernel per layer does it:
//   phase A (thread per node)   route the node (v % P, v / P or host slot), read its two indptr
//                               words, count = min(deg, fanout); CTA scan + chained look-back scan
//                               across tiles gives every node its final (compacted) edge offset.
//   phase B (G lanes per node)  deg <= fanout: copy the list in CSR order (bit-exact contract);
//                               deg >  fanout: Floyd's subset sampling on *positions* -- exactly
//                               `fanout` distinct positions, uniform over all subsets, reading only
//                               the chosen neighbours (KHop3's idea without its shared-memory set,
//                               atomics or shared RNG state).  Counter-based Philox, no state.
//                               Optionally (pipeline) each neighbour goes straight into the hash table.
// Edge order: node-major in input order, slot-minor (cuda_sampling_khop0.cu:192-239 semantics).
// Remote lists (peer HBM over NVLink, pinned host over PCIe) are read with plain loads through the
// pointers in the by-value xg_graph handle.
#include <cstdlib>

#include "sampling.cuh"

namespace xg {

namespace {

// plain variant: 8 CTAs/SM (32 registers) - the big layer is bound by random-access throughput and every resident
// warp counts; pipelined variant: registers for 4 nodes in flight instead
template <int G, int VPL, bool FUSED, bool PIPE>
__global__ void __launch_bounds__(kSampleBlock, VPL > 1 ? 1 : (PIPE ? 5 : 8))
sample_uniform_kernel(const xg_graph g, const uint32_t *__restrict__ input, uint32_t num_input_max,
                      const uint32_t *__restrict__ d_num_input, uint32_t fanout,
                      uint32_t *__restrict__ out_src, uint32_t *__restrict__ out_dst,
                      uint64_t *__restrict__ d_num_out, uint2 rng_key, const uint2 *__restrict__ d_rng_key,
                      uint32_t salt, SampleFusion fusion, uint32_t *ticket, uint64_t *scan_state) {
  static_assert(kSampleBlock == kSampleTile, "phase A maps one thread to one node");
  __shared__ const uint32_t *s_edges[kSampleTile];
  __shared__ uint32_t s_deg[kSampleTile];
  __shared__ uint32_t s_off[kSampleTile];
  __shared__ uint32_t s_src[kSampleTile];
  __shared__ uint32_t s_warp[kSampleBlock / 32];
  __shared__ uint32_t s_tile;
  __shared__ uint64_t s_tile_base;

  uint32_t n = d_num_input ? *d_num_input : num_input_max;
  if (n > num_input_max) n = num_input_max;
  if (d_rng_key) rng_key = *d_rng_key;  // per-batch key lives on the device when the batch is a replayed graph
  const uint32_t num_tiles = (n + kSampleTile - 1) / kSampleTile;
  if (num_tiles == 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *d_num_out = 0;
    return;
  }
  const unsigned lane32 = threadIdx.x & 31;
  const unsigned grp = threadIdx.x / G;         // group index inside the CTA
  const unsigned gl = threadIdx.x % G;          // lane inside the group
  const unsigned grp_shift = (lane32 / G) * G;  // first lane of my group inside the warp
  constexpr unsigned kGroups = kSampleBlock / G;
  constexpr uint32_t kGroupMask = (G == 32) ? 0xffffffffu : ((1u << G) - 1u);

  while (true) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= num_tiles) break;

    // ---------------- phase A: one thread per input node -----------------------------------
    const uint32_t i = tile * kSampleTile + threadIdx.x;
    uint32_t deg = 0, v = 0;
    const uint32_t *edges = nullptr;
    if (i < n) {
      v = input[i];
      const NodeRoute r = route(g, v);
      const uint32_t b = r.indptr[r.row];
      const uint32_t e = r.indptr[r.row + 1];
      deg = e - b;
      edges = r.indices + b;
    }
    const uint32_t cnt = deg < fanout ? deg : fanout;
    uint32_t tile_total;
    const uint32_t excl = block_exclusive_scan<kSampleBlock>(cnt, s_warp, &tile_total);
    s_edges[threadIdx.x] = edges;
    s_deg[threadIdx.x] = deg;
    s_off[threadIdx.x] = excl;
    if (FUSED) s_src[threadIdx.x] = (fusion.src_local && i < n) ? fusion.src_local[i] : i;
    else s_src[threadIdx.x] = v;
    if (threadIdx.x < 32) {
      const uint64_t before = lookback_exclusive(scan_state, tile, tile_total);
      if (threadIdx.x == 0) {
        s_tile_base = before;
        if (tile == num_tiles - 1) *d_num_out = before + tile_total;
      }
    }
    __syncthreads();
    const uint64_t tile_base = s_tile_base;
    if (FUSED && fusion.csc_indptr != nullptr) {
      // edges are node-major in input order, so the compacted offset of node i is indptr[i] of the layer's CSC
      // (destination-major) form: the caller can build the DGL block without a COO -> CSC conversion
      if (i < n) fusion.csc_indptr[i] = (uint32_t)(tile_base + excl);
      if (i == n - 1) fusion.csc_indptr[n] = (uint32_t)(tile_base + excl + cnt);
    }

    // ---------------- phase B: G lanes per node ---------------------------------------------
    if constexpr (VPL == 1 && PIPE) {
      // fanout <= G: a lane owns at most ONE edge of a node (slot gl).  The tile's nodes are walked kPipe at a
      // time in two sweeps - positions + neighbour loads, then the stores - so every lane has kPipe independent
      // DRAM loads in flight instead of one dependent chain per node (that chain, 8-16 times per tile, was 70 % of
      // the kernel: ncu r01b long-scoreboard stalls).  Same Philox counters and draws as the plain loop.
      // Used for layers of at most one wave of tiles (seed-side layers: latency bound); the big input-side layer
      // is bound by random-sector DRAM throughput, where extra registers only cost occupancy, so it keeps the
      // plain loop.
      constexpr int kPipe = 4;
      static_assert((kSampleTile / kGroups) % kPipe == 0, "tile iterations must be a multiple of the pipeline depth");
      for (unsigned j0 = grp; j0 < kSampleTile; j0 += kGroups * kPipe) {
        uint32_t nb[kPipe];
        bool live[kPipe];
#pragma unroll
        for (int q = 0; q < kPipe; ++q) {  // sweep 1: choose the position, issue the neighbour load
          const unsigned j = j0 + q * kGroups;
          const uint32_t ndeg = s_deg[j];
          const uint32_t *nedges = s_edges[j];
          const bool big = ndeg > fanout;
          uint32_t sel = gl;
          if (__any_sync(0xffffffffu, big)) {
            const uint32_t node_pos = tile * kSampleTile + j;
            const uint4 r = Philox::rand4(rng_key, make_uint4(node_pos, gl, salt, 0u));
            const uint32_t jbase = big ? (ndeg - fanout) : 0u;
            const uint32_t draw = bounded(r.x, jbase + gl + 1u);
            uint32_t mine_sel = kEmpty;
            for (unsigned jl = 0; jl < (unsigned)G; ++jl) {
              if (jl >= fanout) break;  // uniform across the warp
              const uint32_t t = __shfl_sync(0xffffffffu, draw, jl, G);
              const unsigned hits = (__ballot_sync(0xffffffffu, mine_sel == t) >> grp_shift) & kGroupMask;
              if (gl == jl) mine_sel = hits ? (ndeg - fanout + jl) : t;
            }
            if (big) sel = mine_sel;
          }
          live[q] = gl < (big ? fanout : ndeg);
          nb[q] = live[q] ? nedges[sel] : 0u;
        }
#pragma unroll
        for (int q = 0; q < kPipe; ++q) {  // sweep 2: write the edges
          if (!live[q]) continue;
          const unsigned j = j0 + q * kGroups;
          const uint64_t e = tile_base + s_off[j] + gl;
          out_dst[e] = nb[q];
          out_src[e] = s_src[j];
        }
      }
    } else {
      for (unsigned j = grp; j < kSampleTile; j += kGroups) {
        const uint32_t ndeg = s_deg[j];
        const uint32_t *nedges = s_edges[j];
        const uint64_t obase = tile_base + s_off[j];
        const uint32_t src = s_src[j];
        const bool big = ndeg > fanout;

        uint32_t sel[VPL];  // chosen positions (big) for slot k*G + gl
        if (__any_sync(0xffffffffu, big)) {
          // Floyd: for slot s = 0..fanout-1, J = deg - fanout + s, t ~ U[0, J]; take t unless already
          // taken, else take J.  Draws are made lane-parallel up front, conflicts resolved in order.
          uint32_t draw[VPL];
          {
            const uint32_t node_pos = tile * kSampleTile + j;
            const uint4 r = Philox::rand4(rng_key, make_uint4(node_pos, gl, salt, 0u));
            const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
            const uint32_t j0 = big ? (ndeg - fanout) : 0u;
  #pragma unroll
            for (int k = 0; k < VPL; ++k) {
              uint32_t rk = rr[k & 3];
              if (VPL > 4 && k >= 4) {  // more than four slots per lane: extra Philox blocks
                const uint4 r2 = Philox::rand4(rng_key, make_uint4(node_pos, gl, salt, (uint32_t)(k >> 2)));
                const uint32_t rr2[4] = {r2.x, r2.y, r2.z, r2.w};
                rk = rr2[k & 3];
              }
              draw[k] = bounded(rk, j0 + (uint32_t)(k * G) + gl + 1u);
              sel[k] = kEmpty;
            }
          }
  #pragma unroll
          for (int k = 0; k < VPL; ++k) {
            for (unsigned jl = 0; jl < (unsigned)G; ++jl) {
              const uint32_t s = (uint32_t)k * G + jl;
              if (s >= fanout) break;  // uniform across the warp
              const uint32_t t = __shfl_sync(0xffffffffu, draw[k], jl, G);
              bool mine = false;
  #pragma unroll
              for (int kk = 0; kk <= k; ++kk) mine |= (sel[kk] == t);  // unset slots hold kEmpty != t
              const unsigned hits = (__ballot_sync(0xffffffffu, mine) >> grp_shift) & kGroupMask;
              if (gl == jl) sel[k] = hits ? (ndeg - fanout + s) : t;
            }
          }
        }

        if (big) {
  #pragma unroll
          for (int k = 0; k < VPL; ++k) {
            const uint32_t s = (uint32_t)k * G + gl;
            if (s < fanout) {
              const uint32_t nbr = nedges[sel[k]];
              const uint64_t e = obase + s;
              out_dst[e] = nbr;
              out_src[e] = src;
            }
          }
        } else {
          for (uint32_t s = gl; s < ndeg; s += G) {
            const uint32_t nbr = nedges[s];
            const uint64_t e = obase + s;
            out_dst[e] = nbr;
            out_src[e] = src;
          }
        }
      }
    }
    __syncthreads();  // shared staging is reused by the next tile
  }
}

template <int G, int VPL>
int launch_variant(const xg_graph &g, const uint32_t *d_input, size_t num_input_max,
                   const uint32_t *d_num_input, size_t fanout, uint32_t *d_out_src, uint32_t *d_out_dst,
                   uint64_t *d_num_out, uint2 key, const uint2 *d_key, uint32_t salt,
                   const SampleFusion &fusion, const ScanSpace &scan, cudaStream_t s) {
  const size_t tiles = (num_input_max + kSampleTile - 1) / kSampleTile;
  static const int ctas_per_sm = [] {
    // 8 x 256 threads = full occupancy at 32 registers; measured on B200 (0.9 M nodes -> 4.3 M edges): 4 CTAs/SM
    // 0.316 ms, 6: 0.255, 8: 0.229, 12 (more tiles than residency): 0.229
    const char *e = std::getenv("XG_SAMPLE_CTAS");
    return e ? std::atoi(e) : 8;
  }();
  const unsigned grid = grid_for(tiles, ctas_per_sm);
  static const int force_pipe = [] {
    const char *e = std::getenv("XG_SAMPLE_PIPE");  // experiment knob: 1 = always pipelined, 0 = never
    return e ? std::atoi(e) : -1;
  }();
  const bool pipe = VPL == 1 && (force_pipe >= 0 ? force_pipe == 1 : tiles <= (size_t)kNumSM * 6);  // one wave at the pipelined variant's residency
#define XG_LAUNCH_SAMPLE(F, P)                                                                              \
  sample_uniform_kernel<G, VPL, F, P><<<grid, kSampleBlock, 0, s>>>(                                        \
      g, d_input, (uint32_t)num_input_max, d_num_input, (uint32_t)fanout, d_out_src, d_out_dst, d_num_out, \
      key, d_key, salt, fusion, scan.ticket, scan.state)
  if (fusion.enabled) {
    if (pipe) XG_LAUNCH_SAMPLE(true, true);
    else XG_LAUNCH_SAMPLE(true, false);
  } else {
    if (pipe) XG_LAUNCH_SAMPLE(false, true);
    else XG_LAUNCH_SAMPLE(false, false);
  }
#undef XG_LAUNCH_SAMPLE
  XG_LAUNCHED();
  return XG_OK;
}

}  // namespace

int launch_sample_uniform(const xg_graph &g, const uint32_t *d_input, size_t num_input_max,
                          const uint32_t *d_num_input, size_t fanout, uint32_t *d_out_src,
                          uint32_t *d_out_dst, uint64_t *d_num_out, uint64_t seed, uint64_t batch_key,
                          const uint2 *d_rng_key, uint32_t salt, const SampleFusion &fusion,
                          const ScanSpace &scan, cudaStream_t s) {
  XG_REQUIRE(fanout >= 1 && fanout <= 512, "fanout must be in [1, 512]");
  XG_REQUIRE(num_input_max < (1ull << 31), "too many input nodes");
  XG_REQUIRE(num_input_max * fanout < (1ull << 31), "num_input * fanout must stay below 2^31");
  const size_t tiles = (num_input_max + kSampleTile - 1) / kSampleTile;
  XG_REQUIRE(tiles <= scan.max_tiles, "scan workspace too small");
  if (num_input_max == 0) {
    XG_CUDA(cudaMemsetAsync(d_num_out, 0, sizeof(uint64_t), s));
    return XG_OK;
  }
  const uint2 key = philox_key(seed, batch_key);
#define XG_SAMPLE(G, VPL)                                                                         \
  return launch_variant<G, VPL>(g, d_input, num_input_max, d_num_input, fanout, d_out_src, d_out_dst, \
                                d_num_out, key, d_rng_key, salt, fusion, scan, s)
  if (fanout <= 4) XG_SAMPLE(4, 1);
  if (fanout <= 8) XG_SAMPLE(8, 1);
  if (fanout <= 16) XG_SAMPLE(16, 1);
  if (fanout <= 32) XG_SAMPLE(32, 1);
  if (fanout <= 64) XG_SAMPLE(32, 2);
  if (fanout <= 128) XG_SAMPLE(32, 4);
  if (fanout <= 256) XG_SAMPLE(32, 8);
  XG_SAMPLE(32, 16);
#undef XG_SAMPLE
}

}  // namespace xg

// =================================================================================================
// C-ABI
// =================================================================================================
using namespace xg;

namespace {
struct SampleWorkspace {  // per-device scratch for the standalone operator
  void *scan_mem = nullptr;
  size_t scan_tiles = 0;
};
SampleWorkspace g_ws[XG_MAX_PARTS * 4];
}  // namespace

extern "C" int xg_sample_khop(const xg_graph *g, int sample_type, const uint32_t *d_input,
                              size_t num_input, size_t fanout, uint32_t *d_out_src,
                              uint32_t *d_out_dst, size_t *d_num_out, uint64_t seed,
                              uint64_t batch_key, void *stream) {
  XG_REQUIRE(g != nullptr, "graph handle is NULL");
  XG_REQUIRE(sample_type == XG_KHOP0 || sample_type == XG_KHOP2 || sample_type == XG_KHOP3,
             "xg_sample_khop handles KHop0 / KHop2 / KHop3");
  XG_REQUIRE(d_num_out != nullptr, "d_num_out is NULL");
  XG_REQUIRE(g->num_part <= XG_MAX_PARTS, "too many partitions");
  int dev = 0;
  XG_CUDA(cudaGetDevice(&dev));
  SampleWorkspace &ws = g_ws[dev % (XG_MAX_PARTS * 4)];
  const size_t tiles = (num_input + kSampleTile - 1) / kSampleTile + 1;
  cudaStream_t s = as_stream(stream);
  if (ws.scan_tiles < tiles) {
    if (ws.scan_mem) {
      XG_CUDA(cudaStreamSynchronize(s));
      XG_CUDA(cudaFree(ws.scan_mem));
    }
    size_t want = tiles + tiles / 2 + 64;
    XG_CUDA(cudaMalloc(&ws.scan_mem, (want + 1) * sizeof(uint64_t)));
    ws.scan_tiles = want;
  }
  XG_CUDA(cudaMemsetAsync(ws.scan_mem, 0, (tiles + 1) * sizeof(uint64_t), s));
  ScanSpace scan{reinterpret_cast<uint32_t *>(ws.scan_mem), reinterpret_cast<uint64_t *>(ws.scan_mem) + 1,
                 ws.scan_tiles};
  SampleFusion none{};
  none.enabled = false;
  static_assert(sizeof(size_t) == sizeof(uint64_t), "size_t must be 64-bit");
  return launch_sample_uniform(*g, d_input, num_input, nullptr, fanout, d_out_src, d_out_dst,
                               reinterpret_cast<uint64_t *>(d_num_out), seed, batch_key, nullptr, kSampleSalt, none,
                               scan, s);
}
