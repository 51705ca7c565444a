// KHop1, weighted KHop (alias / prefix-sum / hash-dedup) samplers.
//
// Reference behaviour being replaced:
//   GPUSampleKHop1                cuda/cuda_sampling_khop1.cu:42-236
//   GPUSampleWeightedKHop         cuda/cuda_sampling_weighted_khop.cu:41-76,132-
//   GPUSampleWeightedKHopPrefix   cuda/cuda_sampling_weighted_khop_prefix.cu:41-92,148-
//   GPUSampleWeightedKHopHashDedup cuda/cuda_sampling_weighted_khop_hash_dedup.cu:40-116,203-
// Contract kept:
//   KHop1 / weighted / prefix : `fanout` draws WITH replacement per input node; the (src, dst) pairs are
//     stably sorted by the GLOBAL src id (draw order survives inside a src), then only ADJACENT equal
//     pairs collapse (the last of a run is kept); nodes without neighbours contribute nothing.
//   hash-dedup : node-major in input order; deg <= fanout -> all neighbours; otherwise alias draws are
//     rejected until `fanout` DISTINCT ids were seen (bounded here: 16 * fanout draws, the reference
//     spins forever when fewer distinct outcomes exist).
// How it is done here: the reference radix-sorts all N * fanout pairs (64-bit of payload each) after
// drawing.  Sorting the N input ids FIRST (radix_sort_pairs below over N (id, position) pairs, fanout x
// fewer keys) and drawing in sorted order yields exactly the sorted pair list, because the sort key is
// the src only and the sort is stable.  Draws are a pure function of (input position, slot): counter-
// based Philox, so the result does not depend on the sort, the grid or the schedule.  The adjacent-
// dedup compaction is one kernel with a chained look-back scan instead of count + DeviceScan + compact.
#include "sampling.cuh"

namespace xg {
namespace {

__device__ __forceinline__ uint32_t resolve_count(size_t n_max, const uint32_t *d_n) {
  uint32_t n = d_n ? *d_n : (uint32_t)n_max;
  return n < n_max ? n : (uint32_t)n_max;
}

// keys[i] = input id (EMPTY beyond the live count so that the sort parks those at the end), vals[i] = i
__global__ void __launch_bounds__(256)
sort_prep_kernel(const uint32_t *__restrict__ input, size_t n_max, const uint32_t *__restrict__ d_n,
                 uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
  const uint32_t n = resolve_count(n_max, d_n);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_max; i += (size_t)gridDim.x * blockDim.x) {
    keys[i] = i < n ? input[i] : kEmpty;
    vals[i] = (uint32_t)i;
  }
}

// ---- stable LSD radix sort of (id, position) pairs: 9-bit digits, three small kernels per pass ------------------------
// (r01 called cub::DeviceRadixSort here.)  N is one layer's input count (<= ~1.4 M on the benchmark shapes) and the keys
// are node ids, so ceil(bits(num_node) / 9) passes do (3 for every shape up to 134 M nodes).  What matters is that the
// sort is exact and STABLE - equal ids keep input order, which is what makes the drawing order "ascending global src,
// draw order inside" (cuda_sampling_khop1.cu:42-236).  Only the first n (device-side count) elements take part.
//   hist      : one CTA per tile of kRadixTile consecutive elements -> hist[bin][tile]   (bin-major)
//   bin scan  : one CTA per bin: exclusive scan of its row over the tiles (coalesced) + the bin's total
//   scatter   : tile -> output.  Warp w owns elements [256 w, 256 w + 256) of the tile, 32 at a time; an element's slot
//               = first slot of (bin, tile) + elements of the bin in earlier warps + in earlier rounds of my warp
//               + earlier lanes of my round (__match_any_sync) - i.e. its rank in input order.
constexpr int kRadixBits = 9;
constexpr int kRadixBins = 1 << kRadixBits;
constexpr int kRadixBlock = kRadixBins;  // thread t owns bin t in the bookkeeping steps
constexpr int kRadixWarps = kRadixBlock / 32;
constexpr int kRadixRounds = 8;          // per warp: 8 x 32 elements
constexpr int kRadixTile = kRadixBlock * kRadixRounds;

__global__ void __launch_bounds__(kRadixBlock)
radix_hist_kernel(const uint32_t *__restrict__ keys, size_t n_max, const uint32_t *__restrict__ d_n, int shift,
                  uint32_t *__restrict__ hist, uint32_t tiles) {
  __shared__ uint32_t h[kRadixBins];
  const uint32_t n = resolve_count(n_max, d_n);
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * kRadixTile;
#pragma unroll
  for (int r = 0; r < kRadixRounds; ++r) {
    const uint32_t i = base + r * kRadixBlock + threadIdx.x;
    if (i < n) atomicAdd(&h[(keys[i] >> shift) & (kRadixBins - 1)], 1u);
  }
  __syncthreads();
  hist[(size_t)threadIdx.x * tiles + blockIdx.x] = h[threadIdx.x];
}

__global__ void __launch_bounds__(256)
radix_bin_scan_kernel(uint32_t *__restrict__ hist, uint32_t tiles, uint32_t *__restrict__ bin_total) {
  __shared__ uint32_t s_warp[8];
  uint32_t *row = hist + (size_t)blockIdx.x * tiles;
  uint32_t carry = 0;
  for (uint32_t t0 = 0; t0 < tiles; t0 += 256) {
    const uint32_t t = t0 + threadIdx.x;
    const uint32_t v = t < tiles ? row[t] : 0u;
    uint32_t total;
    const uint32_t excl = block_exclusive_scan<256>(v, s_warp, &total);
    if (t < tiles) row[t] = carry + excl;
    carry += total;
  }
  if (threadIdx.x == 0) bin_total[blockIdx.x] = carry;
}

__global__ void __launch_bounds__(kRadixBlock)
radix_scatter_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, size_t n_max,
                     const uint32_t *__restrict__ d_n, int shift, const uint32_t *__restrict__ offs,
                     const uint32_t *__restrict__ bin_total, uint32_t tiles, uint32_t *__restrict__ keys_out,
                     uint32_t *__restrict__ vals_out) {
  __shared__ uint32_t s_cnt[kRadixWarps][kRadixBins];  // counts, then first output slot, of (warp, bin)
  __shared__ uint32_t s_warp[kRadixWarps];
  const uint32_t n = resolve_count(n_max, d_n);
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t base = blockIdx.x * kRadixTile;
  if (base >= n) return;  // uniform: the CTA's tile is beyond the live elements
  uint32_t total_unused;
  const uint32_t bin_base = block_exclusive_scan<kRadixBlock>(bin_total[threadIdx.x], s_warp, &total_unused);
#pragma unroll
  for (int w = 0; w < kRadixWarps; ++w) s_cnt[w][threadIdx.x] = 0;
  __syncthreads();
  uint32_t key[kRadixRounds], val[kRadixRounds];
  // ---- the warp's histogram of its 256 elements ----
#pragma unroll
  for (int r = 0; r < kRadixRounds; ++r) {
    const uint32_t i = base + warp * (32 * kRadixRounds) + r * 32 + lane;
    const bool live = i < n;
    key[r] = live ? keys[i] : 0u;
    val[r] = live ? vals[i] : 0u;
    const uint32_t bin = live ? ((key[r] >> shift) & (kRadixBins - 1)) : kRadixBins;  // dead lanes match each other only
    const unsigned peers = __match_any_sync(0xffffffffu, bin);
    if (live && (peers & ((1u << lane) - 1u)) == 0) s_cnt[warp][bin] += __popc(peers);  // first lane of the digit group
    __syncwarp();
  }
  __syncthreads();
  {  // thread t = bin t: counts -> first output slot of (warp, bin)
    uint32_t run = offs[(size_t)threadIdx.x * tiles + blockIdx.x] + bin_base;
#pragma unroll
    for (int w = 0; w < kRadixWarps; ++w) {
      const uint32_t c = s_cnt[w][threadIdx.x];
      s_cnt[w][threadIdx.x] = run;
      run += c;
    }
  }
  __syncthreads();
  // ---- scatter in input order ----
#pragma unroll
  for (int r = 0; r < kRadixRounds; ++r) {
    const uint32_t i = base + warp * (32 * kRadixRounds) + r * 32 + lane;
    const bool live = i < n;
    const uint32_t bin = live ? ((key[r] >> shift) & (kRadixBins - 1)) : kRadixBins;
    const unsigned peers = __match_any_sync(0xffffffffu, bin);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    uint32_t slot = 0;
    if (live) {
      slot = s_cnt[warp][bin] + rank;
      keys_out[slot] = key[r];
      vals_out[slot] = val[r];
    }
    __syncwarp();
    if (live && rank == 0) s_cnt[warp][bin] += __popc(peers);
    __syncwarp();
  }
}

// sorts the first n (= min(n_max, *d_n)) pairs of (keys_a, vals_a) by key; key_bits = number of significant key bits.
// Ping-pongs a -> b -> a ...; *result_in_b says where the sorted pairs are.
int radix_sort_pairs(uint32_t *keys_a, uint32_t *vals_a, uint32_t *keys_b, uint32_t *vals_b, size_t n_max,
                     const uint32_t *d_n, int key_bits, uint32_t *hist, bool *result_in_b, cudaStream_t s) {
  const uint32_t tiles = (uint32_t)((n_max + kRadixTile - 1) / kRadixTile);
  uint32_t *bin_total = hist + (size_t)tiles * kRadixBins;
  const int passes = (key_bits + kRadixBits - 1) / kRadixBits;
  for (int pass = 0; pass < passes; ++pass) {
    const uint32_t *kin = (pass & 1) ? keys_b : keys_a, *vin = (pass & 1) ? vals_b : vals_a;
    uint32_t *kout = (pass & 1) ? keys_a : keys_b, *vout = (pass & 1) ? vals_a : vals_b;
    radix_hist_kernel<<<tiles, kRadixBlock, 0, s>>>(kin, n_max, d_n, pass * kRadixBits, hist, tiles);
    XG_LAUNCHED();
    radix_bin_scan_kernel<<<kRadixBins, 256, 0, s>>>(hist, tiles, bin_total);
    XG_LAUNCHED();
    radix_scatter_kernel<<<tiles, kRadixBlock, 0, s>>>(kin, vin, n_max, d_n, pass * kRadixBits, hist, bin_total, tiles, kout,
                                                       vout);
    XG_LAUNCHED();
  }
  *result_in_b = (passes & 1) != 0;
  return XG_OK;
}

// One thread per (sorted node k, slot s).  MODE: XG_KHOP1, XG_WEIGHTED_KHOP, XG_WEIGHTED_KHOP_PREFIX.
template <int MODE>
__global__ void __launch_bounds__(256)
draw_with_replacement_kernel(const xg_graph g, const WeightTables w, const uint32_t *__restrict__ ids,
                             const uint32_t *__restrict__ pos, size_t n_max, const uint32_t *__restrict__ d_n,
                             uint32_t fanout, uint32_t *__restrict__ tmp_src, uint32_t *__restrict__ tmp_dst,
                             uint2 rng_key, const uint2 *__restrict__ d_rng_key, uint32_t salt) {
  const uint32_t n = resolve_count(n_max, d_n);
  if (d_rng_key) rng_key = *d_rng_key;
  const uint64_t tasks = (uint64_t)n * fanout;
  for (uint64_t t = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; t < tasks; t += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t k = (uint32_t)(t / fanout), slot = (uint32_t)(t - (uint64_t)k * fanout);
    const uint32_t v = ids[k];
    const NodeRoute r = route(g, v);
    const uint32_t off = r.indptr[r.row];
    const uint32_t len = r.indptr[r.row + 1] - off;
    uint32_t src = kEmpty, dst = kEmpty;
    if (len != 0) {
      const uint4 rnd = Philox::rand4(rng_key, make_uint4(pos[k], slot, salt, 0u));
      src = v;
      if (MODE == XG_KHOP1) {
        dst = r.indices[off + rnd.x % len];  // `curand % len`, cuda_sampling_khop1.cu:66
      } else if (MODE == XG_WEIGHTED_KHOP) {
        const uint32_t j = off + rnd.x % len;  // cuda_sampling_weighted_khop.cu:65-71
        dst = (uniform_open0(rnd.y) < w.prob[j]) ? r.indices[j] : w.alias[j];
      } else {  // inverse CDF: smallest j with prefix[j] >= x (cuda_sampling_weighted_khop_prefix.cu:67-88)
        const float x = uniform_open0(rnd.x) * w.prefix[off + len - 1];
        uint32_t lo = off, hi = off + len - 1;
        if (x <= w.prefix[off]) {
          hi = off;
        } else {
          while (hi - lo >= 2) {
            const uint32_t mid = (uint32_t)(((uint64_t)lo + hi) >> 1);
            if (w.prefix[mid] >= x) hi = mid;
            else lo = mid;
          }
        }
        dst = r.indices[hi];
      }
    }
    tmp_src[t] = src;
    tmp_dst[t] = dst;
  }
}

// Thread per input node, node-major (no sort): cuda_sampling_weighted_khop_hash_dedup.cu:59-116
constexpr int kHashDedupMaxFanout = 64;
__global__ void __launch_bounds__(128)
draw_hash_dedup_kernel(const xg_graph g, const WeightTables w, const uint32_t *__restrict__ input, size_t n_max,
                       const uint32_t *__restrict__ d_n, uint32_t fanout, uint32_t *__restrict__ tmp_src,
                       uint32_t *__restrict__ tmp_dst, uint2 rng_key, const uint2 *__restrict__ d_rng_key,
                       uint32_t salt) {
  const uint32_t n = resolve_count(n_max, d_n);
  if (d_rng_key) rng_key = *d_rng_key;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t v = input[i];
    const NodeRoute r = route(g, v);
    const uint32_t off = r.indptr[r.row];
    const uint32_t len = r.indptr[r.row + 1] - off;
    uint32_t *src = tmp_src + (size_t)i * fanout, *dst = tmp_dst + (size_t)i * fanout;
    uint32_t have = 0;
    if (len <= fanout) {
      for (; have < len; ++have) {
        src[have] = v;
        dst[have] = r.indices[off + have];
      }
    } else {
      uint32_t seen[kHashDedupMaxFanout];
      const uint32_t budget = kHashDedupTriesPerSlot * fanout;
      for (uint32_t tries = 0; have < fanout && tries < budget; ++tries) {
        const uint4 rnd = Philox::rand4(rng_key, make_uint4(i, tries, salt, 1u));
        const uint32_t j = off + rnd.x % len;
        uint32_t cand = r.indices[j];
        if (uniform_open0(rnd.y) > w.prob[j]) cand = w.alias[j];  // `r > prob` here, `r < prob` in the plain sampler
        bool dup = false;
        for (uint32_t q = 0; q < have; ++q) dup |= (seen[q] == cand);
        if (dup) continue;
        seen[have] = cand;
        src[have] = v;
        dst[have] = cand;
        ++have;
      }
    }
    for (; have < fanout; ++have) {
      src[have] = kEmpty;
      dst[have] = kEmpty;
    }
  }
}

// Stable compaction of the padded draw list with one chained scan.  ADJ: keep slot t only if it is the
// LAST of a run of equal (src, dst) pairs (count_edge / compact_edge of cuda_sampling_khop1.cu:74-128);
// otherwise keep every non-padding slot (cuda_sampling_weighted_khop_hash_dedup.cu:118-200).
constexpr int kCompactBlock = 256, kCompactItems = 4, kCompactTile = kCompactBlock * kCompactItems;

template <bool ADJ>
__global__ void __launch_bounds__(kCompactBlock)
compact_pairs_kernel(const uint32_t *__restrict__ tmp_src, const uint32_t *__restrict__ tmp_dst,
                     const uint32_t *__restrict__ node_pos /* sorted k -> input position, or NULL */,
                     uint32_t fanout, size_t n_max, const uint32_t *__restrict__ d_n, bool emit_local,
                     const uint32_t *__restrict__ src_local, uint32_t *__restrict__ out_src,
                     uint32_t *__restrict__ out_dst, uint64_t *__restrict__ d_num_out, uint32_t *ticket,
                     uint64_t *scan_state) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_cnt[kCompactItems * (kCompactBlock / 32)];
  __shared__ uint64_t s_prefix;
  constexpr int kWarps = kCompactBlock / 32;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t total = (uint64_t)resolve_count(n_max, d_n) * fanout;
  const uint32_t num_tiles = (uint32_t)((total + kCompactTile - 1) / kCompactTile);
  if (num_tiles == 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *d_num_out = 0;
    return;
  }
  while (true) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= num_tiles) break;
    const uint64_t base = (uint64_t)tile * kCompactTile;
    uint32_t src[kCompactItems], dst[kCompactItems], before[kCompactItems];
    bool keep[kCompactItems];
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
      const uint64_t t = base + (uint64_t)k * kCompactBlock + threadIdx.x;
      src[k] = kEmpty;
      dst[k] = kEmpty;
      keep[k] = false;
      if (t < total) {
        src[k] = tmp_src[t];
        dst[k] = tmp_dst[t];
        keep[k] = src[k] != kEmpty;
        if (ADJ && keep[k] && t + 1 < total) keep[k] = (src[k] != tmp_src[t + 1]) || (dst[k] != tmp_dst[t + 1]);
      }
      const unsigned b = __ballot_sync(0xffffffffu, keep[k]);
      before[k] = __popc(b & ((1u << lane) - 1));
      if (lane == 0) s_cnt[k * kWarps + warp] = __popc(b);
    }
    __syncthreads();
    if (warp == 0) {
      static_assert(kCompactItems * kWarps == 32, "32 partial counts expected");
      const uint32_t c = s_cnt[lane];
      uint32_t incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (unsigned)o) incl += v;
      }
      const uint32_t sum = __shfl_sync(0xffffffffu, incl, 31);
      s_cnt[lane] = incl - c;
      const uint64_t excl = lookback_exclusive(scan_state, tile, sum);
      if (lane == 0) {
        s_prefix = excl;
        if (tile == num_tiles - 1) *d_num_out = excl + sum;
      }
    }
    __syncthreads();
    const uint64_t tile_out = s_prefix;
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
      if (!keep[k]) continue;
      const uint64_t t = base + (uint64_t)k * kCompactBlock + threadIdx.x;
      const uint64_t o = tile_out + s_cnt[k * kWarps + warp] + before[k];
      uint32_t s_out = src[k];
      if (emit_local) {
        const uint32_t node = (uint32_t)(t / fanout);
        const uint32_t p = node_pos ? node_pos[node] : node;
        s_out = src_local ? src_local[p] : p;
      }
      out_src[o] = s_out;
      out_dst[o] = dst[k];
    }
    __syncthreads();
  }
}

}  // namespace

int ext_reserve(ExtWorkspace &ws, size_t num_nodes, size_t slots_per_node, cudaStream_t s) {
  const size_t slots = num_nodes * slots_per_node;
  if (ws.cap_nodes < num_nodes) {
    XG_CUDA(cudaStreamSynchronize(s));
    for (uint32_t **p : {&ws.keys_in, &ws.keys_out, &ws.vals_in, &ws.vals_out}) {
      if (*p) XG_CUDA(cudaFree(*p));
      *p = nullptr;
    }
    if (ws.sort_temp) XG_CUDA(cudaFree(ws.sort_temp));
    ws.sort_temp = nullptr;
    const size_t cap = num_nodes + num_nodes / 4 + 1024;
    for (uint32_t **p : {&ws.keys_in, &ws.keys_out, &ws.vals_in, &ws.vals_out}) XG_CUDA(cudaMalloc(p, cap * sizeof(uint32_t)));
    const size_t bytes = (((cap + kRadixTile - 1) / kRadixTile) + 1) * kRadixBins * sizeof(uint32_t);  // hist[bin][tile] + bin totals
    XG_CUDA(cudaMalloc(&ws.sort_temp, bytes ? bytes : 16));
    ws.sort_bytes = bytes;
    ws.cap_nodes = cap;
  }
  if (ws.cap_slots < slots) {
    XG_CUDA(cudaStreamSynchronize(s));
    if (ws.tmp_src) XG_CUDA(cudaFree(ws.tmp_src));
    if (ws.tmp_dst) XG_CUDA(cudaFree(ws.tmp_dst));
    ws.tmp_src = ws.tmp_dst = nullptr;
    const size_t cap = slots + slots / 4 + 1024;
    XG_CUDA(cudaMalloc(&ws.tmp_src, cap * sizeof(uint32_t)));
    XG_CUDA(cudaMalloc(&ws.tmp_dst, cap * sizeof(uint32_t)));
    ws.cap_slots = cap;
  }
  const size_t tiles = (slots + kCompactTile - 1) / kCompactTile + 2;
  if (ws.scan_tiles < tiles) {
    XG_CUDA(cudaStreamSynchronize(s));
    if (ws.scan_mem) XG_CUDA(cudaFree(ws.scan_mem));
    ws.scan_mem = nullptr;
    const size_t want = tiles + tiles / 4 + 64;
    XG_CUDA(cudaMalloc(&ws.scan_mem, (want + 1) * sizeof(uint64_t)));
    ws.scan_tiles = want;
  }
  return XG_OK;
}

void ext_release(ExtWorkspace &ws) {
  cudaFree(ws.keys_in); cudaFree(ws.keys_out); cudaFree(ws.vals_in); cudaFree(ws.vals_out);
  cudaFree(ws.tmp_src); cudaFree(ws.tmp_dst); cudaFree(ws.sort_temp); cudaFree(ws.scan_mem);
  ws = ExtWorkspace{};
}

int launch_sample_ext(int sample_type, const xg_graph &g, const WeightTables &w, const uint32_t *d_input,
                      size_t num_input_max, const uint32_t *d_num_input, size_t fanout, bool emit_local,
                      const uint32_t *src_local, uint32_t *d_out_src, uint32_t *d_out_dst, uint64_t *d_num_out,
                      uint2 rng_key, const uint2 *d_rng_key, uint32_t salt, ExtWorkspace &ws, cudaStream_t s) {
  XG_REQUIRE(fanout >= 1 && fanout <= 4096, "fanout must be in [1, 4096]");
  XG_REQUIRE(num_input_max * fanout < (1ull << 31), "num_input * fanout must stay below 2^31");
  const bool weighted = sample_type != XG_KHOP1;
  if (weighted) {
    XG_REQUIRE(g.num_part == 0, "weighted samplers take the flat CSR their tables are aligned with");
    if (sample_type == XG_WEIGHTED_KHOP_PREFIX) XG_REQUIRE(w.prefix != nullptr, "prob_prefix_table is NULL");
    else XG_REQUIRE(w.prob != nullptr && w.alias != nullptr, "prob_table / alias_table is NULL");
  }
  if (sample_type == XG_WEIGHTED_KHOP_HASH_DEDUP)
    XG_REQUIRE(fanout <= (size_t)kHashDedupMaxFanout, "hash-dedup sampler supports fanout <= 64");
  if (num_input_max == 0) {
    XG_CUDA(cudaMemsetAsync(d_num_out, 0, sizeof(uint64_t), s));
    return XG_OK;
  }
  XG_REQUIRE(ws.cap_nodes >= num_input_max && ws.cap_slots >= num_input_max * fanout, "workspace not reserved");
  const size_t slots = num_input_max * fanout;
  const size_t tiles = (slots + kCompactTile - 1) / kCompactTile;
  XG_REQUIRE(tiles + 1 <= ws.scan_tiles, "scan workspace too small");
  XG_CUDA(cudaMemsetAsync(ws.scan_mem, 0, (tiles + 2) * sizeof(uint64_t), s));
  auto *ticket = reinterpret_cast<uint32_t *>(ws.scan_mem);
  auto *state = reinterpret_cast<uint64_t *>(ws.scan_mem) + 1;
  const unsigned grid_slots = grid_for((slots + 255) / 256, 8);

  if (sample_type == XG_WEIGHTED_KHOP_HASH_DEDUP) {
    draw_hash_dedup_kernel<<<grid_for((num_input_max + 127) / 128, 16), 128, 0, s>>>(
        g, w, d_input, num_input_max, d_num_input, (uint32_t)fanout, ws.tmp_src, ws.tmp_dst, rng_key, d_rng_key, salt);
    XG_LAUNCHED();
    compact_pairs_kernel<false><<<grid_for(tiles, 8), kCompactBlock, 0, s>>>(
        ws.tmp_src, ws.tmp_dst, nullptr, (uint32_t)fanout, num_input_max, d_num_input, emit_local, src_local,
        d_out_src, d_out_dst, d_num_out, ticket, state);
    XG_LAUNCHED();
    return XG_OK;
  }

  // sort the N input ids (stable, by id); padding keys park at the end
  sort_prep_kernel<<<grid_for((num_input_max + 255) / 256, 8), 256, 0, s>>>(d_input, num_input_max, d_num_input,
                                                                          ws.keys_in, ws.vals_in);
  XG_LAUNCHED();
  // keys are node ids: bits(num_node - 1) significant bits (32 when the handle does not say how many nodes there are)
  int key_bits = 32;
  if (g.num_node > 1) key_bits = 32 - __builtin_clz((unsigned)(g.num_node - 1));
  bool in_b = false;
  if (int rc = radix_sort_pairs(ws.keys_in, ws.vals_in, ws.keys_out, ws.vals_out, num_input_max, d_num_input, key_bits,
                                static_cast<uint32_t *>(ws.sort_temp), &in_b, s))
    return rc;
  const uint32_t *sorted_ids = in_b ? ws.keys_out : ws.keys_in, *sorted_pos = in_b ? ws.vals_out : ws.vals_in;
  switch (sample_type) {
    case XG_KHOP1:
      draw_with_replacement_kernel<XG_KHOP1><<<grid_slots, 256, 0, s>>>(
          g, w, sorted_ids, sorted_pos, num_input_max, d_num_input, (uint32_t)fanout, ws.tmp_src, ws.tmp_dst, rng_key,
          d_rng_key, salt);
      break;
    case XG_WEIGHTED_KHOP:
      draw_with_replacement_kernel<XG_WEIGHTED_KHOP><<<grid_slots, 256, 0, s>>>(
          g, w, sorted_ids, sorted_pos, num_input_max, d_num_input, (uint32_t)fanout, ws.tmp_src, ws.tmp_dst, rng_key,
          d_rng_key, salt);
      break;
    case XG_WEIGHTED_KHOP_PREFIX:
      draw_with_replacement_kernel<XG_WEIGHTED_KHOP_PREFIX><<<grid_slots, 256, 0, s>>>(
          g, w, sorted_ids, sorted_pos, num_input_max, d_num_input, (uint32_t)fanout, ws.tmp_src, ws.tmp_dst, rng_key,
          d_rng_key, salt);
      break;
    default:
      XG_REQUIRE(false, "not a with-replacement sampler type");
  }
  XG_LAUNCHED();
  compact_pairs_kernel<true><<<grid_for(tiles, 8), kCompactBlock, 0, s>>>(
      ws.tmp_src, ws.tmp_dst, sorted_pos, (uint32_t)fanout, num_input_max, d_num_input, emit_local, src_local,
      d_out_src, d_out_dst, d_num_out, ticket, state);
  XG_LAUNCHED();
  return XG_OK;
}

}  // namespace xg

// =================================================================================================
// C-ABI
// =================================================================================================
using namespace xg;

namespace {
ExtWorkspace g_ext_ws[XG_MAX_PARTS * 4];
}

extern "C" {

int xg_sample_khop1(const xg_graph *g, const uint32_t *d_input, size_t num_input, size_t fanout,
                    uint32_t *d_out_src, uint32_t *d_out_dst, size_t *d_num_out, uint64_t seed, uint64_t batch_key,
                    void *stream) {
  XG_REQUIRE(g && d_num_out, "NULL argument");
  int dev = 0;
  XG_CUDA(cudaGetDevice(&dev));
  ExtWorkspace &ws = g_ext_ws[dev % (XG_MAX_PARTS * 4)];
  cudaStream_t s = as_stream(stream);
  int rc = ext_reserve(ws, num_input, fanout, s);
  if (rc) return rc;
  return launch_sample_ext(XG_KHOP1, *g, WeightTables{}, d_input, num_input, nullptr, fanout, false, nullptr, d_out_src,
                           d_out_dst, reinterpret_cast<uint64_t *>(d_num_out), philox_key(seed, batch_key), nullptr,
                           kSampleSalt, ws, s);
}

int xg_sample_weighted_khop(const xg_graph *g, int sample_type, const float *d_prob_table,
                            const uint32_t *d_alias_table, const float *d_prob_prefix_table, const uint32_t *d_input,
                            size_t num_input, size_t fanout, uint32_t *d_out_src, uint32_t *d_out_dst,
                            size_t *d_num_out, uint64_t seed, uint64_t batch_key, void *stream) {
  XG_REQUIRE(g && d_num_out, "NULL argument");
  XG_REQUIRE(sample_type == XG_WEIGHTED_KHOP || sample_type == XG_WEIGHTED_KHOP_PREFIX ||
                 sample_type == XG_WEIGHTED_KHOP_HASH_DEDUP,
             "sample_type must be a weighted KHop variant");
  int dev = 0;
  XG_CUDA(cudaGetDevice(&dev));
  ExtWorkspace &ws = g_ext_ws[dev % (XG_MAX_PARTS * 4)];
  cudaStream_t s = as_stream(stream);
  int rc = ext_reserve(ws, num_input, fanout, s);
  if (rc) return rc;
  WeightTables w;
  w.prob = d_prob_table;
  w.alias = d_alias_table;
  w.prefix = d_prob_prefix_table;
  return launch_sample_ext(sample_type, *g, w, d_input, num_input, nullptr, fanout, false, nullptr, d_out_src,
                           d_out_dst, reinterpret_cast<uint64_t *>(d_num_out), philox_key(seed, batch_key), nullptr,
                           kSampleSalt, ws, s);
}

/* the 4 x 32 random bits a sampler kernel sees for counter (c0, c1, c2, c3): lets a test replay a kernel's
 * draws through the CPU oracle (host-only helper; same code path as the device, common.cuh Philox) */
void xg_philox4x32(uint64_t seed, uint64_t batch_key, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                   uint32_t out[4]) {
  const uint4 r = Philox::rand4(philox_key(seed, batch_key), make_uint4(c0, c1, c2, c3));
  out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
uint32_t xg_sample_salt(void) { return kSampleSalt; }

}  // extern "C"
