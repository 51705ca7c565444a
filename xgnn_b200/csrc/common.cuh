// Shared device/host helpers for the xgnn_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>
#include <string>

#include "../../include/xgnn_b200.h"

namespace xg {

constexpr uint32_t kEmpty = XG_EMPTY_KEY;
constexpr int kNumSM = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// ---------------------------------------------------------------------------------------------
// error plumbing
void set_error(const std::string &msg);
extern std::atomic<uint64_t> g_launches;

#define XG_CUDA(expr)                                                                      \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::xg::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " @" + __FILE__ + \
                      ":" + std::to_string(__LINE__));                                     \
      return XG_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)

#define XG_REQUIRE(cond, msg)                                  \
  do {                                                         \
    if (!(cond)) {                                             \
      ::xg::set_error(std::string("invalid argument: ") + msg); \
      return XG_ERR_INVALID;                                   \
    }                                                          \
  } while (0)

#define XG_LAUNCHED()                                     \
  do {                                                    \
    ::xg::g_launches.fetch_add(1, std::memory_order_relaxed); \
    XG_CUDA(cudaGetLastError());                          \
  } while (0)

inline cudaStream_t as_stream(void *s) { return static_cast<cudaStream_t>(s); }

// persistent-style grid: enough CTAs to fill the chip `waves_per_sm` deep, never more than tiles
inline unsigned grid_for(size_t tiles, int ctas_per_sm) {
  size_t cap = static_cast<size_t>(kNumSM) * ctas_per_sm;
  if (tiles < 1) tiles = 1;
  return static_cast<unsigned>(tiles < cap ? tiles : cap);
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11), counter-based: no state arrays, reproducible per
// (key, counter).  Replaces the per-thread curandState pool of cuda_random_states.cu:64-109.
struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  __host__ __device__ static inline uint4 rand4(uint2 key, uint4 ctr) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
      uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
      uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
#else
      uint64_t p0 = (uint64_t)M0 * ctr.x, p1 = (uint64_t)M1 * ctr.z;
      uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
      uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
      ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
      key.x += W0;
      key.y += W1;
    }
    return ctr;
  }
};

// uniform integer in [0, n) from 32 random bits (multiply-shift; bias < n / 2^32)
__host__ __device__ inline uint32_t bounded(uint32_t r, uint32_t n) {
#ifdef __CUDA_ARCH__
  return __umulhi(r, n);
#else
  return (uint32_t)(((uint64_t)r * n) >> 32);
#endif
}

// ---------------------------------------------------------------------------------------------
// Single-pass chained scan ("decoupled look-back", Merrill & Garland 2016) over per-tile
// aggregates.  One 64-bit word per tile: [63:62] status, [61:0] value.  Status words must be zero
// before the launch.  Tiles must be started in index order (dynamic ticket or blockIdx order with
// all earlier tiles resident or finished), which both launch styles used here guarantee.
constexpr uint64_t kScanAggregate = 1ull << 62;
constexpr uint64_t kScanInclusive = 2ull << 62;
constexpr uint64_t kScanValueMask = (1ull << 62) - 1;

__device__ inline uint64_t ld_relaxed(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline void st_relaxed(uint64_t *p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Called by ONE warp of the CTA (all 32 lanes). Publishes this tile's aggregate, walks back over
// predecessors, publishes the inclusive prefix and returns the exclusive prefix (sum of all
// earlier tiles' aggregates) to every lane.
__device__ inline uint64_t lookback_exclusive(uint64_t *state, uint32_t tile, uint64_t aggregate) {
  const unsigned lane = threadIdx.x & 31;
  if (tile == 0) {
    if (lane == 0) st_relaxed(state, kScanInclusive | aggregate);
    return 0;
  }
  if (lane == 0) st_relaxed(state + tile, kScanAggregate | aggregate);
  uint64_t exclusive = 0;
  int64_t base = (int64_t)tile - 1;
  while (true) {
    int64_t idx = base - lane;
    uint64_t w = 0;
    // spin until this lane's predecessor has published something
    if (idx >= 0) {
      do { w = ld_relaxed(state + idx); } while ((w >> 62) == 0);
    } else {
      w = kScanInclusive;  // virtual tile before 0: inclusive prefix 0
    }
    unsigned incl = __ballot_sync(0xffffffffu, (w >> 62) == 2);
    // lanes at or before the first inclusive predecessor contribute
    int first = incl ? (__ffs(incl) - 1) : 32;
    uint64_t v = (lane <= (unsigned)first) ? (w & kScanValueMask) : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    exclusive += v;
    if (incl) break;
    base -= 32;
  }
  if (lane == 0) st_relaxed(state + tile, kScanInclusive | (exclusive + aggregate));
  return exclusive;
}

// block-wide exclusive scan of one uint32 per thread, BLOCK threads (multiple of 32, <= 1024).
// Returns the exclusive prefix; *total (all threads) receives the block sum.
template <int BLOCK>
__device__ inline uint32_t block_exclusive_scan(uint32_t v, uint32_t *smem_warp /*[BLOCK/32]*/,
                                                uint32_t *total) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= (unsigned)o) incl += n;
  }
  if (lane == 31) smem_warp[warp] = incl;
  __syncthreads();
  uint32_t wsum = 0, all = 0;
#pragma unroll
  for (int w = 0; w < BLOCK / 32; ++w) {
    uint32_t s = smem_warp[w];
    if (w < (int)warp) wsum += s;
    all += s;
  }
  __syncthreads();
  *total = all;
  return wsum + incl - v;
}
#endif  // __CUDACC__

}  // namespace xg
