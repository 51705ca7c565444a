// Probe (one process, two GPUs): how many NVLink REQUEST bytes does a peer-row gather cost, and does a 512-byte bulk copy
// (TMA engine, cp.async.bulk) ask for a row with fewer / larger requests than 32 lanes x 16-byte loads do?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/nvlink_req_probe tools/nvlink_req_probe.cu
//   /tmp/nvlink_req_probe                  # GB/s one way and both ways at once, per variant
//   ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,gpu__time_duration.sum /tmp/nvlink_req_probe --one-way
// Random 512-byte rows out of a 4 GiB table in the OTHER GPU's HBM -> a dense local array.  Not product code.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x)                                                                                  \
  do {                                                                                         \
    cudaError_t e_ = (x);                                                                      \
    if (e_ != cudaSuccess) {                                                                   \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));      \
      exit(1);                                                                                 \
    }                                                                                          \
  } while (0)

constexpr int kRow = 512;
constexpr int kWarps = 8;

// ---- variant A: every lane loads 16 bytes (a warp = one row = 4 x 128-byte sectors), U rows in flight per warp ----
template <int U>
__global__ void __launch_bounds__(kWarps * 32) gather_ldg(const char *__restrict__ table, const uint32_t *__restrict__ idx,
                                                         size_t n, char *__restrict__ out) {
  const unsigned lane = threadIdx.x & 31;
  const size_t warp = (size_t)blockIdx.x * kWarps + (threadIdx.x >> 5), nw = (size_t)gridDim.x * kWarps;
  for (size_t r0 = warp * U; r0 < n; r0 += nw * U) {
    int4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (r0 + u < n) {
        const int4 *p = reinterpret_cast<const int4 *>(table + (size_t)idx[r0 + u] * kRow) + lane;
        asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p));
      }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (r0 + u < n) reinterpret_cast<int4 *>(out + (r0 + u) * kRow)[lane] = v[u];
  }
}

// ---- variant B: one lane asks the TMA engine for the whole row (cp.async.bulk 512 B -> shared), S rows in flight per warp ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int S>
__global__ void __launch_bounds__(kWarps * 32) gather_bulk(const char *__restrict__ table, const uint32_t *__restrict__ idx,
                                                          size_t n, char *__restrict__ out) {
  extern __shared__ __align__(128) char smem[];
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  char *ring = smem + (size_t)wib * S * kRow;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)kWarps * S * kRow) + wib * S;
  const size_t warp = (size_t)blockIdx.x * kWarps + wib, nw = (size_t)gridDim.x * kWarps;
  if (lane == 0)
    for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + s)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const size_t my_rows = warp < n ? (n - warp + nw - 1) / nw : 0;
  auto issue = [&](size_t k) {  // lane 0 only
    const int s = (int)(k % S);
    const char *src = table + (size_t)idx[warp + k * nw] * kRow;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bars + s)), "r"(kRow) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(ring + s * kRow)),
                 "l"(src), "r"(kRow), "r"(smem_u32(bars + s))
                 : "memory");
  };
  if (lane == 0)
    for (size_t k = 0; k < (size_t)S && k < my_rows; ++k) issue(k);
  for (size_t k = 0; k < my_rows; ++k) {
    const int s = (int)(k % S);
    const uint32_t parity = (uint32_t)((k / S) & 1);
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(smem_u32(bars + s)), "r"(parity) : "memory");
    const int4 v = reinterpret_cast<const int4 *>(ring + s * kRow)[lane];
    reinterpret_cast<int4 *>(out + (warp + k * nw) * kRow)[lane] = v;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // my generic read of the slot before the engine rewrites it
    __syncwarp();
    if (lane == 0 && k + S < my_rows) issue(k + S);
  }
}

// ---- variant C: the rows never touch a register.  One warp = one CTA; a stage = 32 rows (16 KB): every lane asks the TMA
// engine for its row (cp.async.bulk global -> shared, one mbarrier per stage), and when the stage has landed every lane
// hands its row back to the engine (cp.async.bulk shared -> global).  ST stages per warp.
template <int ST>
__global__ void __launch_bounds__(32) gather_bulk2(const char *__restrict__ table, const uint32_t *__restrict__ idx, size_t n,
                                                   char *__restrict__ out) {
  extern __shared__ __align__(128) char smem[];
  const unsigned lane = threadIdx.x;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)ST * 32 * kRow);
  const size_t warp = blockIdx.x, nw = gridDim.x;
  if (lane == 0)
    for (int s = 0; s < ST; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + s)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const size_t groups = (n + 31) / 32;
  const size_t mine = warp < groups ? (groups - warp + nw - 1) / nw : 0;
  auto issue = [&](size_t k) {
    const int s = (int)(k % ST);
    const size_t g = warp + k * nw;
    const uint32_t rows_in = (uint32_t)((n - g * 32) < 32 ? (n - g * 32) : 32);
    if (lane == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bars + s)), "r"(rows_in * kRow) : "memory");
    __syncwarp();
    if (lane < rows_in) {
      const char *src = table + (size_t)idx[g * 32 + lane] * kRow;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_u32(smem + ((size_t)s * 32 + lane) * kRow)),
                   "l"(src), "r"(kRow), "r"(smem_u32(bars + s))
                   : "memory");
    }
  };
  for (size_t k = 0; k < (size_t)ST && k < mine; ++k) issue(k);
  for (size_t k = 0; k < mine; ++k) {
    const int s = (int)(k % ST);
    const uint32_t parity = (uint32_t)((k / ST) & 1);
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(smem_u32(bars + s)), "r"(parity) : "memory");
    const size_t g = warp + k * nw;
    const size_t row = g * 32 + lane;
    if (row < n)
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + row * kRow),
                   "r"(smem_u32(smem + ((size_t)s * 32 + lane) * kRow)), "r"(kRow)
                   : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // the stage handed back one step ago has been read
    __syncwarp();
    if (k >= 1 && k - 1 + ST < mine) issue(k - 1 + ST);
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

struct Side {
  int dev;
  char *table = nullptr, *out = nullptr;
  uint32_t *idx = nullptr;
  cudaStream_t st;
  cudaEvent_t e0, e1;
};

int main(int argc, char **argv) {
  bool one_way = false, local = false;
  for (int i = 1; i < argc; ++i) one_way |= !strcmp(argv[i], "--one-way");
  for (int i = 1; i < argc; ++i) local |= !strcmp(argv[i], "--local");  // rows from the GPU's OWN table: the HBM side
  if (local) one_way = true;
  const size_t table_rows = (size_t)8 << 20, n = (size_t)2 << 20;  // 4 GiB table, 1 GiB gathered
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) { printf("needs two GPUs\n"); return 0; }
  Side sd[2];
  std::vector<uint32_t> h(n);
  uint64_t x = 88172645463325252ull;
  for (auto &v : h) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; v = (uint32_t)(x % table_rows); }
  for (int d = 0; d < 2; ++d) {
    sd[d].dev = d;
    CK(cudaSetDevice(d));
    CK(cudaDeviceEnablePeerAccess(1 - d, 0));
    CK(cudaMalloc(&sd[d].table, table_rows * kRow));
    CK(cudaMemset(sd[d].table, d + 1, table_rows * kRow));
    CK(cudaMalloc(&sd[d].out, n * kRow));
    CK(cudaMalloc(&sd[d].idx, n * 4));
    CK(cudaMemcpy(sd[d].idx, h.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaStreamCreate(&sd[d].st));
    CK(cudaEventCreate(&sd[d].e0));
    CK(cudaEventCreate(&sd[d].e1));
  }
  constexpr int S = 16;
  const size_t smem_bulk = (size_t)kWarps * S * kRow + (size_t)kWarps * S * 8;
  constexpr int ST = 4;
  const size_t smem_bulk2 = (size_t)ST * 32 * kRow + ST * 8;
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaFuncSetAttribute(gather_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bulk));
    CK(cudaFuncSetAttribute(gather_bulk2<ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bulk2));
  }
  auto launch = [&](int variant, int d) {
    CK(cudaSetDevice(d));
    const char *remote = local ? sd[d].table : sd[1 - d].table;
    if (variant == 0) gather_ldg<4><<<148 * 8, kWarps * 32, 0, sd[d].st>>>(remote, sd[d].idx, n, sd[d].out);
    else if (variant == 1) gather_ldg<8><<<148 * 4, kWarps * 32, 0, sd[d].st>>>(remote, sd[d].idx, n, sd[d].out);
    else if (variant == 2) gather_bulk<S><<<148 * 3, kWarps * 32, smem_bulk, sd[d].st>>>(remote, sd[d].idx, n, sd[d].out);
    else gather_bulk2<ST><<<148 * 3, 32, smem_bulk2, sd[d].st>>>(remote, sd[d].idx, n, sd[d].out);
    CK(cudaGetLastError());
  };
  const char *names[4] = {"ldg 16 B/lane, 4 rows in flight x 64 warps/SM", "ldg 16 B/lane, 8 rows in flight x 32 warps/SM",
                          "cp.async.bulk 512 B/row, 16 rows in flight x 24 warps/SM",
                          "bulk in + bulk out, 4 x 32-row stages x 3 warps/SM"};
  for (int variant = 0; variant < 4; ++variant) {
    for (int both = 0; both < (one_way ? 1 : 2); ++both) {
      float best[2] = {1e9f, 1e9f};
      for (int it = 0; it < (one_way ? 2 : 5); ++it) {
        for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
        for (int d = 0; d <= both; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(sd[d].e0, sd[d].st)); }
        for (int d = 0; d <= both; ++d) launch(variant, d);
        for (int d = 0; d <= both; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(sd[d].e1, sd[d].st)); }
        for (int d = 0; d <= both; ++d) {
          CK(cudaSetDevice(d));
          CK(cudaEventSynchronize(sd[d].e1));
          float ms = 0;
          CK(cudaEventElapsedTime(&ms, sd[d].e0, sd[d].e1));
          if (it > 0 && ms < best[d]) best[d] = ms;
        }
      }
      // check a few rows: every byte of a row pulled from device 1 - d is (2 - d)
      CK(cudaSetDevice(0));
      unsigned char probe[kRow];
      CK(cudaMemcpy(probe, sd[0].out + (n - 1) * kRow, kRow, cudaMemcpyDeviceToHost));
      bool ok = true;
      for (int i = 0; i < kRow; ++i) ok &= probe[i] == (local ? 1 : 2);
      const double gb = (double)n * kRow / 1e9;
      if (both) printf("%-58s both ways: %.1f / %.1f GB/s per direction  %s\n", names[variant], gb / best[0] * 1e3, gb / best[1] * 1e3, ok ? "" : "WRONG DATA");
      else printf("%-58s %s: %.1f GB/s of rows (x2 = HBM traffic when local)  %s\n", names[variant], local ? "local HBM" : "one way  ", gb / best[0] * 1e3, ok ? "" : "WRONG DATA");
      CK(cudaMemset(sd[0].out, 0, n * kRow));
    }
  }
  return 0;
}
