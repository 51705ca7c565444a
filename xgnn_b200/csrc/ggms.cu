// GGMS store plumbing on the host side: cache-rank shuffle, the
// node-prefix rule for the topology cache, partition builders and CUDA-IPC handle exchange.
//
// Reference behaviour being replaced:
//   GPUCacheManager ctor             cuda/cuda_cache_manager_host.cc:133-254  (rank shuffle + table)
//   DistEngine::SampleDataCopy       dist/dist_engine.cc:198-244              (num_cache_node rule)
//   DistGraph::_DatasetPartition     cuda/dist_graph.cu:228-272               (single-threaded CPU loops)
//   DistGraph::_DataIpcShare         cuda/dist_graph.cu:274-307               (IPC handles + barrier)
// The partition builders here run on the GPU (one pass over the CSR) instead of on one CPU thread.
#include <algorithm>
#include <cstring>
#include <random>

#include "common.cuh"

namespace xg {
namespace {

// part CSR offsets: out_indptr[j + 1] - out_indptr[j] = deg(part + j * P), scanned by the caller's
// second pass; here: degrees only
__global__ void __launch_bounds__(256)
part_degree_kernel(const uint32_t *__restrict__ indptr, uint32_t part, uint32_t num_part, size_t n_rows,
                   uint32_t *__restrict__ deg_out) {
  for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n_rows; j += (size_t)gridDim.x * blockDim.x) {
    const size_t v = (size_t)part + j * num_part;
    deg_out[j] = indptr[v + 1] - indptr[v];
  }
}

// copy the neighbour lists of nodes part, part + P, ... into the rebased CSR (8 lanes per node)
__global__ void __launch_bounds__(256)
part_copy_kernel(const uint32_t *__restrict__ indptr, const uint32_t *__restrict__ indices, uint32_t part,
                 uint32_t num_part, size_t n_rows, const uint32_t *__restrict__ part_indptr,
                 uint32_t *__restrict__ part_indices) {
  constexpr int G = 8;
  const size_t groups = ((size_t)gridDim.x * blockDim.x) / G;
  const unsigned gl = threadIdx.x % G;
  for (size_t j = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) / G; j < n_rows; j += groups) {
    const size_t v = (size_t)part + j * num_part;
    const uint32_t b = indptr[v], e = indptr[v + 1], ob = part_indptr[j];
    for (uint32_t k = gl; k < e - b; k += G) part_indices[ob + k] = indices[b + k];
  }
}

}  // namespace
}  // namespace xg

using namespace xg;

extern "C" {

/* std::shuffle(nodes, nodes + n, std::mt19937(n)) -- cuda_cache_manager_host.cc:169-171. Every process
 * computes the same order, so rank r -> part r % P needs no communication. */
int xg_shuffle_cache_ranks(uint32_t *h_nodes, size_t num_cached) {
  XG_REQUIRE(h_nodes != nullptr || num_cached == 0, "h_nodes is NULL");
  std::mt19937 eg(static_cast<uint32_t>(num_cached));
  std::shuffle(h_nodes, h_nodes + num_cached, eg);
  return XG_OK;
}

/* num_cache_node of the topology: smallest n with indptr[n] >= floor(E * ratio)
 * (dist/dist_engine.cc:223-232); a node-id prefix. */
uint32_t xg_num_cache_node(const uint32_t *h_indptr, uint32_t num_node, double ratio) {
  if (!h_indptr) return 0;
  const uint64_t num_edge = h_indptr[num_node];
  const uint64_t want = static_cast<uint64_t>(num_edge * ratio);
  const uint32_t *p = std::lower_bound(h_indptr, h_indptr + num_node + 1, want,
                                       [](uint32_t a, uint64_t b) { return (uint64_t)a < b; });
  size_t n = static_cast<size_t>(p - h_indptr);
  return static_cast<uint32_t>(n > num_node ? num_node : n);
}

/* rows of partition `part`: ceil((num_cache_node - part) / P) (dist_graph.cu:242-262) */
size_t xg_part_num_rows(uint32_t num_cache_node, uint32_t part, uint32_t num_part) {
  if (num_part == 0 || part >= num_cache_node) return 0;
  return ((size_t)num_cache_node - part + num_part - 1) / num_part;
}

/* _DatasetPartition step 1: degrees of the part's rows (d_indptr may be device or pinned host memory).
 * The caller scans them (exclusive) into the part's indptr. */
int xg_part_degrees(const uint32_t *d_indptr, uint32_t num_cache_node, uint32_t part, uint32_t num_part,
                    uint32_t *d_deg_out, void *stream) {
  const size_t n = xg_part_num_rows(num_cache_node, part, num_part);
  if (n == 0) return XG_OK;
  XG_REQUIRE(d_indptr && d_deg_out, "NULL argument");
  part_degree_kernel<<<grid_for((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(d_indptr, part, num_part, n,
                                                                                  d_deg_out);
  XG_LAUNCHED();
  return XG_OK;
}

/* _DatasetPartition step 2: neighbour lists into the rebased CSR */
int xg_part_copy_indices(const uint32_t *d_indptr, const uint32_t *d_indices, uint32_t num_cache_node, uint32_t part,
                         uint32_t num_part, const uint32_t *d_part_indptr, uint32_t *d_part_indices, void *stream) {
  const size_t n = xg_part_num_rows(num_cache_node, part, num_part);
  if (n == 0) return XG_OK;
  XG_REQUIRE(d_indptr && d_indices && d_part_indptr && d_part_indices, "NULL argument");
  part_copy_kernel<<<grid_for((n * 8 + 255) / 256, 16), 256, 0, as_stream(stream)>>>(
      d_indptr, d_indices, part, num_part, n, d_part_indptr, d_part_indices);
  XG_LAUNCHED();
  return XG_OK;
}

/* ---- raw device memory + CUDA IPC (DistGraph::_DataIpcShare, dist_graph.cu:274-307) ----------------
 * Shards that peers map must come from their own cudaMalloc (an IPC handle names a whole allocation). */
int xg_device_malloc(void **out, size_t bytes) {
  XG_REQUIRE(out != nullptr, "out is NULL");
  // Sizes are rounded up to whole 2 MiB pages.  Measured on B200 / driver 580 (tools/p2p_probe4.py,
  // profiles/r01c_p2p_probe.md): a multi-GB cudaMalloc whose size is NOT a multiple of 2 MiB is served to an
  // IPC-importing peer at 50-90 GB/s (random 512-B rows over NVLink) instead of 750 GB/s; local access and
  // rounded sizes are unaffected.
  constexpr size_t kPage = size_t(2) << 20;
  const size_t want = bytes ? (bytes + kPage - 1) / kPage * kPage : kPage;
  XG_CUDA(cudaMalloc(out, want));
  return XG_OK;
}
int xg_device_free(void *p) {
  if (p) XG_CUDA(cudaFree(p));
  return XG_OK;
}
int xg_ipc_export(const void *d_ptr, unsigned char out_handle[64]) {
  XG_REQUIRE(d_ptr && out_handle, "NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  cudaIpcMemHandle_t h;
  XG_CUDA(cudaIpcGetMemHandle(&h, const_cast<void *>(d_ptr)));
  std::memcpy(out_handle, &h, 64);
  return XG_OK;
}
int xg_ipc_import(const unsigned char handle[64], void **out) {
  XG_REQUIRE(handle && out, "NULL argument");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, 64);
  XG_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return XG_OK;
}
int xg_ipc_close(void *p) {
  if (p) XG_CUDA(cudaIpcCloseMemHandle(p));
  return XG_OK;
}
/* single-process multi-GPU (tests): let `device` load from `peer` */
int xg_enable_peer_access(int device, int peer) {
  int can = 0;
  XG_CUDA(cudaDeviceCanAccessPeer(&can, device, peer));
  XG_REQUIRE(can, "devices are not P2P capable");
  int prev = 0;
  XG_CUDA(cudaGetDevice(&prev));
  XG_CUDA(cudaSetDevice(device));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    e = cudaSuccess;
  }
  cudaSetDevice(prev);
  XG_CUDA(e);
  return XG_OK;
}

}  // extern "C"
