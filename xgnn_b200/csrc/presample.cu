// Presampling cache-rank policy: rank nodes by how often they appear in a minibatch's
// input_nodes over `presample_epoch` epochs of real sampling.
//
// Reference behaviour being replaced: dist::PreSampler (samgraph/common/dist/pre_sampler.cc:40-165; CUDA twin
// cuda/pre_sampler.cc).  There every batch's input_nodes is copied D2H, counted by an OpenMP loop into a
// host array of 64-bit {low: node id, high: frequency} words, and the array is sorted with
// __gnu_parallel::sort(std::greater<uint64_t>) - frequency descending, ties by node id descending.
//
// Here the counts never leave the GPU: xg_presample_accumulate adds one batch's unique node list into
// u32 freq[V] (the list is duplicate-free, so plain atomics never contend), and xg_presample_rank packs the
// same 64-bit words and sorts them with one descending radix sort (cub, init-time only), giving the
// identical order.
#include <cub/cub.cuh>

#include "common.cuh"

namespace xg {
namespace {

__global__ void __launch_bounds__(256)
freq_accumulate_kernel(uint32_t *__restrict__ freq, const uint32_t *__restrict__ nodes, size_t n_max,
                       const uint32_t *__restrict__ d_n, uint32_t num_node) {
  size_t n = d_n ? (*d_n < n_max ? *d_n : n_max) : n_max;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t v = nodes[i];
    if (v < num_node) atomicAdd(freq + v, 1u);
  }
}

__global__ void __launch_bounds__(256)
freq_pack_kernel(const uint32_t *__restrict__ freq, uint64_t *__restrict__ keys, size_t num_node) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < num_node; i += (size_t)gridDim.x * blockDim.x)
    keys[i] = ((uint64_t)freq[i] << 32) | (uint64_t)i;  // pre_sampler.cc:48-52: low word id, high word count
}

__global__ void __launch_bounds__(256)
freq_unpack_kernel(const uint64_t *__restrict__ keys, uint32_t *__restrict__ rank_nodes,
                   uint32_t *__restrict__ rank_freq, size_t num_node) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < num_node; i += (size_t)gridDim.x * blockDim.x) {
    const uint64_t k = keys[i];
    rank_nodes[i] = (uint32_t)k;                       // GetRankNode, pre_sampler.cc:141-152
    if (rank_freq) rank_freq[i] = (uint32_t)(k >> 32);  // GetFreq, pre_sampler.cc:131-139
  }
}

}  // namespace
}  // namespace xg

using namespace xg;

extern "C" {

int xg_presample_accumulate(uint32_t *d_freq, uint32_t num_node, const uint32_t *d_nodes, size_t num_nodes_max,
                            const uint32_t *d_num_nodes, void *stream) {
  XG_REQUIRE(d_freq != nullptr, "d_freq is NULL");
  if (num_nodes_max == 0) return XG_OK;
  XG_REQUIRE(d_nodes != nullptr, "d_nodes is NULL");
  freq_accumulate_kernel<<<grid_for((num_nodes_max + 255) / 256, 8), 256, 0, as_stream(stream)>>>(
      d_freq, d_nodes, num_nodes_max, d_num_nodes, num_node);
  XG_LAUNCHED();
  return XG_OK;
}

int xg_presample_rank(const uint32_t *d_freq, uint32_t num_node, uint32_t *d_rank_nodes, uint32_t *d_rank_freq,
                      void *stream) {
  XG_REQUIRE(d_freq && d_rank_nodes, "NULL argument");
  if (num_node == 0) return XG_OK;
  cudaStream_t s = as_stream(stream);
  uint64_t *keys = nullptr, *sorted = nullptr;
  void *temp = nullptr;
  size_t temp_bytes = 0;
  XG_CUDA(cudaMalloc(&keys, (size_t)num_node * sizeof(uint64_t)));
  cudaError_t e = cudaMalloc(&sorted, (size_t)num_node * sizeof(uint64_t));
  if (e == cudaSuccess)
    e = cub::DeviceRadixSort::SortKeysDescending(nullptr, temp_bytes, keys, sorted, (int64_t)num_node, 0, 64, s);
  if (e == cudaSuccess) e = cudaMalloc(&temp, temp_bytes ? temp_bytes : 16);
  if (e == cudaSuccess) {
    const unsigned grid = grid_for(((size_t)num_node + 255) / 256, 8);
    // every exit goes through the cleanup below (no early return with temp / sorted / keys allocated)
    freq_pack_kernel<<<grid, 256, 0, s>>>(d_freq, keys, num_node);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cub::DeviceRadixSort::SortKeysDescending(temp, temp_bytes, keys, sorted, (int64_t)num_node, 0, 64, s);
    if (e == cudaSuccess) {
      freq_unpack_kernel<<<grid, 256, 0, s>>>(sorted, d_rank_nodes, d_rank_freq, num_node);
      g_launches.fetch_add(1, std::memory_order_relaxed);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  }
  cudaFree(temp);
  cudaFree(sorted);
  cudaFree(keys);
  XG_CUDA(e);
  return XG_OK;
}

}  // extern "C"

// =================================================================================================
// SAMGRAPH_SANITY_CHECK (reference: GPUBatchSanityCheck / GPUSanityCheckList, cuda/cuda_sanity_check.cu:28-88,
// called from GPUShuffler::GetBatch, cuda/cuda_shuffler.cc:145-152; CPUSanityCheckNoDuplicate,
// cpu/cpu_sanity_check.cc:29-45).  The reference asserts inside the kernel; here violations are COUNTED
// (one launch per list, no printf from the device) and the caller decides what to do with the count.
//   seeds:  no kEmptyKey, every id < num_node, and - with an epoch map - not seen before in this epoch
//   edges:  row[e] < num_src, col[e] < num_dst
namespace xg {
namespace {

__global__ void __launch_bounds__(256)
sanity_ids_kernel(const uint32_t *__restrict__ ids, size_t n, uint32_t num_node, uint32_t *epoch_map,
                  unsigned long long *violations) {
  unsigned long long bad = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t v = ids[i];
    if (v == kEmpty || v >= num_node) {
      ++bad;
    } else if (epoch_map != nullptr && atomicExch(epoch_map + v, 1u) != 0u) {
      ++bad;  // "duplicate batch input", cuda_sanity_check.cu:37-41
    }
  }
  if (bad) atomicAdd(violations, bad);
}

__global__ void __launch_bounds__(256)
sanity_edges_kernel(const uint32_t *__restrict__ row, const uint32_t *__restrict__ col, size_t n, uint32_t num_src,
                    uint32_t num_dst, unsigned long long *violations) {
  unsigned long long bad = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    if (row[i] >= num_src || col[i] >= num_dst) ++bad;
  if (bad) atomicAdd(violations, bad);
}

}  // namespace
}  // namespace xg

extern "C" {

int xg_sanity_check_ids(const uint32_t *d_ids, size_t n, uint32_t num_node, uint32_t *d_epoch_map,
                        uint64_t *d_violations, void *stream) {
  XG_REQUIRE(d_violations != nullptr, "d_violations is NULL");
  if (n == 0) return XG_OK;
  XG_REQUIRE(d_ids != nullptr, "d_ids is NULL");
  sanity_ids_kernel<<<grid_for((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(
      d_ids, n, num_node, d_epoch_map, reinterpret_cast<unsigned long long *>(d_violations));
  XG_LAUNCHED();
  return XG_OK;
}

int xg_sanity_check_edges(const uint32_t *d_row, const uint32_t *d_col, size_t n, uint32_t num_src, uint32_t num_dst,
                          uint64_t *d_violations, void *stream) {
  XG_REQUIRE(d_violations != nullptr, "d_violations is NULL");
  if (n == 0) return XG_OK;
  XG_REQUIRE(d_row && d_col, "NULL edge list");
  sanity_edges_kernel<<<grid_for((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(
      d_row, d_col, n, num_src, num_dst, reinterpret_cast<unsigned long long *>(d_violations));
  XG_LAUNCHED();
  return XG_OK;
}

}  // extern "C"
