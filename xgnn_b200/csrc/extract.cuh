// Row gather over the GGMS feature store; kernels in extract.cu.
#pragma once
#include "common.cuh"

namespace xg {

enum GatherMode : int {
  kByNode = 0,   // idx = node id; cache_table decides cache vs host
  kHostRow = 1,  // idx = row of store.host (miss list / plain GPUExtract)
  kCacheRank = 2 // idx = cache rank (hit list)
};

// out[dst_idx ? dst_idx[i] : i] = row(src_idx[i]) for i < min(num_rows, *d_num_rows).
// d_num_rows (device, may be NULL) lets a whole minibatch be enqueued before its counts are known.
int launch_gather(const xg_feature_store &st, int mode, const uint32_t *src_idx, const uint32_t *dst_idx,
                  size_t num_rows, const uint32_t *d_num_rows, void *out, uint64_t *stats, cudaStream_t s);

}  // namespace xg
