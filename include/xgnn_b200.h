/*
 * xgnn_b200 -- C-ABI of the B200-native sample -> unique/remap -> extract path.
 *
 * Drop-in boundary.  Two layers live in ONE shared library
 * (xgnn_b200/lib/libxgnn_b200.so):
 *
 *   1. include/samgraph_abi.h : the reference's own `samgraph_*` symbols
 *      (samgraph/common/operation.h:28-116) -- what samgraph/common/__init__.py binds with ctypes.
 *   2. this file              : operator-level entry points (`xg_*`), one per reference operator on
 *      the hot path, for callers that bind single operators (tests, bench, a maintainer replacing
 *      one .cu file at a time).  Each declaration cites the reference interface it replaces.
 *
 * Conventions
 *   - plain C: pointers + sizes; no torch / C++ types.  `d_` = device pointer (or UVA-mapped
 *     pinned host pointer where stated), `h_` = host pointer.  `stream` is a cudaStream_t passed as
 *     void* (NULL = legacy default stream).
 *   - IdType is uint32_t (constant.h:28); XG_EMPTY_KEY = 0xFFFFFFFF (constant.h:75).
 *   - every function returns 0 on success, a negative XG_ERR_* otherwise; xg_last_error() gives the
 *     message.  (The reference CHECK-aborts; the `samgraph_*` layer keeps that behaviour, this
 *     layer reports instead so that tests can assert on failures.)
 *   - no function synchronises the stream unless it has a host out-parameter, which is said so.
 */
#ifndef XGNN_B200_H
#define XGNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XG_EMPTY_KEY 0xFFFFFFFFu
#define XG_MAX_PARTS 16 /* >= kMaxDevice GPUs that can share a GGMS store; slot [num_part] = host */

enum {
  XG_OK = 0,
  XG_ERR_CUDA = -1,
  XG_ERR_INVALID = -2,
  XG_ERR_CAPACITY = -3,
  XG_ERR_UNSUPPORTED = -4,
};

/* DataType numbering of samgraph/common/common.h:38-47 */
enum { XG_F32 = 0, XG_F64 = 1, XG_F16 = 2, XG_U8 = 3, XG_I32 = 4, XG_I8 = 5, XG_I64 = 6 };

/* SampleType numbering of samgraph/common/common.h:50-59 */
enum {
  XG_KHOP0 = 0, XG_KHOP1 = 1, XG_WEIGHTED_KHOP = 2, XG_RANDOM_WALK = 3,
  XG_WEIGHTED_KHOP_PREFIX = 4, XG_KHOP2 = 5, XG_WEIGHTED_KHOP_HASH_DEDUP = 6, XG_KHOP3 = 7,
};

const char *xg_last_error(void);
int xg_version(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
uint64_t xg_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Topology handle: replaces DeviceNormalGraph / DeviceDistGraph (cuda/dist_graph.h:114-180).
 * Passed BY VALUE into kernels (no device-side pointer table to chase).
 *   flat CSR      : num_part = 0, part_indptr[0] / part_indices[0] = whole graph.
 *   GGMS sharded  : node v < num_cache_node -> part v % num_part, row v / num_part (rebased CSR);
 *                   otherwise slot [num_part] = the whole CSR in pinned host memory (UVA).
 * Pointers may be local HBM, peer HBM (P2P-mapped) or pinned host memory.
 */
typedef struct xg_graph {
  const uint32_t *part_indptr[XG_MAX_PARTS + 1];
  const uint32_t *part_indices[XG_MAX_PARTS + 1];
  uint32_t num_part;
  uint32_t num_cache_node;
  uint32_t num_node;
  uint32_t reserved;
} xg_graph;

/* Feature store handle: replaces `const void* src` of GPUExtract (cuda/cuda_extraction.cu:74),
 * the direct-address cache table of GPUCacheManager (cuda/cuda_cache_manager_host.cc:197-229) and
 * DeviceDistFeature (cuda/dist_graph.h:182-213) in one by-value struct.
 *   cache_table == NULL : every row comes from `host` (which may just as well be device memory).
 *   otherwise r = cache_table[node]; r == XG_EMPTY_KEY -> `host` row `node`;
 *             else part r % num_part, row r / num_part.
 */
typedef struct xg_feature_store {
  const void *part[XG_MAX_PARTS];
  const void *host;
  const uint32_t *cache_table;
  uint32_t num_part;
  uint32_t local_part_mask; /* bit p set: part p is in this GPU's HBM (statistics only) */
  uint64_t row_bytes;       /* dim * sizeof(dtype) */
} xg_feature_store;

/* ------------------------------------------------------------------------------------------------
 * Ordered hash table (unique + remap): replaces cuda::OrderedHashTable, cuda/cuda_hashtable.h:103-153.
 * Deterministic where the reference is racy: among duplicate occurrences the FIRST in input order
 * wins, so ids equal CPUHashTable0's (cpu/cpu_hashtable0.cc:37-47) bit for bit.
 */
typedef struct xg_hashtable xg_hashtable;

/* OrderedHashTable(size, ctx, stream, scale), cuda_hashtable.cu:699-727. max_items = most unique ids
 * one batch can hold (PredictNumNodes, common.cc:488-497). */
int xg_hashtable_create(xg_hashtable **out, size_t max_items, int device);
int xg_hashtable_destroy(xg_hashtable *t);
/* Reset, cuda_hashtable.cu:739-742 */
int xg_hashtable_reset(xg_hashtable *t, void *stream);
/* FillWithUnique, cuda_hashtable.cu:1045-1064: caller guarantees d_input has no duplicates */
int xg_hashtable_fill_with_unique(xg_hashtable *t, const uint32_t *d_input, size_t num_input, void *stream);
/* FillWithDupRevised, cuda_hashtable.cu:850-912 */
int xg_hashtable_fill_with_dup_revised(xg_hashtable *t, const uint32_t *d_input, size_t num_input, void *stream);
/* FillWithDuplicates, cuda_hashtable.cu:744-837: also copies the whole unique list (old prefix + new)
 * to d_unique and returns its length in *h_num_unique (host; synchronises the stream, as the
 * reference does). d_unique needs NumItems()+num_input entries. */
int xg_hashtable_fill_with_duplicates(xg_hashtable *t, const uint32_t *d_input, size_t num_input,
                                      uint32_t *d_unique, uint32_t *h_num_unique, void *stream);
/* FillWithDupMutable, cuda_hashtable.cu:914-974: like DupRevised, but d_input is overwritten with a
 * per-item scratch word (the slot position), as the reference allows itself to clobber it. */
int xg_hashtable_fill_with_dup_mutable(xg_hashtable *t, uint32_t *d_input, size_t num_input, void *stream);
/* RefUnique / CopyUnique / NumItems, cuda_hashtable.cu:839-848, cuda_hashtable.h:141.
 * NumItems synchronises the table's last fill (the count lives on the device). */
int xg_hashtable_ref_unique(xg_hashtable *t, const uint32_t **d_unique, uint32_t *h_num_unique);
int xg_hashtable_copy_unique(xg_hashtable *t, uint32_t *d_unique, void *stream);
size_t xg_hashtable_num_items(xg_hashtable *t);
/* DeviceOrderedHashTable::SearchO2N(id)->local for an array of ids, cuda_hashtable.h:57-60.
 * Absent ids give XG_EMPTY_KEY (the reference would spin forever). */
int xg_hashtable_lookup(xg_hashtable *t, const uint32_t *d_ids, size_t n, uint32_t *d_local, void *stream);
/* GPUMapEdges(global_src, new_global_src, global_dst, new_global_dst, num_edges, table, ctx, stream),
 * cuda/cuda_mapping.cu:68-80 (same argument order). */
int xg_map_edges(xg_hashtable *t, const uint32_t *d_src, uint32_t *d_new_src, const uint32_t *d_dst,
                 uint32_t *d_new_dst, size_t num_edges, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Samplers.  Common contract (cuda_sampling_khop0.cu:243-335): for each input node, in input order,
 * its sampled edges (out_src = the node, out_dst = neighbour), padding removed, node-major /
 * slot-minor; *d_num_out (device, size_t) receives the edge count.  Out arrays need
 * num_input * fanout entries.  RNG: Philox4x32-10 keyed by (seed, batch_key) with counter
 * (node position, lane) -- no state arrays (replaces GPURandomStates, cuda_random_states.cu:64-109).
 */
/* GPUSampleKHop0<GraphType>, cuda_sampling_khop0.cu:243-335 and GPUSampleKHop3<GraphType>,
 * cuda_sampling_khop3.cu:234-318 and GPUSampleKHop2, cuda_sampling_khop2.cu:297-375 : uniform
 * without replacement; all neighbours in CSR order when degree <= fanout.  The three reference
 * variants differ only in how the deg > fanout subset is drawn; one kernel implements that law here
 * (position sampling, reads exactly min(deg, fanout) neighbours). `sample_type` is accepted for all
 * three (XG_KHOP0 / XG_KHOP2 / XG_KHOP3). */
int xg_sample_khop(const xg_graph *g, int sample_type, const uint32_t *d_input, size_t num_input,
                   size_t fanout, uint32_t *d_out_src, uint32_t *d_out_dst, size_t *d_num_out,
                   uint64_t seed, uint64_t batch_key, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Extraction.
 */
/* GPUExtract(dst, src, index, num_index, dim, dtype, ctx, stream), cuda/cuda_extraction.cu:74-116.
 * `src` may be device memory or pinned/registered host memory (UVA), as in DoGPULabelExtract
 * (dist/dist_loops.cc:938-974). */
int xg_extract(void *d_dst, const void *src, const uint32_t *d_index, size_t num_index, size_t dim,
               int dtype, void *stream);
/* GetMissCacheIndex + GPUExtractMissData + CombineCacheData in ONE launch
 * (cuda/cuda_cache_manager_device.cu:355-441,573-625,628-731): out[i] = row of d_nodes[i], routed per
 * xg_feature_store.  d_stats (optional, 4 x uint64 on device) accumulates rows served from
 * {local HBM, peer HBM, host} and total. */
int xg_extract_ggms(const xg_feature_store *store, const uint32_t *d_nodes, size_t num_nodes,
                    void *d_out, uint64_t *d_stats, void *stream);
/* GPUCacheManager::GetMissCacheIndex, cuda_cache_manager_device.cu:355-441: stable split.
 * Outputs need num_nodes entries each; counts are written to the host (synchronises). */
int xg_get_miss_cache_index(const uint32_t *d_cache_table, const uint32_t *d_nodes, size_t num_nodes,
                            uint32_t *d_miss_src, uint32_t *d_miss_dst, size_t *h_num_miss,
                            uint32_t *d_cache_src, uint32_t *d_cache_dst, size_t *h_num_cache,
                            void *stream);
/* GPUExtractMissData (cuda_cache_manager_device.cu:573-625): out[dst[i]] = host_feat[src[i]] */
int xg_extract_miss_data(void *d_out, const void *host_feat, const uint32_t *d_miss_src,
                         const uint32_t *d_miss_dst, size_t num_miss, size_t row_bytes, void *stream);
/* CombineCacheData (cuda_cache_manager_device.cu:628-731): out[dst[i]] = cache row rank src[i],
 * routed rank % P / rank / P over store->part[] */
int xg_combine_cache_data(void *d_out, const xg_feature_store *store, const uint32_t *d_cache_src,
                          const uint32_t *d_cache_dst, size_t num_cache, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* XGNN_B200_H */
