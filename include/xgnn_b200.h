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
 * B200 extension (num_replicated > 0): the hottest cache ranks r < num_replicated are REPLICATED in every GPU's
 * own HBM (`replica`, row r) and only the remaining ranks r' = r - num_replicated are partitioned
 * (part r' % num_part, row r' / num_part).  In the reference's vocabulary this is a cache split into a
 * clique_size = 1 tier and a clique_size = P tier (run_config `clique_size`, cuda/dist_graph.cu:392-480);
 * num_replicated = 0 is exactly the reference's DeviceDistFeature routing (cuda/dist_graph.h:203-207).
 */
typedef struct xg_feature_store {
  const void *part[XG_MAX_PARTS];
  const void *host;
  const uint32_t *cache_table;
  uint32_t num_part;
  uint32_t local_part_mask; /* bit p set: part p is in this GPU's HBM (statistics only) */
  uint64_t row_bytes;       /* dim * sizeof(dtype) */
  const void *replica;      /* rows of cache ranks [0, num_replicated) in THIS GPU's HBM, or NULL */
  uint32_t num_replicated;
  uint32_t compact_bitwidth; /* 0: rank routing above.  b > 0 (needs num_replicated == 0): a cache_table value is
                              * (part << (32 - b)) | row, the ICS'22 clique placement's compact map
                              * (cuda/ics22_song_dist_feature.h:38-47, cuda_cache_manager_host.cc:203-222) */
  const void *replica_words; /* optional accelerator for the replicated tier, or NULL: uint2[ceil(num_node / 32)] from
                              * xg_build_replica_words.  Needs the replica laid out by node id
                              * (cache_table[v] = number of replicated nodes with a smaller id). */
  uint32_t replica_words_cover_all; /* 1 (as reported by xg_build_replica_words): every node is cached and the sharded tier is
                                     * in node-id order as well, cache_table[v] = num_replicated + number of NON-replicated
                                     * nodes with a smaller id - then no lookup reads the table at all */
  uint32_t host_row_mask; /* 0: `host` has one row per node.  2^k - 1: SAMGRAPH_EMPTY_FEAT=k mock features - `host` has 2^k
                           * rows and a host-resident node v reads row v & mask (gpu_mock_extract, cuda/cuda_extraction.cu:
                           * 51-72; extract_miss_data's mask, cuda_cache_manager_device.cu:589-592) */
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
 * three (XG_KHOP0 / XG_KHOP2 / XG_KHOP3).
 * CONCURRENCY of the standalone operators xg_sample_khop, xg_sample_khop1, xg_sample_weighted_khop,
 * xg_sample_random_walk and xg_get_miss_cache_index: each keeps ONE look-back scan workspace per device, so calls
 * on the same device must be issued on ONE stream (or be otherwise ordered); the reference's operators have the same
 * shape of restriction (one workspace pool per device, one sampler stream).  The minibatch engine (xg_sampler_*) and
 * xg_hashtable_* carry their own per-object workspaces and have no such limit. */
int xg_sample_khop(const xg_graph *g, int sample_type, const uint32_t *d_input, size_t num_input,
                   size_t fanout, uint32_t *d_out_src, uint32_t *d_out_dst, size_t *d_num_out,
                   uint64_t seed, uint64_t batch_key, void *stream);

/* GPUSampleKHop1, cuda_sampling_khop1.cu:130-236: `fanout` draws with replacement per node; output grouped
 * by ascending GLOBAL src id (stable), adjacent equal (src,dst) pairs collapsed. */
int xg_sample_khop1(const xg_graph *g, const uint32_t *d_input, size_t num_input, size_t fanout,
                    uint32_t *d_out_src, uint32_t *d_out_dst, size_t *d_num_out, uint64_t seed, uint64_t batch_key,
                    void *stream);
/* GPUSampleWeightedKHop (cuda_sampling_weighted_khop.cu:132-), ...Prefix (cuda_sampling_weighted_khop_prefix.cu
 * :148-), ...HashDedup (cuda_sampling_weighted_khop_hash_dedup.cu:203-), selected by sample_type
 * (XG_WEIGHTED_KHOP / _PREFIX / _HASH_DEDUP).  Tables are aligned with the flat CSR's indices
 * (prob_table.bin f32, alias_table.bin u32 node ids, prob_prefix_table.bin f32); pass NULL for the ones the
 * variant does not use.  The graph must be flat (num_part == 0), as in the reference. */
int xg_sample_weighted_khop(const xg_graph *g, int sample_type, const float *d_prob_table,
                            const uint32_t *d_alias_table, const float *d_prob_prefix_table, const uint32_t *d_input,
                            size_t num_input, size_t fanout, uint32_t *d_out_src, uint32_t *d_out_dst,
                            size_t *d_num_out, uint64_t seed, uint64_t batch_key, void *stream);
/* GPUSampleRandomWalk<GraphType> + FrequencyHashmap::GetTopK, cuda_sampling_random_walk.cu:116-165,
 * cuda_frequency_hashmap.cu:643-841.  Out arrays need num_input * K entries; d_out_data = visit counts. */
int xg_sample_random_walk(const xg_graph *g, const uint32_t *d_input, size_t num_input, size_t random_walk_length,
                          double random_walk_restart_prob, size_t num_random_walk, size_t K, uint32_t *d_out_src,
                          uint32_t *d_out_dst, uint32_t *d_out_data, size_t *d_num_out, uint64_t seed,
                          uint64_t batch_key, void *stream);
/* Host-only test hooks: the four 32-bit words the kernels draw for Philox counter (c0,c1,c2,c3) under
 * (seed, batch_key), and the third counter word the standalone samplers use.  Counters: KHop1 / weighted
 * (input position, slot, salt, 0); hash-dedup (input position, try, salt, 1); random walk (input position,
 * walk, step, salt). */
void xg_philox4x32(uint64_t seed, uint64_t batch_key, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                   uint32_t out[4]);
uint32_t xg_sample_salt(void);

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
/* xg_get_miss_cache_index for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Same split, same output arrays (num_nodes entries each); the two counts go to caller memory on the DEVICE
 * (d_counts[0] = misses, d_counts[1] = cache hits) and the call neither synchronises nor allocates: its scan scratch is
 * a file-scope __device__ array of XG_NOALLOC_MAX_NODES / 1024 + 4 words of 8 bytes (128 KB) that every GPU loading
 * the library holds until exit.  num_nodes > XG_NOALLOC_MAX_NODES -> XG_ERR_INVALID (xg_last_error says so).
 * All calls on one device share that array: calls that overlap on the GPU (different streams without an ordering
 * between them) race; calls on one stream, or ordered by events, are safe. */
#define XG_NOALLOC_MAX_NODES (1u << 24)
int xg_get_miss_cache_index_noalloc(const uint32_t *d_cache_table, const uint32_t *d_nodes, size_t num_nodes,
                                    uint32_t *d_miss_src, uint32_t *d_miss_dst, uint32_t *d_cache_src,
                                    uint32_t *d_cache_dst, uint64_t *d_counts, void *stream);
/* GPUMockExtract (cuda/cuda_extraction.cu:118-160; CPU twin cpu/cpu_extraction.cc:47-64): dst[i] = src[index[i] & (2^k - 1)],
 * the SAMGRAPH_EMPTY_FEAT=k mock-feature mode of the reference's evaluation scripts; k = 0 is plain GPUExtract */
int xg_mock_extract(void *d_dst, const void *src, const uint32_t *d_index, size_t num_index, size_t row_bytes,
                    unsigned empty_feat_bits, void *stream);
/* GPUExtractMissData (cuda_cache_manager_device.cu:573-625): out[dst[i]] = host_feat[src[i]] */
int xg_extract_miss_data(void *d_out, const void *host_feat, const uint32_t *d_miss_src,
                         const uint32_t *d_miss_dst, size_t num_miss, size_t row_bytes, void *stream);
/* CombineCacheData (cuda_cache_manager_device.cu:628-731): out[dst[i]] = cache row rank src[i],
 * routed rank % P / rank / P over store->part[] */
int xg_combine_cache_data(void *d_out, const xg_feature_store *store, const uint32_t *d_cache_src,
                          const uint32_t *d_cache_dst, size_t num_cache, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Minibatch engine: one call enqueues what dist::DoGPUSample (dist/dist_loops.cc:62-368),
 * DoGPULabelExtract (:938-974), DoArch6GetCacheMissIndex (:1015-1074) and
 * DoArch6GPUCacheFeatureCopy (:1209-1285) do for one Task -- hash table Reset, seed fill, per layer
 * sample + unique + remap, then the routed feature gather and the label gather -- as ONE replayed
 * CUDA graph with every count kept on the device; the only host round trip is the final 128-byte
 * read of the counts.
 */
#define XG_MAX_LAYERS 8
#define XG_SAMPLER_NO_GRAPH 1u /* flags: launch the kernels one by one instead of replaying a graph */
#define XG_SAMPLER_STAGE_TIMING 2u /* flags: (implies NO_GRAPH) CUDA events around every stage */
#define XG_SAMPLER_SINGLE_STREAM 4u /* flags: sample and extract phases on ONE stream (default: the extract phase of
                                       batch i runs on its own stream while batch i+1 is sampled) */

/* stages of one minibatch, for xg_batch_stage_ms */
enum {
  XG_STAGE_RESET_SEEDS = 0, /* Reset + FillWithDupRevised(seeds) */
  XG_STAGE_SAMPLE = 1,      /* per layer: sample kernel */
  XG_STAGE_INSERT = 2,      /* per layer: every sampled neighbour into the hash table (first occurrence wins) */
  XG_STAGE_NUMBER = 3,      /* per layer: number first occurrences (unique list append) + neighbour local ids */
  XG_STAGE_INPUT_NODES = 4, /* input_nodes copy + counts */
  XG_STAGE_FEATURE = 5,     /* routed feature gather */
  XG_STAGE_LABEL = 6,       /* label gather */
};

typedef struct xg_sampler_config {
  uint32_t num_layer;
  uint32_t fanout[XG_MAX_LAYERS]; /* indexed by layer; consumed last -> first (dist_loops.cc:120-123) */
  uint32_t batch_size;            /* most seeds one minibatch can carry */
  int32_t sample_type;            /* any SampleType; KHop0/2/3 run the fused sample+insert kernel */
  uint32_t num_slots;             /* minibatches that may be alive at once; 0 -> 2 */
  uint32_t flags;
  uint32_t feat_dim;
  int32_t feat_dtype;             /* XG_F32 ...; row_bytes of the store must equal dim * sizeof(dtype) */
  uint64_t seed;
  uint64_t max_input_nodes;       /* 0 -> PredictNumNodes(batch_size, fanout) (common.cc:488-497),
                                     clipped to the number of nodes */
  void *stream;                   /* cudaStream_t all work is enqueued on; NULL -> a private stream */
  /* sample_type == XG_RANDOM_WALK (operation.cc:163-175): fanout[l] = num_neighbor (K) for every layer */
  uint32_t random_walk_length;
  uint32_t num_random_walk;
  double random_walk_restart_prob;
  /* weighted samplers: tables aligned with the flat CSR (NULL otherwise) */
  const float *prob_table;
  const uint32_t *alias_table;
  const float *prob_prefix_table;
} xg_sampler_config;

/* One finished Task (common.h:246-285). Pointers are device pointers owned by the batch; they stay
 * valid until xg_batch_release. Layer l = TrainGraph graphs[l]: graphs[num_layer-1] is the seed-side
 * block, graphs[0] the input-side block (num_src[0] == num_input). */
typedef struct xg_batch_view {
  uint64_t key;
  uint32_t num_layer;
  uint32_t num_output; /* seeds */
  uint32_t num_input;  /* input_nodes */
  uint32_t reserved;
  const uint32_t *row[XG_MAX_LAYERS];  /* local id of the sampled neighbour, < num_src */
  const uint32_t *col[XG_MAX_LAYERS];  /* local id of the expanded node,     < num_dst */
  const uint32_t *data[XG_MAX_LAYERS]; /* per-edge payload (random-walk visit counts) or NULL */
  uint64_t num_src[XG_MAX_LAYERS];
  uint64_t num_dst[XG_MAX_LAYERS];
  uint64_t num_edge[XG_MAX_LAYERS];
  const uint32_t *input_nodes;
  const uint32_t *output_nodes;
  const void *input_feat;      /* [num_input, feat_dim] or NULL when no feature store was given */
  const int64_t *output_label; /* [num_output] or NULL */
  /* KHop0/2/3 only (edges node-major in input order), else NULL (also NULL for the seed layer of a batch whose
     seeds contain duplicates: they share a local id, so that layer is not destination-major): csc_indptr[l][0..num_dst[l]] are the edge offsets
     of the layer's destination nodes, i.e. (csc_indptr, row) is the block in CSC form - what DGL's
     create_unitgraph_from_coo + format conversion computes on the caller side (torch/adapter.py:98-124). */
  const uint32_t *csc_indptr[XG_MAX_LAYERS];
} xg_batch_view;

typedef struct xg_sampler xg_sampler;
typedef struct xg_batch xg_batch;

/* `store` and `label` may be NULL (sampling only). `label` = int64[num_node], device or pinned host. */
int xg_sampler_create(xg_sampler **out, const xg_graph *graph, const xg_feature_store *store,
                      const void *label, const xg_sampler_config *cfg, int device);
int xg_sampler_destroy(xg_sampler *s);
/* Enqueue one minibatch for `num_seeds` seeds (host pointer, or device pointer when
 * seeds_on_device != 0). Returns at once; *out is claimed until xg_batch_release. */
int xg_sampler_submit(xg_sampler *s, const uint32_t *seeds, size_t num_seeds, int seeds_on_device,
                      uint64_t batch_key, xg_batch **out);
/* Block until the batch is complete, then describe it. XG_ERR_CAPACITY if it outgrew its buffers. */
int xg_batch_wait(xg_batch *b, xg_batch_view *view);
int xg_batch_release(xg_batch *b);
/* stream the sampler enqueues on (for events) and kernels per minibatch (for launch accounting) */
void *xg_sampler_stream(xg_sampler *s);
/* stream of the extract phase (feature/label gather, count read-back); == xg_sampler_stream in single-stream
   mode. A batch is complete when xg_batch_wait returns; for stream-ordered timing record on this stream. */
void *xg_sampler_extract_stream(xg_sampler *s);
/* diagnostics (env XG_TIMELINE=1 at sampler creation): device timestamps, in ms since the sampler was created,
   of {sample phase begin, sample phase end, extract phase begin, extract phase end} of a finished batch */
int xg_batch_timeline(xg_batch *b, float ms[4]);
uint32_t xg_sampler_launches_per_batch(xg_sampler *s);
/* device time of one stage of a finished batch (sampler created with XG_SAMPLER_STAGE_TIMING);
 * `layer` is ignored for the stages that are not per layer */
int xg_batch_stage_ms(xg_batch *b, int stage, int layer, float *ms);
/* PredictNumNodes, common.cc:488-497 */
size_t xg_predict_num_nodes(size_t batch_size, const uint32_t *fanout, size_t num_layer);

/* ------------------------------------------------------------------------------------------------
 * GGMS store plumbing (host side).
 */
/* std::shuffle(nodes, nodes + n, std::mt19937(n)): the cache-rank order of GPUCacheManager,
 * cuda/cuda_cache_manager_host.cc:162-171 (same libstdc++ algorithm, hence the same order). */
int xg_shuffle_cache_ranks(uint32_t *h_nodes, size_t num_cached);
/* Rank/select words of the replicated tier: word w = { bit i: node 32w+i is replicated (cache_table < num_replicated),
 * number of replicated nodes with id < 32w }.  8 B per 32 nodes, so L2-resident where the 4 B x V table is not: the
 * gather resolves replicated nodes from these words alone and only reads the table for the rest.  Fails with
 * XG_ERR_INVALID when the replica is not in node-id order.  *h_cover_all (optional) = 1 when, in addition, every node is
 * cached and the sharded tier is in node-id order too (see xg_feature_store.replica_words_cover_all).
 * Init-time (allocates scratch, synchronises `stream`). */
int xg_build_replica_words(const uint32_t *d_cache_table, uint32_t num_node, uint32_t num_replicated, void *d_words,
                           uint32_t *h_cover_all, void *stream);

/* ICS'22 clique feature placement (comparison baseline): ICS22SongPlacementSolver, cuda/ics22_song_dist_feature.cc:22-167,
 * seen from ONE worker.  h_rank_nodes / h_rank_weight: all nodes by weight descending (ties: node id descending) and
 * their weights - the reference uses the in-degree (ics22_song_dist_feature.cc:196-213); xg_presample_accumulate over
 * `indices` + xg_presample_rank produce exactly this pair.  Every GPU owns num_cached row slots; clique_size GPUs
 * (worker / clique_size is the clique) split the coldest slots among themselves while
 * weight(candidate) >= alpha * weight(slot's node).  Outputs, host memory:
 *   h_cache_table u32[num_node]  (owner GPU << (32 - compact_bitwidth)) | row, XG_EMPTY_KEY = not cached
 *                                (the compact map of cuda_cache_manager_host.cc:203-222; xg_feature_store.compact_bitwidth)
 *   h_slot_nodes  u32[num_cached] node whose row sits in slot i of THIS worker's cache
 *   h_total_cached               distinct nodes cached by the clique (GetRealCachedNodeNum) */
int xg_ics22_placement(const uint32_t *h_rank_nodes, const uint32_t *h_rank_weight, uint32_t num_node,
                       uint32_t num_cached, uint32_t num_dev, uint32_t clique_size, double alpha, uint32_t worker,
                       uint32_t compact_bitwidth, uint32_t *h_cache_table, uint32_t *h_slot_nodes,
                       uint32_t *h_total_cached);
/* smallest n with indptr[n] >= floor(E * ratio): DistEngine::SampleDataCopy, dist/dist_engine.cc:223-232 */
uint32_t xg_num_cache_node(const uint32_t *h_indptr, uint32_t num_node, double ratio);
/* DistGraph::_DatasetPartition, cuda/dist_graph.cu:228-272, as two device passes around the caller's
 * exclusive scan: rows = ceil((num_cache_node - part) / P); row j is node part + j * P. */
size_t xg_part_num_rows(uint32_t num_cache_node, uint32_t part, uint32_t num_part);
int xg_part_degrees(const uint32_t *d_indptr, uint32_t num_cache_node, uint32_t part, uint32_t num_part,
                    uint32_t *d_deg_out, void *stream);
int xg_part_copy_indices(const uint32_t *d_indptr, const uint32_t *d_indices, uint32_t num_cache_node, uint32_t part,
                         uint32_t num_part, const uint32_t *d_part_indptr, uint32_t *d_part_indices, void *stream);
/* DistGraph::_DataIpcShare, cuda/dist_graph.cu:274-307: shards come from their own cudaMalloc so that
 * a 64-byte CUDA IPC handle names exactly one shard; peers map it and load from it inside kernels. */
int xg_device_malloc(void **out, size_t bytes);
int xg_device_free(void *p);
int xg_ipc_export(const void *d_ptr, unsigned char out_handle[64]);
int xg_ipc_import(const unsigned char handle[64], void **out);
int xg_ipc_close(void *p);
int xg_enable_peer_access(int device, int peer);

/* Presampling cache-rank policy (kCacheByPreSample): dist::PreSampler, dist/pre_sampler.cc:40-165.
 * accumulate: freq[v] += 1 for every v of one batch's input_nodes (count from *d_num_nodes when given, else
 * num_nodes_max).  rank: nodes ordered by frequency descending, ties by node id descending - the order of
 * __gnu_parallel::sort(std::greater<>) over {low: id, high: freq} words (pre_sampler.cc:113-119); d_rank_freq may be
 * NULL (GetFreq, :131-139).  rank synchronises the stream. */
int xg_presample_accumulate(uint32_t *d_freq, uint32_t num_node, const uint32_t *d_nodes, size_t num_nodes_max,
                            const uint32_t *d_num_nodes, void *stream);
int xg_presample_rank(const uint32_t *d_freq, uint32_t num_node, uint32_t *d_rank_nodes, uint32_t *d_rank_freq,
                      void *stream);

/* SAMGRAPH_SANITY_CHECK: GPUBatchSanityCheck / GPUSanityCheckList (cuda/cuda_sanity_check.cu:28-88).  Violations are
 * added to *d_violations (device u64) instead of asserting in the kernel.  ids: kEmptyKey, id >= num_node, or - when
 * d_epoch_map (u32[num_node], zeroed at the start of an epoch) is given - an id already seen this epoch.
 * edges: row[e] >= num_src or col[e] >= num_dst. */
int xg_sanity_check_ids(const uint32_t *d_ids, size_t n, uint32_t num_node, uint32_t *d_epoch_map,
                        uint64_t *d_violations, void *stream);
int xg_sanity_check_edges(const uint32_t *d_row, const uint32_t *d_col, size_t n, uint32_t num_src, uint32_t num_dst,
                          uint64_t *d_violations, void *stream);

/* DistAlignedShuffler::ReShuffle, dist/dist_shuffler_aligned.cc:75-121: one epoch's Fisher-Yates pass with
 * std::default_random_engine(seed = epoch), in place over the padded train set (host memory).  The engine
 * applies it epoch after epoch to the same array, as the reference does. */
int xg_aligned_shuffle_epoch(uint32_t *h_data, size_t n, uint64_t epoch);
/* PartitionSolver::solve, cuda/dist_graph.cu:728-777 (search: :40-222; standalone twin
 * utility/nvlink_placement_solver/solver.cc).  bandwidth = num_gpu x num_gpu row-major GB/s, 0 = no P2P
 * path (the diagonal is the local bandwidth).  Outputs, all num_gpu x num_gpu row-major ints except
 * out_num_parts[num_gpu]:  out_part_ids[g][0..out_num_parts[g]) = parts GPU g stores (ascending, rest -1)
 * = GroupConfig::part_ids;  out_access[g][p] = GPU that g reads part p from = GroupConfig::ctx_group
 * (cuda/dist_graph.h:217-229).  Same search order and tie-breaks as the reference, hence the same
 * placement; on an all-to-all matrix (NVSwitch) the result is the identity placement. */
int xg_partition_solve(const double *bandwidth, int num_gpu, int *out_num_parts, int *out_part_ids,
                       int *out_access);
/* PartitionSolver::DetectTopo_child, cuda/dist_graph.cu:779-884: out_bandwidth[d][p] = GB/s of a timed
 * 128 MiB copy p -> d (0 when the pair has no P2P path); needs the GPUs. */
int xg_detect_topology(int num_gpu, double *out_bandwidth);

/* ------------------------------------------------------------------------------------------------
 * Synthetic datasets of the BASELINE.json shapes, generated in HBM (stand-in for datagen/ and
 * utility/data-process/, which need the network). Pure functions of (node, slot, seed); a shard is
 * the arithmetic sequence v = begin + i * stride (the GGMS partition v % P == p). numpy twins:
 * tests/synth.py.
 *   degree(v)       = min(max_deg, floor(c * sqrt(num_node / (v + 1))))
 *   neighbour(v, j) = floor(num_node * u^2), u = hash(v, j, seed) / 2^32
 *   feature(v, col) = ((v * K1 + col * K2) >> 40) % 2048 / 8 - 128      (exact in fp16 / fp32)
 */
int xg_synth_degrees(uint32_t *d_deg, uint64_t begin, uint64_t stride, size_t n, uint64_t num_node, double c,
                     uint32_t max_deg, void *stream);
/* d_indptr: the shard's own (rebased) CSR offsets, n + 1 entries */
int xg_synth_indices(uint32_t *d_indices, const uint32_t *d_indptr, uint64_t begin, uint64_t stride, size_t n,
                     uint64_t num_node, uint64_t seed, void *stream);
/* rows = d_row_ids[i] when given, else begin + i * stride; dtype XG_F32 or XG_F16 */
int xg_synth_features(void *d_out, const uint32_t *d_row_ids, uint64_t begin, uint64_t stride, size_t n_rows,
                      uint32_t dim, int dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* XGNN_B200_H */
