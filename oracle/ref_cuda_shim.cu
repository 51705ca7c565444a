// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Glue that lets the reference's own CUDA samplers be compiled *in place* from /root/reference
// (oracle/Makefile, target `refcuda`) into oracle/_ref/libxgnn_ref_cuda.so and driven from
// tests/ as a second oracle for the sampler families whose CPU files are empty stubs
// (cpu/cpu_sampling_khop1.cc:24-27, cpu_sampling_weighted_khop.cc:24-28,
// cpu_sampling_random_walk.cc:24-28): KHop1, weighted KHop (alias / prefix / hash-dedup),
// random walk + FrequencyHashmap::GetTopK, plus the GPU KHop0 / KHop3 kernels.
//
// Nothing here restates reference logic.  This file provides
//   * a cudaMalloc-backed `Device` (the reference's GPUDevice pulls in its workspace pool,
//     engine and run-config machinery: cuda/cuda_device.cc),
//   * no-op `Profiler` methods (profiler.cc needs the Engine singleton),
//   * a flat extern "C" interface: host arrays in, host arrays out.
//
// The reference seeds curand from the wall clock (cuda/cuda_random_states.cu:106-108), so its
// output is not reproducible: tests compare distributions (two-sample chi-square) and the
// cases where the result does not depend on the draws.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <tuple>
#include <vector>

#include "common/common.h"
#include "common/constant.h"
#include "common/cuda/cuda_cache_manager.h"
#include "common/cuda/cuda_frequency_hashmap.h"
#include "common/cuda/cuda_function.h"
#include "common/cuda/cuda_hashtable.h"
#include "common/cuda/cuda_random_states.h"
#include "common/cuda/dist_graph.h"
#include "common/cuda/ics22_song_dist_feature.h"
#include "common/device.h"
#include "common/logging.h"
#include "common/profiler.h"
#include "common/run_config.h"

#define SHIM_CUDA(x)                                                                      \
  do {                                                                                    \
    cudaError_t e_ = (x);                                                                 \
    if (e_ != cudaSuccess) {                                                              \
      fprintf(stderr, "ref_cuda_shim %s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      abort();                                                                            \
    }                                                                                     \
  } while (0)

namespace samgraph {
namespace common {

namespace {
// GPU contexts: cudaMalloc; CPU contexts (the host-side solvers in ref_host_shim.cc): malloc.
class ShimDevice : public Device {
 public:
  void SetDevice(Context ctx) override {
    if (ctx.device_type == kGPU) SHIM_CUDA(cudaSetDevice(ctx.device_id));
  }
  void *AllocDataSpace(Context ctx, size_t nbytes, size_t) override {
    void *p = nullptr;
    if (nbytes == 0) nbytes = 64;
    if (ctx.device_type == kGPU) {
      SHIM_CUDA(cudaMalloc(&p, nbytes));
    } else if (posix_memalign(&p, 64, nbytes) != 0) {
      abort();
    }
    return p;
  }
  void FreeDataSpace(Context ctx, void *ptr) override {
    if (ctx.device_type == kGPU) SHIM_CUDA(cudaFree(ptr));
    else free(ptr);
  }
  size_t WorkspaceActualSize(Context, void *) override { return 0; }
  void CopyDataFromTo(const void *from, size_t from_offset, void *to, size_t to_offset,
                      size_t nbytes, Context ctx_from, Context ctx_to, StreamHandle stream) override {
    char *dst = static_cast<char *>(to) + to_offset;
    const char *src = static_cast<const char *>(from) + from_offset;
    if (ctx_from.device_type != kGPU && ctx_to.device_type != kGPU) {
      memcpy(dst, src, nbytes);
      return;
    }
    SHIM_CUDA(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDefault, static_cast<cudaStream_t>(stream)));
    SHIM_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  }
  void StreamSync(Context ctx, StreamHandle stream) override {
    if (ctx.device_type == kGPU) SHIM_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  }
  void SyncDevice(Context ctx) override {
    if (ctx.device_type == kGPU) SHIM_CUDA(cudaDeviceSynchronize());
  }
};
ShimDevice g_dev;
}  // namespace

Device *Device::Get(Context) { return &g_dev; }
void *Device::AllocWorkspace(Context ctx, size_t nbytes, double) {
  return AllocDataSpace(ctx, nbytes, kTempAllocaAlignment);
}
void Device::FreeWorkspace(Context ctx, void *ptr, size_t) { FreeDataSpace(ctx, ptr); }
StreamHandle Device::CreateStream(Context) { return nullptr; }
void Device::FreeStream(Context, StreamHandle) {}
void Device::SyncStreamFromTo(Context, StreamHandle, StreamHandle) {}

Profiler::Profiler() {}
Profiler &Profiler::Get() {
  static Profiler p;
  return p;
}
void Profiler::LogStep(uint64_t, LogStepItem, double) {}
void Profiler::LogStepAdd(uint64_t, LogStepItem, double) {}
void Profiler::LogEpochAdd(uint64_t, LogEpochItem, double) {}
void Profiler::LogInit(LogInitItem, double) {}
void Profiler::LogInitAdd(LogInitItem, double) {}
void Profiler::ResetStepEpoch() {}

}  // namespace common
}  // namespace samgraph

using namespace samgraph::common;

namespace {

template <typename T>
T *to_device(const T *h, size_t n) {
  T *d = nullptr;
  SHIM_CUDA(cudaMalloc(&d, (n ? n : 1) * sizeof(T)));
  if (n) SHIM_CUDA(cudaMemcpy(d, h, n * sizeof(T), cudaMemcpyHostToDevice));
  return d;
}

struct RefCudaGraph {
  int device = 0;
  uint32_t num_node = 0;
  size_t num_edge = 0;
  uint32_t *indptr = nullptr, *indices = nullptr, *alias = nullptr;
  float *prob = nullptr, *prefix = nullptr;
  // curand states live as long as the engine in the reference (dist/dist_engine.cc:431-433) and every kernel
  // stores them back, so batch k continues where batch k-1 stopped: one object per (sampler, fanout, batch size).
  std::map<std::tuple<int, size_t, size_t>, cuda::GPURandomStates *> states;
  // optional GGMS partitioning of the same graph: DeviceDistGraph's pointer tables (cuda/dist_graph.h:114-151)
  uint32_t num_part = 0, num_cache_node = 0;
  uint32_t **d_part_indptr = nullptr, **d_part_indices = nullptr;
  std::vector<uint32_t *> part_bufs;
};

}  // namespace

// The one thing the partitioned feature cache adds to CombineCacheData is WHERE a cache rank lives: the reference's
// combine_cache_data_for_partition<T, Feat> (cuda_cache_manager_device.cu:277-299, file-local) reads
// Feat::Get<T>(rank, col).  This kernel calls the reference's own device accessors - DeviceDistFeature::Get
// (cuda/dist_graph.h:193-207: part rank % P, row rank / P) and DeviceICS22SongDistFeature::Get
// (cuda/ics22_song_dist_feature.h:27-47: owner in the top bits, or the device_map / new_idx_map pair) - for every
// (rank, column); the parts are laid out by the caller exactly as DistGraph::FeatureLoad leaves them
// (dist_graph.cu:492-505: part p holds cache_rank_node[p], [p + P], ... in that order).
template <typename T, typename Feat>
__global__ void shim_dist_feature_rows(Feat feat, const samgraph::common::IdType *ranks, size_t n, uint32_t dim, T *out) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n * dim; i += (size_t)gridDim.x * blockDim.x) {
    out[i] = feat.template Get<T>(ranks[i / dim], (samgraph::common::IdType)(i % dim));
  }
}

extern "C" {

// Returns 0 when no CUDA device is usable (callers skip).
int refcu_available() {
  int n = 0;
  return cudaGetDeviceCount(&n) == cudaSuccess && n > 0;
}

void *refcu_graph_create(int device, const uint32_t *indptr, const uint32_t *indices,
                         uint32_t num_node, size_t num_edge, const float *prob_table,
                         const uint32_t *alias_table, const float *prob_prefix_table) {
  SHIM_CUDA(cudaSetDevice(device));
  auto *g = new RefCudaGraph();
  g->device = device;
  g->num_node = num_node;
  g->num_edge = num_edge;
  g->indptr = to_device(indptr, static_cast<size_t>(num_node) + 1);
  g->indices = to_device(indices, num_edge);
  if (prob_table) g->prob = to_device(prob_table, num_edge);
  if (alias_table) g->alias = to_device(alias_table, num_edge);
  if (prob_prefix_table) g->prefix = to_device(prob_prefix_table, num_edge);
  return g;
}

void refcu_graph_destroy(void *h) {
  auto *g = static_cast<RefCudaGraph *>(h);
  cudaFree(g->indptr);
  cudaFree(g->indices);
  cudaFree(g->prob);
  cudaFree(g->alias);
  cudaFree(g->prefix);
  for (auto *b : g->part_bufs) cudaFree(b);
  cudaFree(g->d_part_indptr);
  cudaFree(g->d_part_indices);
  for (auto &kv : g->states) delete kv.second;
  delete g;
}

// Attach a GGMS partitioning of the graph: part p = CSR of the nodes v < num_cache_node with v % P == p (rebased, as
// DistGraph::_DatasetPartition leaves it, dist_graph.cu:228-272); slot P = the whole CSR (pinned host memory in the
// reference, the device copy here: the routing is the same).  After this call KHop0 / KHop3 / random walk run through
// the reference's DeviceDistGraph instead of DeviceNormalGraph.
void refcu_graph_set_parts(void *h, uint32_t num_part, uint32_t num_cache_node, const uint32_t *indptr_concat,
                           const size_t *indptr_off, const uint32_t *indices_concat, const size_t *indices_off) {
  auto *g = static_cast<RefCudaGraph *>(h);
  SHIM_CUDA(cudaSetDevice(g->device));
  std::vector<uint32_t *> ip(num_part + 1), ix(num_part + 1);
  for (uint32_t p = 0; p < num_part; ++p) {
    ip[p] = to_device(indptr_concat + indptr_off[p], indptr_off[p + 1] - indptr_off[p]);
    ix[p] = to_device(indices_concat + indices_off[p], indices_off[p + 1] - indices_off[p]);
    g->part_bufs.push_back(ip[p]);
    g->part_bufs.push_back(ix[p]);
  }
  ip[num_part] = g->indptr;
  ix[num_part] = g->indices;
  g->d_part_indptr = to_device(ip.data(), ip.size());
  g->d_part_indices = to_device(ix.data(), ix.size());
  g->num_part = num_part;
  g->num_cache_node = num_cache_node;
}

// Forget the curand states: the next call re-seeds from the wall clock like a fresh engine.
void refcu_reset_states(void *h) {
  auto *g = static_cast<RefCudaGraph *>(h);
  for (auto &kv : g->states) delete kv.second;
  g->states.clear();
}

// One call of a reference GPU sampler.  sample_type uses the reference numbering (common.h:50-59):
// 0 KHop0, 1 KHop1, 2 WeightedKHop, 3 RandomWalk, 4 WeightedKHopPrefix, 5 KHop2,
// 6 WeightedKHopHashDedup, 7 KHop3.  For random walk `fanout` is K (num_neighbor) and
// out_data receives the visit counts.  out_* are host arrays of capacity num_input * fanout.
// Returns the number of output edges.
size_t refcu_sample(void *h, int sample_type, const uint32_t *input, size_t num_input,
                    size_t fanout, size_t random_walk_length, double restart_prob,
                    size_t num_random_walk, uint32_t *out_src, uint32_t *out_dst,
                    uint32_t *out_data) {
  auto *g = static_cast<RefCudaGraph *>(h);
  SHIM_CUDA(cudaSetDevice(g->device));
  Context ctx = GPU(g->device);
  RunConfig::num_random_walk = num_random_walk;
  RunConfig::random_walk_length = random_walk_length;
  RunConfig::random_walk_restart_prob = restart_prob;

  const size_t cap = num_input * fanout;
  uint32_t *d_input = to_device(input, num_input);
  uint32_t *d_src = nullptr, *d_dst = nullptr, *d_data = nullptr;
  size_t *d_num_out = nullptr;
  SHIM_CUDA(cudaMalloc(&d_src, (cap ? cap : 1) * 4));
  SHIM_CUDA(cudaMalloc(&d_dst, (cap ? cap : 1) * 4));
  SHIM_CUDA(cudaMalloc(&d_data, (cap ? cap : 1) * 4));
  SHIM_CUDA(cudaMalloc(&d_num_out, sizeof(size_t)));
  SHIM_CUDA(cudaMemset(d_num_out, 0, sizeof(size_t)));

  // GPURandomStates sizes itself from (sample_type, fanout list, batch size): with a one-layer
  // list {fanout} and batch = num_input it covers exactly the threads one call launches.
  std::vector<size_t> fan{fanout};
  auto key = std::make_tuple(sample_type, fanout, num_input);
  if (!g->states.count(key)) {
    g->states[key] = new cuda::GPURandomStates(static_cast<SampleType>(sample_type), fan, num_input, ctx);
    SHIM_CUDA(cudaDeviceSynchronize());
  }
  cuda::GPURandomStates &states = *g->states[key];
  StreamHandle stream = nullptr;
  cuda::DeviceNormalGraph graph(g->indptr, g->indices, g->num_node);
  if (g->num_part > 0 && (sample_type == kKHop0 || sample_type == kKHop3 || sample_type == kRandomWalk)) {
    cuda::DeviceDistGraph dgraph(g->d_part_indptr, g->d_part_indices, g->num_part, g->num_cache_node, g->num_node);
    if (sample_type == kKHop0) {
      cuda::GPUSampleKHop0(dgraph, d_input, num_input, fanout, d_src, d_dst, d_num_out, ctx, stream, &states, 0);
    } else if (sample_type == kKHop3) {
      cuda::GPUSampleKHop3(dgraph, d_input, num_input, fanout, d_src, d_dst, d_num_out, ctx, stream, &states, 0);
    } else {
      cuda::FrequencyHashmap fmap(num_input, num_random_walk * random_walk_length, ctx);
      cuda::GPUSampleRandomWalk(dgraph, d_input, num_input, random_walk_length, restart_prob, num_random_walk, fanout,
                                d_src, d_dst, d_data, d_num_out, &fmap, ctx, stream, &states, 0);
      SHIM_CUDA(cudaDeviceSynchronize());
    }
  } else
  switch (sample_type) {
    case kKHop0:
      cuda::GPUSampleKHop0(graph, d_input, num_input, fanout, d_src, d_dst, d_num_out, ctx, stream,
                           &states, 0);
      break;
    case kKHop1:
      cuda::GPUSampleKHop1(g->indptr, g->indices, d_input, num_input, fanout, d_src, d_dst,
                           d_num_out, ctx, stream, &states, 0);
      break;
    case kWeightedKHop:
      cuda::GPUSampleWeightedKHop(g->indptr, g->indices, g->prob, g->alias, d_input, num_input,
                                  fanout, d_src, d_dst, d_num_out, ctx, stream, &states, 0);
      break;
    case kWeightedKHopPrefix:
      cuda::GPUSampleWeightedKHopPrefix(g->indptr, g->indices, g->prefix, d_input, num_input,
                                        fanout, d_src, d_dst, d_num_out, ctx, stream, &states, 0);
      break;
    case kKHop2:
      cuda::GPUSampleKHop2(g->indptr, g->indices, d_input, num_input, fanout, d_src, d_dst,
                           d_num_out, ctx, stream, &states, 0);
      break;
    case kWeightedKHopHashDedup:
      cuda::GPUSampleWeightedKHopHashDedup(g->indptr, g->indices, g->prob, g->alias, d_input,
                                           num_input, fanout, d_src, d_dst, d_num_out, ctx, stream,
                                           &states, 0);
      break;
    case kKHop3:
      cuda::GPUSampleKHop3(graph, d_input, num_input, fanout, d_src, d_dst, d_num_out, ctx, stream,
                           &states, 0);
      break;
    case kRandomWalk: {
      // dist/dist_engine.cc:436-442
      cuda::FrequencyHashmap fmap(num_input, num_random_walk * random_walk_length, ctx);
      cuda::GPUSampleRandomWalk(graph, d_input, num_input, random_walk_length, restart_prob,
                                num_random_walk, fanout, d_src, d_dst, d_data, d_num_out, &fmap,
                                ctx, stream, &states, 0);
      SHIM_CUDA(cudaDeviceSynchronize());
      break;
    }
    default:
      fprintf(stderr, "refcu_sample: unknown sample type %d\n", sample_type);
      abort();
  }
  SHIM_CUDA(cudaDeviceSynchronize());
  size_t num_out = 0;
  SHIM_CUDA(cudaMemcpy(&num_out, d_num_out, sizeof(size_t), cudaMemcpyDeviceToHost));
  if (num_out > cap) {
    fprintf(stderr, "refcu_sample: num_out %zu > capacity %zu\n", num_out, cap);
    abort();
  }
  if (out_src) SHIM_CUDA(cudaMemcpy(out_src, d_src, num_out * 4, cudaMemcpyDeviceToHost));
  if (out_dst) SHIM_CUDA(cudaMemcpy(out_dst, d_dst, num_out * 4, cudaMemcpyDeviceToHost));
  if (out_data && sample_type == kRandomWalk)
    SHIM_CUDA(cudaMemcpy(out_data, d_data, num_out * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_input);
  cudaFree(d_src);
  cudaFree(d_dst);
  cudaFree(d_data);
  cudaFree(d_num_out);
  return num_out;
}

// ---------------------------------------------------------------------------------------------
// The reference's single-GPU feature cache: GPUCacheManager (cuda/cuda_cache_manager_host.cc:62-131 builds the
// direct-address table and the cached rows; cuda_cache_manager_device.cu:355-441 GetMissCacheIndex, :573-625
// GPUExtractMissData, :656-731 CombineCacheData).  Both files are compiled in place; this only feeds them.
struct RefCache {
  cuda::GPUCacheManager *mgr = nullptr;
  void *mem = nullptr;        // zeroed storage of the object: its first constructor leaves _dist_graph unset
  void *host_feat = nullptr;  // pinned copy of the feature table: extract_miss_data loads it from the GPU (UVA)
  int device = 0;
  size_t row_bytes = 0;
};

void *refcu_cache_create(int device, const void *feat, size_t num_node, size_t dim, int dtype, size_t elem_bytes,
                         const uint32_t *rank_nodes, double cache_percentage) {
  SHIM_CUDA(cudaSetDevice(device));
  auto *c = new RefCache();
  c->device = device;
  c->row_bytes = dim * elem_bytes;
  SHIM_CUDA(cudaHostAlloc(&c->host_feat, num_node * c->row_bytes, cudaHostAllocMapped));
  memcpy(c->host_feat, feat, num_node * c->row_bytes);
  if (RunConfig::omp_thread_num <= 0) RunConfig::omp_thread_num = 4;
  RunConfig::option_empty_feat = 0;
  c->mem = calloc(1, sizeof(cuda::GPUCacheManager));
  c->mgr = new (c->mem) cuda::GPUCacheManager(GPU(device), GPU(device), c->host_feat, static_cast<DataType>(dtype), dim,
                                              rank_nodes, num_node, cache_percentage);
  SHIM_CUDA(cudaDeviceSynchronize());
  return c;
}

void refcu_cache_destroy(void *h) {
  auto *c = static_cast<RefCache *>(h);
  c->mgr->~GPUCacheManager();
  free(c->mem);
  cudaFreeHost(c->host_feat);
  delete c;
}

// DoArch6GetCacheMissIndex + DoArch6GPUCacheFeatureCopy for one node list (dist/dist_loops.cc:1015-1074, 1209-1285): the
// split, then miss rows from the host table and cached rows from the cache into out_feat[n x row_bytes].
void refcu_cache_extract(void *h, const uint32_t *nodes, size_t n, void *out_feat, uint32_t *miss_src,
                         uint32_t *miss_dst, size_t *num_miss, uint32_t *cache_src, uint32_t *cache_dst,
                         size_t *num_cache) {
  auto *c = static_cast<RefCache *>(h);
  SHIM_CUDA(cudaSetDevice(c->device));
  uint32_t *d_nodes = to_device(nodes, n), *d_idx = nullptr;
  void *d_out = nullptr;
  SHIM_CUDA(cudaMalloc(&d_idx, (n ? n : 1) * 4 * sizeof(uint32_t)));
  SHIM_CUDA(cudaMalloc(&d_out, (n ? n : 1) * c->row_bytes));
  uint32_t *d_ms = d_idx, *d_md = d_idx + n, *d_cs = d_idx + 2 * n, *d_cd = d_idx + 3 * n;
  c->mgr->GetMissCacheIndex(d_ms, d_md, num_miss, d_cs, d_cd, num_cache, d_nodes, n, nullptr);
  c->mgr->GPUExtractMissData(d_out, d_ms, d_md, *num_miss, nullptr);
  c->mgr->CombineCacheData(d_out, d_cs, d_cd, *num_cache, nullptr);
  SHIM_CUDA(cudaDeviceSynchronize());
  SHIM_CUDA(cudaMemcpy(out_feat, d_out, n * c->row_bytes, cudaMemcpyDeviceToHost));
  SHIM_CUDA(cudaMemcpy(miss_src, d_ms, *num_miss * 4, cudaMemcpyDeviceToHost));
  SHIM_CUDA(cudaMemcpy(miss_dst, d_md, *num_miss * 4, cudaMemcpyDeviceToHost));
  SHIM_CUDA(cudaMemcpy(cache_src, d_cs, *num_cache * 4, cudaMemcpyDeviceToHost));
  SHIM_CUDA(cudaMemcpy(cache_dst, d_cd, *num_cache * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_nodes);
  cudaFree(d_idx);
  cudaFree(d_out);
}

// ---------------------------------------------------------------------------------------------
// The reference's GPU OrderedHashTable (cuda/cuda_hashtable.cu:699-1064) and GPUMapEdges (cuda/cuda_mapping.cu:50-80).
// Among duplicate occurrences of one id its winner is whichever CAS lands first (cuda_hashtable.cu:54-79), so tests
// compare sets / prefix / round trips there and exact arrays where the input of a fill has no duplicates.
struct RefTable {
  cuda::OrderedHashTable *t = nullptr;
  int device = 0;
};

void *refcu_table_create(int device, size_t size) {
  SHIM_CUDA(cudaSetDevice(device));
  auto *r = new RefTable();
  r->device = device;
  r->t = new cuda::OrderedHashTable(size, GPU(device), nullptr);
  SHIM_CUDA(cudaDeviceSynchronize());
  return r;
}
void refcu_table_destroy(void *h) {
  auto *r = static_cast<RefTable *>(h);
  delete r->t;
  delete r;
}
void refcu_table_reset(void *h) { static_cast<RefTable *>(h)->t->Reset(nullptr); }
size_t refcu_table_num_items(void *h) { return static_cast<RefTable *>(h)->t->NumItems(); }

// FillWithDuplicates (unique_kind 0) or FillWithUnique (1); unique_out (host, capacity >= NumItems after the fill)
// receives the whole unique list.
size_t refcu_table_fill(void *h, int unique_kind, const uint32_t *input, size_t n, uint32_t *unique_out, size_t cap) {
  auto *r = static_cast<RefTable *>(h);
  SHIM_CUDA(cudaSetDevice(r->device));
  uint32_t *d_in = to_device(input, n), *d_unique = nullptr;
  SHIM_CUDA(cudaMalloc(&d_unique, (cap ? cap : 1) * 4));
  IdType num_unique = 0;
  if (unique_kind == 1) {
    r->t->FillWithUnique(d_in, n, nullptr);
    num_unique = r->t->NumItems();
    r->t->CopyUnique(d_unique, nullptr);
  } else {
    r->t->FillWithDuplicates(d_in, n, d_unique, &num_unique, nullptr);
  }
  SHIM_CUDA(cudaDeviceSynchronize());
  if (num_unique > cap) abort();
  SHIM_CUDA(cudaMemcpy(unique_out, d_unique, (size_t)num_unique * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_in);
  cudaFree(d_unique);
  return num_unique;
}

void refcu_table_map_edges(void *h, const uint32_t *src, const uint32_t *dst, size_t n, uint32_t *new_src,
                           uint32_t *new_dst) {
  auto *r = static_cast<RefTable *>(h);
  SHIM_CUDA(cudaSetDevice(r->device));
  uint32_t *d_src = to_device(src, n), *d_dst = to_device(dst, n), *d_ns = nullptr, *d_nd = nullptr;
  SHIM_CUDA(cudaMalloc(&d_ns, (n ? n : 1) * 4));
  SHIM_CUDA(cudaMalloc(&d_nd, (n ? n : 1) * 4));
  cuda::GPUMapEdges(d_src, d_ns, d_dst, d_nd, n, r->t->DeviceHandle(), GPU(r->device), nullptr);
  SHIM_CUDA(cudaDeviceSynchronize());
  SHIM_CUDA(cudaMemcpy(new_src, d_ns, n * 4, cudaMemcpyDeviceToHost));
  SHIM_CUDA(cudaMemcpy(new_dst, d_nd, n * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_src);
  cudaFree(d_dst);
  cudaFree(d_ns);
  cudaFree(d_nd);
}


// rows (float32, dim columns) of `ranks` through the reference's partitioned-feature accessors.
//   compact_bitwidth == 0 and device_map == NULL: DeviceDistFeature over num_part parts
//   compact_bitwidth != 0: DeviceICS22SongDistFeature, compact ids;  device_map != NULL: its two-map form
// parts_concat / part_off (in ELEMENTS): part p = parts_concat[part_off[p] .. part_off[p + 1])
void refcu_dist_feature_get(int device, const float *parts_concat, const size_t *part_off, uint32_t num_part,
                            uint32_t num_cache_node, uint32_t dim, uint32_t compact_bitwidth,
                            const uint32_t *device_map, const uint32_t *new_idx_map, uint32_t num_map,
                            const uint32_t *ranks, size_t n, float *out) {
  SHIM_CUDA(cudaSetDevice(device));
  std::vector<void *> parts(num_part);
  for (uint32_t p = 0; p < num_part; ++p) parts[p] = to_device(parts_concat + part_off[p], part_off[p + 1] - part_off[p]);
  void **d_parts = to_device(parts.data(), parts.size());
  IdType *d_ranks = to_device(ranks, n), *d_dev = nullptr, *d_idx = nullptr;
  float *d_out = nullptr;
  SHIM_CUDA(cudaMalloc(&d_out, (n * dim ? n * dim : 1) * sizeof(float)));
  if (n) {
    if (compact_bitwidth == 0 && device_map == nullptr) {
      cuda::DeviceDistFeature feat(d_parts, num_part, num_cache_node, dim);
      shim_dist_feature_rows<float><<<296, 256>>>(feat, d_ranks, n, dim, d_out);
    } else {
      if (device_map) {
        d_dev = to_device(device_map, num_map);
        d_idx = to_device(new_idx_map, num_map);
      }
      cuda::DeviceICS22SongDistFeature feat(d_parts, d_dev, d_idx, num_cache_node, dim, compact_bitwidth);
      shim_dist_feature_rows<float><<<296, 256>>>(feat, d_ranks, n, dim, d_out);
    }
    SHIM_CUDA(cudaGetLastError());
  }
  SHIM_CUDA(cudaDeviceSynchronize());
  SHIM_CUDA(cudaMemcpy(out, d_out, n * dim * sizeof(float), cudaMemcpyDeviceToHost));
  for (void *p : parts) cudaFree(p);
  cudaFree(d_parts);
  cudaFree(d_ranks);
  cudaFree(d_dev);
  cudaFree(d_idx);
  cudaFree(d_out);
}

}  // extern "C"
