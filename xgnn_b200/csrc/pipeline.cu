// Minibatch engine (SURVEY.md rows a10 + the extract stage functions).
//
// Reference behaviour being replaced: dist::DoGPUSample (samgraph/common/dist/dist_loops.cc:62-368),
// DoGPULabelExtract (:938-974), DoArch6GetCacheMissIndex (:1015-1074), DoArch6GPUCacheFeatureCopy
// (:1209-1285).  There one Task costs ~35-40 host<->device control crossings per layer (launch +
// StreamSync each), >= 8 workspace alloc/free per layer and a blocking D2H of every count.
//
// Here a Task is ONE stream-ordered chain with every count resident on the device:
//   Reset (targeted clear) -> seed insert -> number -> per layer { sample+insert (fused) -> number ->
//   local-id gather } -> input_nodes copy -> routed feature gather -> label gather -> 1 D2H of counts
// All buffers are sized once for the worst-case batch (PredictNumNodes, common.cc:488-497) so the
// chain has a fixed shape and is captured once per batch slot as a CUDA graph and replayed: one
// launch call per minibatch, no host synchronisation until the consumer asks for the batch.
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "extract.cuh"
#include "hashtable.cuh"
#include "sampling.cuh"

namespace xg {
TableView table_view(xg_hashtable *t);
}
using namespace xg;

namespace {

struct BatchParams {  // written by the host, read by the kernels of one batch
  uint64_t num_seeds;
  uint32_t num_seeds32;
  uint32_t pad;
  uint2 rng_key;
};

struct BatchCounts {  // written by the kernels, read back once
  uint64_t num_edge[XG_MAX_LAYERS];
  uint32_t items[XG_MAX_LAYERS + 2];  // items[0] = 0, [1] = unique seeds, [2 + k] = after k-th expanded layer
  uint32_t overflow;
  uint32_t pad;
};

__global__ void collect_counts_kernel(TableView t, int num_fill, BatchCounts *out) {
  const int i = threadIdx.x;
  if (i <= num_fill) out->items[i] = t.items[i];
  if (i == 0) out->overflow = *t.overflow;
}

size_t elem_bytes(int dtype) {
  switch (dtype) {
    case XG_F32: case XG_I32: return 4;
    case XG_F64: case XG_I64: return 8;
    case XG_F16: return 2;
    case XG_U8: case XG_I8: return 1;
    default: return 0;
  }
}

}  // namespace

struct xg_batch {
  xg_sampler *owner = nullptr;
  int index = 0;
  int lane = 0;  // which sample lane (hash table + scratch + stream) this slot's graph was captured on
  std::atomic<int> claimed{0};
  // device
  uint32_t *seeds = nullptr, *seed_local = nullptr, *input_nodes = nullptr;
  uint32_t *row[XG_MAX_LAYERS] = {}, *col[XG_MAX_LAYERS] = {}, *data[XG_MAX_LAYERS] = {};
  uint32_t *csc_indptr[XG_MAX_LAYERS] = {};  // uniform samplers only: [cap_in[l] + 1]
  void *feat = nullptr;
  int64_t *label = nullptr;
  BatchParams *d_params = nullptr;
  BatchCounts *d_counts = nullptr;
  // pinned host
  BatchParams *h_params = nullptr;
  BatchCounts *h_counts = nullptr;
  uint32_t *h_seeds = nullptr;
  cudaEvent_t done = nullptr;
  cudaEvent_t tl[3] = {};                 // XG_TIMELINE=1: sample begin / extract begin / extract end (timed)
  cudaEvent_t sampled = nullptr;          // the sample phase of this batch has been enqueued up to here
  cudaGraphExec_t graph = nullptr;        // sample phase (whole Task when the sampler runs on one stream)
  cudaGraphExec_t graph_x = nullptr;      // extract phase (two-stream mode)
  uint64_t key = 0;
  uint32_t num_seeds = 0;
  bool recorded = false;  // `done` has been recorded at least once
  std::vector<cudaEvent_t> marks;  // stage boundaries (XG_SAMPLER_STAGE_TIMING only)
  int num_marks = 0;
};

struct xg_sampler {
  int device = 0;
  xg_graph graph{};
  xg_feature_store store{};
  bool has_store = false;
  const void *label = nullptr;
  xg_sampler_config cfg{};
  size_t cap_in[XG_MAX_LAYERS] = {};    // most input nodes of layer l
  size_t cap_edge[XG_MAX_LAYERS] = {};  // most sampled edges of layer l
  size_t cap_input = 0;                 // most input_nodes of a batch
  size_t row_bytes = 0;
  // A sample LANE = one hash table + per-layer scratch + one stream.  Batches of one lane are strictly ordered;
  // a second lane (XG_SAMPLE_LANES=2) lets the sample phases of two consecutive batches run side by side.
  struct Lane {
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    xg_hashtable *table = nullptr;
    TableView view{};
    uint32_t *dst_global = nullptr, *pos = nullptr;  // per-layer scratch, reused layer after layer
    void *scan_mem = nullptr;
    ExtWorkspace ext;  // scratch of the with-replacement / weighted samplers
    std::atomic<bool> dirty{false};  // a batch overflowed: unnumbered keys may be left in the table
  };
  Lane lanes[2];
  int num_lanes = 1;
  int next_lane = 0;
  size_t scan_bytes = 0, scan_tiles_a = 0, scan_tiles_b = 0;
  cudaStream_t stream = nullptr;    // == lanes[0].stream (sample phase of lane 0)
  cudaStream_t xstream = nullptr;   // extract phase: touches only the batch slot's own buffers (== stream in 1-stream mode)
  bool two_streams = false;
  bool timeline = false;            // XG_TIMELINE=1: timed events around both phases of every batch
  cudaEvent_t origin = nullptr;
  bool use_graph = true;
  bool stage_timing = false;
  bool uniform = true;      // KHop0/2/3: fused sample + insert kernel
  WeightTables weights;
  uint32_t launches_per_batch = 0;
  uint32_t launches_phase[2] = {0, 0};
  std::mutex mu;
  std::vector<xg_batch *> slots;
};

namespace {

ScanSpace scan_a(xg_sampler *s, xg_sampler::Lane &ln) {
  auto *w = reinterpret_cast<uint64_t *>(ln.scan_mem);
  return ScanSpace{reinterpret_cast<uint32_t *>(w), w + 1, s->scan_tiles_a};
}
ScanSpace scan_b(xg_sampler *s, xg_sampler::Lane &ln) {
  auto *w = reinterpret_cast<uint64_t *>(ln.scan_mem) + 1 + s->scan_tiles_a;
  return ScanSpace{reinterpret_cast<uint32_t *>(w), w + 1, s->scan_tiles_b};
}

int alloc_slot(xg_sampler *s, xg_batch **out) {
  auto *b = new xg_batch();
  b->owner = s;
  const uint32_t L = s->cfg.num_layer;
  const size_t B = s->cfg.batch_size;
  XG_CUDA(cudaMalloc(&b->seeds, B * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&b->seed_local, B * sizeof(uint32_t)));
  XG_CUDA(cudaMalloc(&b->input_nodes, s->cap_input * sizeof(uint32_t)));
  for (uint32_t l = 0; l < L; ++l) {
    XG_CUDA(cudaMalloc(&b->row[l], s->cap_edge[l] * sizeof(uint32_t)));
    XG_CUDA(cudaMalloc(&b->col[l], s->cap_edge[l] * sizeof(uint32_t)));
    if (s->cfg.sample_type == XG_RANDOM_WALK) XG_CUDA(cudaMalloc(&b->data[l], s->cap_edge[l] * sizeof(uint32_t)));
    if (s->uniform) {
      XG_CUDA(cudaMalloc(&b->csc_indptr[l], (s->cap_in[l] + 1) * sizeof(uint32_t)));
      XG_CUDA(cudaMemset(b->csc_indptr[l], 0, sizeof(uint32_t)));  // an empty layer still has indptr[0] = 0
    }
  }
  if (s->has_store) XG_CUDA(cudaMalloc(&b->feat, s->cap_input * s->row_bytes));
  if (s->label) XG_CUDA(cudaMalloc(&b->label, B * sizeof(int64_t)));
  XG_CUDA(cudaMalloc(&b->d_params, sizeof(BatchParams)));
  XG_CUDA(cudaMalloc(&b->d_counts, sizeof(BatchCounts)));
  XG_CUDA(cudaMemset(b->d_counts, 0, sizeof(BatchCounts)));
  XG_CUDA(cudaHostAlloc(&b->h_params, sizeof(BatchParams), cudaHostAllocDefault));
  XG_CUDA(cudaHostAlloc(&b->h_counts, sizeof(BatchCounts), cudaHostAllocDefault));
  XG_CUDA(cudaHostAlloc(&b->h_seeds, B * sizeof(uint32_t), cudaHostAllocDefault));
  XG_CUDA(cudaEventCreateWithFlags(&b->done, cudaEventDisableTiming));
  XG_CUDA(cudaEventCreateWithFlags(&b->sampled, s->timeline ? cudaEventDefault : cudaEventDisableTiming));
  if (s->timeline)
    for (auto &e : b->tl) XG_CUDA(cudaEventCreate(&e));
  if (s->stage_timing) {
    b->marks.resize(5 + 3 * L);
    for (auto &e : b->marks) XG_CUDA(cudaEventCreate(&e));
  }
  *out = b;
  return XG_OK;
}

void free_slot(xg_batch *b) {
  if (!b) return;
  if (b->graph) cudaGraphExecDestroy(b->graph);
  if (b->graph_x) cudaGraphExecDestroy(b->graph_x);
  if (b->done) cudaEventDestroy(b->done);
  if (b->sampled) cudaEventDestroy(b->sampled);
  for (auto e : b->tl)
    if (e) cudaEventDestroy(e);
  for (auto e : b->marks) cudaEventDestroy(e);
  cudaFree(b->seeds);
  cudaFree(b->seed_local);
  cudaFree(b->input_nodes);
  for (int l = 0; l < XG_MAX_LAYERS; ++l) {
    cudaFree(b->row[l]);
    cudaFree(b->col[l]);
    cudaFree(b->data[l]);
    cudaFree(b->csc_indptr[l]);
  }
  cudaFree(b->feat);
  cudaFree(b->label);
  cudaFree(b->d_params);
  cudaFree(b->d_counts);
  cudaFreeHost(b->h_params);
  cudaFreeHost(b->h_counts);
  cudaFreeHost(b->h_seeds);
  delete b;
}

// The Task in two stream-ordered phases; no host synchronisation, no allocation (capturable).
//   sample phase  (lane stream): Reset .. collect_counts: uses the lane's hash table and scratch
//   extract phase (s->xstream): feature + label gather and the one D2H: reads only slot-owned buffers
// so the extract phase of batch i (HBM/NVLink-bandwidth bound) runs while batch i+1 is being sampled
// (latency bound, ~20 % of DRAM bandwidth).
int enqueue_sample(xg_sampler *s, xg_batch *b) {
  xg_sampler::Lane &ln = s->lanes[b->lane];
  cudaStream_t st = ln.stream;
  const int L = (int)s->cfg.num_layer;
  const size_t B = s->cfg.batch_size;
  // the table's unique list (n2o) IS this slot's input_nodes buffer: the numbering kernels append straight into
  // it, so task->input_nodes needs no copy and the next batch (another slot) cannot overwrite it while the
  // extract phase of this one is still reading
  TableView t = ln.view;
  t.n2o = b->input_nodes;
  t.n2o_cap = (uint32_t)s->cap_input;
  const uint64_t *d_nseed64 = &b->d_params->num_seeds;
  const uint32_t *d_nseed32 = &b->d_params->num_seeds32;

  b->num_marks = 0;
  auto mark = [&]() {
    if (s->stage_timing && b->num_marks < (int)b->marks.size()) cudaEventRecord(b->marks[b->num_marks++], st);
  };
  mark();
  // hash_table->Reset (dist_loops.cc:101): free the slots the previous batch numbered
  int rc = launch_table_clear(t, L + 1, s->cap_input, st);
  if (rc) return rc;
  // FillWithDupRevised(output_nodes) (dist_loops.cc:106-107)
  XG_CUDA(cudaMemsetAsync(ln.scan_mem, 0, s->scan_bytes, st));
  if ((rc = launch_table_insert(t, b->seeds, ln.pos, B, d_nseed64, st))) return rc;
  if ((rc = launch_table_assign(t, b->seeds, ln.pos, B, d_nseed64, 1, scan_b(s, ln), st, b->seed_local))) return rc;
  mark();

  for (int l = L - 1; l >= 0; --l) {  // seeds outward (dist_loops.cc:120-123)
    const int fill = (L - 1 - l) + 2;
    const bool seed_layer = (l == L - 1);
    XG_CUDA(cudaMemsetAsync(ln.scan_mem, 0, s->scan_bytes, st));
    uint64_t *d_num_edge = &b->d_counts->num_edge[l];
    const uint32_t *in_nodes = seed_layer ? b->seeds : t.n2o;
    const uint32_t *d_num_in = seed_layer ? d_nseed32 : &t.items[fill - 1];
    // later inputs ARE the unique list: local id == position
    const uint32_t *src_local = seed_layer ? b->seed_local : nullptr;
    const uint32_t salt = kSampleSalt + (uint32_t)l;
    if (s->uniform) {
      SampleFusion fusion{};
      fusion.src_local = src_local;
      fusion.csc_indptr = b->csc_indptr[l];
      fusion.enabled = true;
      rc = launch_sample_uniform(s->graph, in_nodes, s->cap_in[l], d_num_in, s->cfg.fanout[l], b->col[l], ln.dst_global,
                                 d_num_edge, 0, 0, &b->d_params->rng_key, salt, fusion, scan_a(s, ln), st);
      if (rc) return rc;
    } else {
      if (s->cfg.sample_type == XG_RANDOM_WALK) {
        rc = launch_random_walk(s->graph, in_nodes, s->cap_in[l], d_num_in, s->cfg.random_walk_length,
                                s->cfg.random_walk_restart_prob, s->cfg.num_random_walk, s->cfg.fanout[l], true,
                                src_local, b->col[l], ln.dst_global, b->data[l], d_num_edge, make_uint2(0, 0),
                                &b->d_params->rng_key, salt, scan_a(s, ln), st);
      } else {
        rc = launch_sample_ext(s->cfg.sample_type, s->graph, s->weights, in_nodes, s->cap_in[l], d_num_in,
                               s->cfg.fanout[l], true, src_local, b->col[l], ln.dst_global, d_num_edge,
                               make_uint2(0, 0), &b->d_params->rng_key, salt, ln.ext, st);
      }
      if (rc) return rc;
    }
    mark();
    // the neighbours go into the table in a pass of their own (see SampleFusion): coalesced keys, one L2 atomic each
    if ((rc = launch_table_insert(t, ln.dst_global, ln.pos, s->cap_edge[l], d_num_edge, st))) return rc;
    mark();
    // FillWithDuplicates(out_dst) + GPUMapEdges (dist_loops.cc:279-297): number the new ids and, in the same pass,
    // write the neighbour's local id for every edge (one 4-byte load per edge at the remembered slot)
    if ((rc = launch_table_assign(t, ln.dst_global, ln.pos, s->cap_edge[l], d_num_edge, fill, scan_b(s, ln), st, b->row[l],
                                  /*ids_into_table=*/l != 0)))  // layer 0 is the batch's last fill: nothing looks its ids up
      return rc;
    mark();
  }

  // task->input_nodes = the unique list (dist_loops.cc:357), already in place; counts leave the table before the next Reset
  collect_counts_kernel<<<1, 32, 0, st>>>(t, L + 1, b->d_counts);
  XG_LAUNCHED();
  mark();
  return XG_OK;
}

int enqueue_extract(xg_sampler *s, xg_batch *b) {
  cudaStream_t st = s->xstream;
  const int L = (int)s->cfg.num_layer;
  const size_t B = s->cfg.batch_size;
  const uint32_t *d_nseed32 = &b->d_params->num_seeds32;
  int rc = XG_OK;
  auto mark = [&]() {
    if (s->stage_timing && b->num_marks < (int)b->marks.size()) cudaEventRecord(b->marks[b->num_marks++], st);
  };
  if (s->has_store) {  // DoArch6GetCacheMissIndex + DoArch6GPUCacheFeatureCopy in one launch
    rc = launch_gather(s->store, kByNode, b->input_nodes, nullptr, s->cap_input, &b->d_counts->items[L + 1],
                       b->feat, nullptr, st);
    if (rc) return rc;
  }
  mark();
  if (s->label) {  // DoGPULabelExtract: dim = 1, int64, straight from wherever the labels live
    xg_feature_store ls{};
    ls.host = s->label;
    ls.row_bytes = sizeof(int64_t);
    if ((rc = launch_gather(ls, kHostRow, b->seeds, nullptr, B, d_nseed32, b->label, nullptr, st))) return rc;
  }
  mark();
  XG_CUDA(cudaMemcpyAsync(b->h_counts, b->d_counts, sizeof(BatchCounts), cudaMemcpyDeviceToHost, st));
  return XG_OK;
}

struct DeviceScope {
  int prev = 0, target = 0;
  explicit DeviceScope(int dev) : target(dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceScope() {
    if (prev != target) cudaSetDevice(prev);
  }
};

}  // namespace

extern "C" {

size_t xg_predict_num_nodes(size_t batch_size, const uint32_t *fanout, size_t num_layer) {
  size_t count = batch_size;
  for (size_t i = num_layer; i-- > 0;) count += count * fanout[i];
  return count;
}

int xg_sampler_create(xg_sampler **out, const xg_graph *graph, const xg_feature_store *store,
                      const void *label, const xg_sampler_config *cfg, int device) {
  XG_REQUIRE(out && graph && cfg, "NULL argument");
  XG_REQUIRE(cfg->num_layer >= 1 && cfg->num_layer <= XG_MAX_LAYERS, "num_layer must be in [1, 8]");
  XG_REQUIRE(cfg->batch_size >= 1, "batch_size is 0");
  XG_REQUIRE(cfg->sample_type >= XG_KHOP0 && cfg->sample_type <= XG_KHOP3, "unknown sample_type");
  const bool uniform = cfg->sample_type == XG_KHOP0 || cfg->sample_type == XG_KHOP2 || cfg->sample_type == XG_KHOP3;
  if (cfg->sample_type == XG_RANDOM_WALK) {
    XG_REQUIRE(cfg->random_walk_length >= 1 && cfg->num_random_walk >= 1, "random walk needs length and walks >= 1");
    XG_REQUIRE((uint64_t)cfg->random_walk_length * cfg->num_random_walk <= 512, "walk_length * num_walk must be <= 512");
  }
  if (cfg->sample_type == XG_WEIGHTED_KHOP || cfg->sample_type == XG_WEIGHTED_KHOP_HASH_DEDUP)
    XG_REQUIRE(cfg->prob_table && cfg->alias_table, "weighted KHop needs prob_table and alias_table");
  if (cfg->sample_type == XG_WEIGHTED_KHOP_PREFIX) XG_REQUIRE(cfg->prob_prefix_table, "prob_prefix_table is NULL");
  if (!uniform && cfg->sample_type != XG_KHOP1 && cfg->sample_type != XG_RANDOM_WALK)
    XG_REQUIRE(graph->num_part == 0, "weighted samplers take the flat CSR their tables are aligned with");
  XG_REQUIRE(graph->num_part <= XG_MAX_PARTS, "too many partitions");
  DeviceScope scope(device);
  auto *s = new xg_sampler();
  s->device = device;
  s->graph = *graph;
  s->cfg = *cfg;
  s->label = label;
  s->uniform = uniform;
  s->weights.prob = cfg->prob_table;
  s->weights.alias = cfg->alias_table;
  s->weights.prefix = cfg->prob_prefix_table;
  if (store) {
    s->store = *store;
    s->has_store = true;
    s->row_bytes = store->row_bytes;
    const size_t eb = elem_bytes(cfg->feat_dtype);
    if (eb == 0 || eb * cfg->feat_dim != store->row_bytes) {
      delete s;
      XG_REQUIRE(false, "feat_dim * sizeof(feat_dtype) must equal the store's row_bytes");
    }
  }
  const int L = (int)cfg->num_layer;
  size_t limit = cfg->max_input_nodes ? cfg->max_input_nodes
                                      : xg_predict_num_nodes(cfg->batch_size, cfg->fanout, cfg->num_layer);
  if (graph->num_node && limit > (size_t)graph->num_node + cfg->batch_size)
    limit = (size_t)graph->num_node + cfg->batch_size;
  size_t cur = cfg->batch_size;
  for (int l = L - 1; l >= 0; --l) {
    if (cfg->fanout[l] < 1 || cfg->fanout[l] > 512) {
      delete s;
      XG_REQUIRE(false, "fanout must be in [1, 512]");
    }
    s->cap_in[l] = cur;
    s->cap_edge[l] = cur * cfg->fanout[l];
    if (s->cap_edge[l] >= (1ull << 31)) {
      delete s;
      XG_REQUIRE(false, "a layer may hold at most 2^31 - 1 sampled edges");
    }
    cur = cur + s->cap_edge[l];
    if (cur > limit) cur = limit;
  }
  s->cap_input = cur;
  size_t max_edge = cfg->batch_size, tiles_a = 1, tiles_b = (cfg->batch_size + kAssignTile - 1) / kAssignTile;
  for (int l = 0; l < L; ++l) {
    if (s->cap_edge[l] > max_edge) max_edge = s->cap_edge[l];
    // random walk scans over 64-seed tiles, the k-hop kernels over kSampleTile nodes
    const size_t ta = (s->cap_in[l] + 63) / 64;
    const size_t tb = (s->cap_edge[l] + kAssignTile - 1) / kAssignTile;
    if (ta > tiles_a) tiles_a = ta;
    if (tb > tiles_b) tiles_b = tb;
  }
  s->scan_tiles_a = tiles_a + 1;
  s->scan_tiles_b = tiles_b + 1;
  s->scan_bytes = (2 + s->scan_tiles_a + s->scan_tiles_b) * sizeof(uint64_t);

  const char *env = std::getenv("XG_NO_GRAPH");
  s->stage_timing = (cfg->flags & XG_SAMPLER_STAGE_TIMING) != 0;
  s->use_graph = !(cfg->flags & (XG_SAMPLER_NO_GRAPH | XG_SAMPLER_STAGE_TIMING)) && !(env && env[0] == '1');
  const char *one = std::getenv("XG_SINGLE_STREAM");
  // XG_STAGE_OVERLAP=1 (measurement knob, tools/stage_overlap.py): stage marks with BOTH streams live, so that the marks
  // show how long each sample-phase kernel takes while the previous batch's gather runs beside it
  const char *ov = std::getenv("XG_STAGE_OVERLAP");
  const bool overlap_marks = s->stage_timing && ov && ov[0] == '1';
  s->two_streams = (!s->stage_timing || overlap_marks) && !(cfg->flags & XG_SAMPLER_SINGLE_STREAM) && !(one && one[0] == '1');
  // one lane by default: measured on B200 (profiles/r01h_sweep_lanes.txt) a second lane does not help - with the
  // extract phase already running beside it the memory system is the shared limit (1.30 ms/minibatch with one lane,
  // 1.34 with two); XG_SAMPLE_LANES=2 keeps the experiment reproducible
  s->num_lanes = 1;
  if (const char *nl = std::getenv("XG_SAMPLE_LANES")) s->num_lanes = (s->two_streams && std::atoi(nl) >= 2) ? 2 : 1;
  if (const char *fg = std::getenv("XG_L2_FETCH")) {  // experiment knob: 32 / 64 / 128-byte L2 fetch granularity
    cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)std::atoi(fg));
  }
  const char *tl = std::getenv("XG_TIMELINE");
  s->timeline = tl && tl[0] == '1';
  int prio_low = 0, prio_high = 0;
  XG_CUDA(cudaDeviceGetStreamPriorityRange(&prio_low, &prio_high));
  int rc = XG_OK;
  for (int k = 0; k < s->num_lanes; ++k) {
    xg_sampler::Lane &ln = s->lanes[k];
    if ((rc = xg_hashtable_create(&ln.table, s->cap_input, device))) {
      xg_sampler_destroy(s);
      return rc;
    }
    ln.view = table_view(ln.table);
    XG_CUDA(cudaMalloc(&ln.dst_global, max_edge * sizeof(uint32_t)));
    XG_CUDA(cudaMalloc(&ln.pos, max_edge * sizeof(uint32_t)));
    XG_CUDA(cudaMalloc(&ln.scan_mem, s->scan_bytes));
    if (!uniform && cfg->sample_type != XG_RANDOM_WALK) {
      size_t max_in = 1;
      for (int l = 0; l < L; ++l) max_in = std::max(max_in, s->cap_in[l]);
      // cap_slots must cover the largest layer: reserve by the largest product
      size_t max_slots = 1;
      for (int l = 0; l < L; ++l) max_slots = std::max(max_slots, s->cap_edge[l]);
      if ((rc = ext_reserve(ln.ext, max_in, (max_slots + max_in - 1) / max_in, nullptr))) {
        xg_sampler_destroy(s);
        return rc;
      }
    }
    if (k == 0 && cfg->stream) {
      ln.stream = as_stream(cfg->stream);
    } else {
      // the sample phase is a chain of short latency-bound kernels: let its CTAs go first
      XG_CUDA(cudaStreamCreateWithPriority(&ln.stream, cudaStreamNonBlocking, prio_high));
      ln.own_stream = true;
    }
  }
  s->stream = s->lanes[0].stream;
  const char *xp = std::getenv("XG_EXTRACT_PRIO");  // experiment knob: "high" = extract phase first
  if (s->two_streams)
    XG_CUDA(cudaStreamCreateWithPriority(&s->xstream, cudaStreamNonBlocking, (xp && xp[0] == 'h') ? prio_high : prio_low));
  else s->xstream = s->stream;
  if (s->timeline) {
    XG_CUDA(cudaEventCreate(&s->origin));
    XG_CUDA(cudaEventRecord(s->origin, s->stream));
  }
  const uint32_t nslots = cfg->num_slots ? cfg->num_slots : 2;
  for (uint32_t i = 0; i < nslots; ++i) {
    xg_batch *b = nullptr;
    if ((rc = alloc_slot(s, &b))) {
      xg_sampler_destroy(s);
      return rc;
    }
    b->index = (int)i;
    b->lane = (int)(i % (uint32_t)s->num_lanes);
    s->slots.push_back(b);
  }
  *out = s;
  return XG_OK;
}

int xg_sampler_destroy(xg_sampler *s) {
  if (!s) return XG_OK;
  DeviceScope scope(s->device);
  for (int k = 0; k < 2; ++k)
    if (s->lanes[k].stream) cudaStreamSynchronize(s->lanes[k].stream);
  if (s->two_streams && s->xstream) cudaStreamSynchronize(s->xstream);
  for (auto *b : s->slots) free_slot(b);
  for (int k = 0; k < 2; ++k) {
    xg_sampler::Lane &ln = s->lanes[k];
    cudaFree(ln.dst_global);
    cudaFree(ln.pos);
    cudaFree(ln.scan_mem);
    ext_release(ln.ext);
    if (ln.table) xg_hashtable_destroy(ln.table);
    if (ln.own_stream && ln.stream) cudaStreamDestroy(ln.stream);
  }
  if (s->origin) cudaEventDestroy(s->origin);
  if (s->two_streams && s->xstream) cudaStreamDestroy(s->xstream);
  delete s;
  return XG_OK;
}

int xg_sampler_submit(xg_sampler *s, const uint32_t *seeds, size_t num_seeds, int seeds_on_device,
                      uint64_t batch_key, xg_batch **out) {
  XG_REQUIRE(s && out, "NULL argument");
  XG_REQUIRE(seeds != nullptr || num_seeds == 0, "seeds is NULL");
  XG_REQUIRE(num_seeds <= s->cfg.batch_size, "more seeds than batch_size");
  DeviceScope scope(s->device);
  std::lock_guard<std::mutex> lk(s->mu);
  xg_batch *b = nullptr;
  // alternate the sample lanes: first look for a free slot on the lane whose turn it is, then on any lane
  for (int pass = 0; pass < 2 && !b; ++pass) {
    for (auto *c : s->slots) {
      if (pass == 0 && c->lane != s->next_lane) continue;
      int expected = 0;
      if (c->claimed.compare_exchange_strong(expected, 1)) {
        b = c;
        break;
      }
    }
  }
  if (!b) {
    set_error("every batch slot is still held: release a batch (xg_batch_release) or raise num_slots");
    return XG_ERR_CAPACITY;
  }
  s->next_lane = (b->lane + 1) % s->num_lanes;
  xg_sampler::Lane &ln = s->lanes[b->lane];
  cudaStream_t st = ln.stream;
  // the slot's pinned staging is about to be rewritten: its previous batch must have drained
  if (b->recorded && cudaEventSynchronize(b->done) != cudaSuccess) {
    b->claimed.store(0);
    set_error("cudaEventSynchronize failed on a recycled batch slot");
    return XG_ERR_CUDA;
  }
  if (ln.dirty.exchange(false)) {  // after an overflow the targeted Reset cannot know every used slot
    const size_t cap = (size_t)ln.view.cap;
    if (cudaMemsetAsync(ln.view.slots, 0xFF, cap * sizeof(uint32_t), st) != cudaSuccess ||
        cudaMemsetAsync(ln.view.items, 0, (kMaxFills + 2) * sizeof(uint32_t), st) != cudaSuccess) {
      b->claimed.store(0);
      return XG_ERR_CUDA;
    }
  }
  b->key = batch_key;
  b->num_seeds = (uint32_t)num_seeds;
  b->h_params->num_seeds = num_seeds;
  b->h_params->num_seeds32 = (uint32_t)num_seeds;
  b->h_params->pad = 0;
  b->h_params->rng_key = philox_key(s->cfg.seed, batch_key);
  int rc = XG_OK;
  auto fail = [&](int code) {
    b->claimed.store(0);
    return code;
  };
  if (num_seeds) {
    if (seeds_on_device) {
      if (cudaMemcpyAsync(b->seeds, seeds, num_seeds * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st) != cudaSuccess)
        return fail(XG_ERR_CUDA);
    } else {
      std::memcpy(b->h_seeds, seeds, num_seeds * sizeof(uint32_t));
      if (cudaMemcpyAsync(b->seeds, b->h_seeds, num_seeds * sizeof(uint32_t), cudaMemcpyHostToDevice, st) != cudaSuccess)
        return fail(XG_ERR_CUDA);
    }
  }
  if (cudaMemcpyAsync(b->d_params, b->h_params, sizeof(BatchParams), cudaMemcpyHostToDevice, st) != cudaSuccess)
    return fail(XG_ERR_CUDA);

  cudaStream_t xs = s->xstream;
  // one phase: either replay its graph (captured on first use of the slot) or enqueue it directly
  auto run_phase = [&](cudaStream_t on, cudaGraphExec_t *exec, int (*enqueue)(xg_sampler *, xg_batch *),
                       uint32_t *count) -> int {
    if (!s->use_graph) return enqueue(s, b);
    if (!*exec) {
      const uint64_t before = g_launches.load();
      cudaGraph_t g = nullptr;
      if (cudaStreamBeginCapture(on, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        set_error("cudaStreamBeginCapture failed");
        return XG_ERR_CUDA;
      }
      int r = enqueue(s, b);
      cudaError_t e = cudaStreamEndCapture(on, &g);
      *count = (uint32_t)(g_launches.load() - before);  // kernels recorded into the graph, not launched yet
      g_launches.store(before);
      if (r || e != cudaSuccess || !g) {
        if (!r) set_error(std::string("graph capture failed: ") + cudaGetErrorString(e));
        if (g) cudaGraphDestroy(g);
        return r ? r : XG_ERR_CUDA;
      }
      e = cudaGraphInstantiate(exec, g, 0);
      cudaGraphDestroy(g);
      if (e != cudaSuccess) {
        set_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
        return XG_ERR_CUDA;
      }
    }
    if (cudaGraphLaunch(*exec, on) != cudaSuccess) {
      set_error(std::string("cudaGraphLaunch: ") + cudaGetErrorString(cudaGetLastError()));
      return XG_ERR_CUDA;
    }
    return XG_OK;
  };
  const uint64_t launches_before = g_launches.load();
  if (s->timeline) cudaEventRecord(b->tl[0], st);
  if ((rc = run_phase(st, &b->graph, enqueue_sample, &s->launches_phase[0]))) return fail(rc);
  if (s->two_streams) {  // extract phase of this batch waits for its own sample phase only
    if (cudaEventRecord(b->sampled, st) != cudaSuccess || cudaStreamWaitEvent(xs, b->sampled, 0) != cudaSuccess)
      return fail(XG_ERR_CUDA);
  }
  if (s->timeline) cudaEventRecord(b->tl[1], xs);
  if ((rc = run_phase(xs, &b->graph_x, enqueue_extract, &s->launches_phase[1]))) return fail(rc);
  if (s->use_graph) {
    s->launches_per_batch = s->launches_phase[0] + s->launches_phase[1];
    g_launches.store(launches_before + s->launches_per_batch);
  } else {
    s->launches_per_batch = (uint32_t)(g_launches.load() - launches_before);
  }
  st = xs;
  if (s->timeline) cudaEventRecord(b->tl[2], st);
  if (cudaEventRecord(b->done, st) != cudaSuccess) return fail(XG_ERR_CUDA);
  b->recorded = true;
  *out = b;
  return XG_OK;
}

int xg_batch_wait(xg_batch *b, xg_batch_view *view) {
  XG_REQUIRE(b && b->owner, "batch is NULL");
  XG_REQUIRE(b->claimed.load() == 1, "batch is not in flight");
  xg_sampler *s = b->owner;
  DeviceScope scope(s->device);
  XG_CUDA(cudaEventSynchronize(b->done));
  const BatchCounts &c = *b->h_counts;
  const int L = (int)s->cfg.num_layer;
  bool over = c.overflow != 0;
  for (int l = L - 1; l >= 0 && !over; --l) {
    const int fill = (L - 1 - l) + 2;
    if (l != L - 1 && c.items[fill - 1] > s->cap_in[l]) over = true;
  }
  if (c.items[L + 1] > s->cap_input) over = true;
  if (over) {
    s->lanes[b->lane].dirty.store(true);
    set_error("minibatch outgrew its buffers (max_input_nodes=" + std::to_string(s->cap_input) + ")");
    return XG_ERR_CAPACITY;
  }
  if (!view) return XG_OK;
  std::memset(view, 0, sizeof(*view));
  view->key = b->key;
  view->num_layer = (uint32_t)L;
  view->num_output = b->num_seeds;
  view->num_input = c.items[L + 1];
  for (int l = L - 1; l >= 0; --l) {
    const int fill = (L - 1 - l) + 2;
    view->row[l] = b->row[l];
    view->col[l] = b->col[l];
    view->data[l] = b->data[l];
    // duplicate seeds (the aligned shuffler's padding, dist_shuffler_aligned.cc:53-55) share one local id, so the
    // seed layer's edges are then not destination-major: no CSC form for that layer of that batch
    const bool dup_seeds = (l == L - 1) && c.items[1] != b->num_seeds;
    view->csc_indptr[l] = dup_seeds ? nullptr : b->csc_indptr[l];
    view->num_dst[l] = (l == L - 1) ? b->num_seeds : c.items[fill - 1];  // num_dst = num_input (dist_loops.cc:304-306)
    view->num_src[l] = c.items[fill];
    view->num_edge[l] = c.num_edge[l];
  }
  view->input_nodes = b->input_nodes;
  view->output_nodes = b->seeds;
  view->input_feat = s->has_store ? b->feat : nullptr;
  view->output_label = s->label ? b->label : nullptr;
  return XG_OK;
}

int xg_batch_release(xg_batch *b) {
  if (!b) return XG_OK;
  b->claimed.store(0);
  return XG_OK;
}

int xg_batch_stage_ms(xg_batch *b, int stage, int layer, float *ms) {
  XG_REQUIRE(b && b->owner && ms, "NULL argument");
  xg_sampler *s = b->owner;
  XG_REQUIRE(s->stage_timing, "sampler was not created with XG_SAMPLER_STAGE_TIMING");
  const int L = (int)s->cfg.num_layer;
  int first;  // index of the mark that opens the stage
  switch (stage) {
    case XG_STAGE_RESET_SEEDS: first = 0; break;
    case XG_STAGE_SAMPLE: case XG_STAGE_INSERT: case XG_STAGE_NUMBER:
      XG_REQUIRE(layer >= 0 && layer < L, "layer out of range");
      first = 1 + 3 * (L - 1 - layer) + (stage - XG_STAGE_SAMPLE);
      break;
    case XG_STAGE_INPUT_NODES: first = 1 + 3 * L; break;
    case XG_STAGE_FEATURE: first = 2 + 3 * L; break;
    case XG_STAGE_LABEL: first = 3 + 3 * L; break;
    default: XG_REQUIRE(false, "unknown stage");
  }
  XG_REQUIRE(first + 1 < b->num_marks, "stage was not recorded");
  DeviceScope scope(s->device);
  XG_CUDA(cudaEventSynchronize(b->marks[first + 1]));
  XG_CUDA(cudaEventElapsedTime(ms, b->marks[first], b->marks[first + 1]));
  return XG_OK;
}

/* XG_TIMELINE=1 only: ms since sampler creation of {sample begin, sample end, extract begin, extract end} */
int xg_batch_timeline(xg_batch *b, float ms[4]) {
  XG_REQUIRE(b && b->owner && ms, "NULL argument");
  xg_sampler *s = b->owner;
  XG_REQUIRE(s->timeline && s->two_streams, "set XG_TIMELINE=1 (two-stream mode) before creating the sampler");
  DeviceScope scope(s->device);
  XG_CUDA(cudaEventSynchronize(b->tl[2]));
  XG_CUDA(cudaEventElapsedTime(&ms[0], s->origin, b->tl[0]));
  XG_CUDA(cudaEventElapsedTime(&ms[1], s->origin, b->sampled));
  XG_CUDA(cudaEventElapsedTime(&ms[2], s->origin, b->tl[1]));
  XG_CUDA(cudaEventElapsedTime(&ms[3], s->origin, b->tl[2]));
  return XG_OK;
}

void *xg_sampler_stream(xg_sampler *s) { return s ? s->stream : nullptr; }
void *xg_sampler_extract_stream(xg_sampler *s) { return s ? s->xstream : nullptr; }
uint32_t xg_sampler_launches_per_batch(xg_sampler *s) { return s ? s->launches_per_batch : 0; }

}  // extern "C"
