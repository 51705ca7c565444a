// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Pins the two host-side cache-rank routines nothing else in the reference exposes:
//   * ICS22SongPlacementSolver   cuda/ics22_song_dist_feature.cc:22-167 (file-local function)
//   * dist::PreSampler           dist/pre_sampler.cc:40-165 (count, 64-bit sort, GetRankNode/GetFreq)
// Both reference files are compiled IN PLACE by #including them from /root/reference into this
// translation unit (that is the only way to reach a function in an anonymous namespace); no line
// of them is restated here.  What this file adds is the scaffolding PreSampler::DoPreSample()
// expects around itself and that cannot be built here (DistEngine, GPUShuffler, DoGPUSample need
// the whole engine): a DistEngine whose methods are empty, a shuffler that hands out numbered
// batches, and a DoGPUSample that answers with the input-node list the caller supplied for that
// batch.  The counting loop, the sort and the rank extraction that run are the reference's.
//
// Linked into oracle/_ref/libxgnn_ref_cuda.so (oracle/Makefile, target `refcuda`); the two entry
// points below never touch a GPU, so CPU tests may call them.
#include "common/cuda/ics22_song_dist_feature.cc"  // NOLINT(build/include)
#include "common/dist/pre_sampler.cc"              // NOLINT(build/include)

#include <cstring>

namespace samgraph {
namespace common {

Engine *Engine::_engine = nullptr;  // engine.cc:60

namespace {
struct PresampleFeed {
  const uint32_t *nodes = nullptr;
  const size_t *offsets = nullptr;
  size_t next = 0, count = 0;
} g_feed;
}  // namespace

namespace cuda {
// Stand-in for cuda/cuda_shuffler.cc (which needs the engine's GPU context): same step count
// rule (:52-56), batches carry no data -- DoGPUSample below supplies the sampled node lists.
GPUShuffler::GPUShuffler(TensorPtr input, size_t num_epoch, size_t batch_size, bool drop_last) {
  _num_data = input->Shape().front();
  _num_epoch = num_epoch;
  _cur_epoch = 0;
  _data = input;
  _drop_last = drop_last;
  _batch_size = batch_size;
  _num_step = drop_last ? _num_data / batch_size : (_num_data + batch_size - 1) / batch_size;
  _initialized = false;
  _cur_step = _num_step;
}
TensorPtr GPUShuffler::GetBatch(StreamHandle) {
  _cur_step++;
  if (_cur_step >= _num_step) {
    if (_initialized) _cur_epoch++;
    _initialized = true;
    _cur_step = 0;
  }
  if (_cur_epoch >= _num_epoch) return nullptr;
  return _data;
}
}  // namespace cuda

namespace dist {
DistEngine::DistEngine() {
  _sampler_ctx = CPU(CPU_CLIB_MALLOC_DEVICE);
  _trainer_ctx = CPU(CPU_CLIB_MALLOC_DEVICE);
  _sample_stream = nullptr;
  _num_step = 1;
  _engine = this;
}
void DistEngine::Init() {}
void DistEngine::Start() {}
void DistEngine::Shutdown() {}
void DistEngine::RunSampleOnce() {}
void DistEngine::ArchCheck() {}
std::unordered_map<std::string, Context> DistEngine::GetGraphFileCtx() { return {}; }

// dist/dist_loops.cc:62-368 in the real engine; here: the caller's list for this batch.
void DoGPUSample(TaskPtr task) {
  if (g_feed.next >= g_feed.count) {
    fprintf(stderr, "ref_host_shim: DoPreSample asked for more batches than supplied\n");
    abort();
  }
  const size_t lo = g_feed.offsets[g_feed.next], hi = g_feed.offsets[g_feed.next + 1];
  g_feed.next++;
  // a Tensor frees its buffer (common.cc:68-75): hand it a copy, never the caller's array
  const Context cpu = CPU(CPU_CLIB_MALLOC_DEVICE);
  task->input_nodes = hi > lo ? Tensor::CopyBlob(g_feed.nodes + lo, kI32, {hi - lo}, cpu, cpu, "presample feed")
                              : Tensor::Empty(kI32, {0}, cpu, "presample feed");
}
}  // namespace dist

}  // namespace common
}  // namespace samgraph

using namespace samgraph::common;

extern "C" {

// ICS22SongPlacementSolver over `num_dev` contexts with device ids 0..num_dev-1.
//   sample_prob f64[num_node]; outputs (host): ranking u32[num_node],
//   device_map / new_idx_map u32[num_dev][num_node], cached_nodes u32[num_dev][num_cached].
// Returns total_cached_node.
uint32_t ref_ics22_solve(const double *sample_prob, uint32_t num_node, uint32_t num_cached,
                         uint32_t num_dev, double alpha, uint32_t clique_size, uint32_t *ranking,
                         uint32_t *device_map, uint32_t *new_idx_map, uint32_t *cached_nodes) {
  std::vector<double> prob(sample_prob, sample_prob + num_node);
  std::vector<Context> ctxes;
  for (uint32_t d = 0; d < num_dev; d++) ctxes.push_back(GPU(d));
  std::vector<TensorPtr> dev_map_vec, idx_map_vec, cached_vec;
  TensorPtr ranking_tensor = Tensor::Empty(kI32, {num_node}, CPU(CPU_CLIB_MALLOC_DEVICE), "ranking");
  IdType total = 0;
  if (RunConfig::omp_thread_num <= 0) RunConfig::omp_thread_num = 4;
  cuda::ICS22SongPlacementSolver(prob, num_cached, ctxes, alpha, clique_size, kI32, dev_map_vec,
                                 idx_map_vec, cached_vec, ranking_tensor, total);
  memcpy(ranking, ranking_tensor->Data(), sizeof(uint32_t) * num_node);
  for (uint32_t d = 0; d < num_dev; d++) {
    memcpy(device_map + static_cast<size_t>(d) * num_node, dev_map_vec[d]->Data(), 4ul * num_node);
    memcpy(new_idx_map + static_cast<size_t>(d) * num_node, idx_map_vec[d]->Data(), 4ul * num_node);
    if (num_cached)
      memcpy(cached_nodes + static_cast<size_t>(d) * num_cached, cached_vec[d]->Data(), 4ul * num_cached);
  }
  return total;
}

// dist::PreSampler: `epochs` x `num_step` batches whose sampled input-node lists are
// nodes[offsets[k] .. offsets[k+1]) in call order.  Outputs: rank_nodes / rank_freq u32[num_node].
void ref_presample_rank(uint32_t num_node, int epochs, size_t num_step, const uint32_t *nodes,
                        const size_t *offsets, uint32_t *rank_nodes, uint32_t *rank_freq) {
  static dist::DistEngine *engine = new dist::DistEngine();
  (void)engine;
  if (RunConfig::omp_thread_num <= 0) RunConfig::omp_thread_num = 4;
  RunConfig::presample_epoch = epochs;
  RunConfig::cache_policy = kCacheByPreSample;
  g_feed.nodes = nodes;
  g_feed.offsets = offsets;
  g_feed.next = 0;
  g_feed.count = static_cast<size_t>(epochs) * num_step;
  // a train set of num_step ids with batch size 1 gives the shuffler exactly num_step steps
  std::vector<uint32_t> train(num_step ? num_step : 1, 0);
  const Context cpu = CPU(CPU_CLIB_MALLOC_DEVICE);
  TensorPtr input = Tensor::CopyBlob(train.data(), kI32, {train.size()}, cpu, cpu, "train");
  {
    dist::PreSampler ps(input, 1, num_node);
    ps.DoPreSample();
    TensorPtr r = ps.GetRankNode();
    TensorPtr f = ps.GetFreq();
    memcpy(rank_nodes, r->Data(), 4ul * num_node);
    memcpy(rank_freq, f->Data(), 4ul * num_node);
  }
}

}  // extern "C"
