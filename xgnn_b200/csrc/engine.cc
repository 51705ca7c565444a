// Host orchestration behind the reference's C-ABI (SURVEY.md rows a1, a10, a15, a16 and boundary (b)).
//
// Reference behaviour being replaced (paths relative to samgraph/common/):
//   operation.cc:49-584            the samgraph_* entry points, config keys and their CHECKs
//   engine.cc:109-443              dataset loader (meta.txt + *.bin, anonymous shared mappings for fork)
//   dist/dist_engine.cc:91-700     DistEngine::Init / SampleInit / TrainInit (arch6)
//   dist/dist_shuffler_aligned.cc  per-GPU slice of a globally shuffled train set
//   dist/dist_loops_arch6.cc       RunSampleSubLoopOnce + RunCacheDataCopySubLoopOnce
//   cuda/dist_graph.cu:228-547     GGMS GraphLoad / FeatureLoad / IPC share
//   cuda/cuda_cache_manager_host.cc:133-254  cache ranking -> direct-address table
//   graph_pool.cc, profiler.cc     batch hand-off and the log items the Python scripts read back
//
// What is different by design: a Task is one xg_sampler_submit (one replayed CUDA graph, pipeline.cu), so the
// reference's task queues between stage threads, its WorkspacePool and its per-stage host syncs have no
// counterpart; partition builders run on the GPU straight from the registered host mapping; IPC handles
// travel through one MAP_SHARED block created in data_init, guarded by a process-shared pthread barrier.
#include <fcntl.h>
#include <pthread.h>
#include <semaphore.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <random>
#include <sstream>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/samgraph_abi.h"
#include "common.cuh"

namespace sam {

// ---------------------------------------------------------------------------------------------------
// CHECK: print file:line + condition, abort (logging.h:32-46, logging.cc:69-73)
[[noreturn]] void fatal(const char *file, int line, const std::string &msg) {
  std::fprintf(stderr, "[%s:%d] FATAL: %s\n", file, line, msg.c_str());
  std::fflush(stderr);
  std::abort();
}
#define SAM_CHECK(cond)                                              \
  do {                                                               \
    if (!(cond)) ::sam::fatal(__FILE__, __LINE__, "Check failed: " #cond); \
  } while (0)
#define SAM_CHECK_MSG(cond, msg)                                                              \
  do {                                                                                        \
    if (!(cond)) ::sam::fatal(__FILE__, __LINE__, std::string("Check failed: " #cond " ") + (msg)); \
  } while (0)
#define SAM_XG(expr)                                                                    \
  do {                                                                                  \
    int _rc = (expr);                                                                   \
    if (_rc != 0) ::sam::fatal(__FILE__, __LINE__, std::string(#expr ": ") + xg_last_error()); \
  } while (0)
#define SAM_CUDA(expr)                                                                        \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) ::sam::fatal(__FILE__, __LINE__, std::string(#expr ": ") + cudaGetErrorString(_e)); \
  } while (0)

int log_level() {
  static int lvl = [] {
    const char *e = std::getenv("SAMGRAPH_LOG_LEVEL");
    if (!e) return 2;  // warning
    std::string s(e);
    if (s == "debug") return 0;
    if (s == "info") return 1;
    if (s == "warn" || s == "warning") return 2;
    return 3;
  }();
  return lvl;
}
void info(const char *fmt, ...) {
  if (log_level() > 1) return;
  va_list ap;
  va_start(ap, fmt);
  std::fprintf(stderr, "[samgraph] ");
  std::vfprintf(stderr, fmt, ap);
  std::fprintf(stderr, "\n");
  va_end(ap);
}

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------------
// RunConfig (run_config.h:31-133, operation.cc:49-326)
enum RunArch { kArch0 = 0, kArch1, kArch2, kArch3, kArch4, kArch5, kArch6, kArch7, kArch9 = 9 };
enum CachePolicy { kCacheByDegree = 0, kCacheByHeuristic, kCacheByPreSample, kCacheByDegreeHop, kCacheByPreSampleStatic,
                   kCacheByFakeOptimal, kDynamicCache, kCacheByRandom };

struct RunConfig {
  bool is_configured = false;
  std::unordered_map<std::string, std::string> raw;
  std::string dataset_path;
  int run_arch = 0, sample_type = 0, cache_policy = 0;
  size_t batch_size = 0, num_epoch = 0, num_layer = 0;
  double cache_percentage = 0.0;
  int presample_epoch = 0;            // operation.cc:184-189
  double replicate_percentage = 0.0;  // `ggms_replicate_percentage`: hot tier replicated per GPU (0 = reference layout)
  std::vector<size_t> fanout;
  size_t random_walk_length = 0, num_random_walk = 0, num_neighbor = 0;
  double random_walk_restart_prob = 0.0;
  int sampler_device = 0, trainer_device = 0;
  size_t num_worker = 1;
  bool use_dist_graph = false, part_cache = false, gpu_extract = false;
  size_t num_train_worker = 0;         // arch5 (num_worker = num_sample_worker there)
  bool use_ics22_song_solver = false;  // operation.cc:213-230
  int clique_size = 0;
  double ics22_song_alpha = 0.0;
  int ics22_compact_bitwidth = 3;      // run_config.cc:99-100 (compact mode is the reference default)
  double dist_graph_percentage = 0.0;
  int omp_thread_num = 1;
};
RunConfig g_cfg;

int parse_ctx(const std::string &ctx) {  // "cuda:N" -> N ; "cpu" -> -1   (common.cc Context(std::string))
  if (ctx.rfind("cuda:", 0) == 0) return std::stoi(ctx.substr(5));
  SAM_CHECK_MSG(ctx.rfind("cpu", 0) == 0, "context must be cuda:N or cpu, got " + ctx);
  return -1;
}

// ---------------------------------------------------------------------------------------------------
// Dataset (engine.cc:109-443): meta.txt + raw little-endian arrays
struct Mapping {
  void *ptr = nullptr;
  size_t bytes = 0;
};
struct Dataset {
  size_t num_node = 0, num_edge = 0, num_class = 0, feat_dim = 0;
  size_t num_train = 0, num_valid = 0, num_test = 0;
  int feat_dtype = XG_F32;
  const uint32_t *indptr = nullptr, *indices = nullptr, *train_set = nullptr, *ranking = nullptr, *alias = nullptr;
  const void *feat = nullptr;
  const int64_t *label = nullptr;
  const float *prob = nullptr, *prefix = nullptr;
  std::vector<uint32_t> ranking_owned;
  uint32_t *ranking_rw = nullptr;  // kCacheByPreSample: shared, written by worker 0 at sample_init
  std::vector<Mapping> maps;
  // SAMGRAPH_EMPTY_FEAT = k (constant.cc kEnvEmptyFeat, run_config.cc:138): mock features - the host table has 2^k rows
  // and node v reads row v & (2^k - 1) everywhere (engine.cc:198-208, cuda_extraction.cu:51-72)
  unsigned empty_feat = 0;
  size_t feat_table_rows() const { return empty_feat ? (size_t(1) << empty_feat) : num_node; }
  uint32_t feat_row_mask() const { return empty_feat ? (uint32_t)((size_t(1) << empty_feat) - 1) : 0u; }
  static size_t elem_bytes(int dtype) {  // GetDataTypeBytes, common.cc
    switch (dtype) {
      case XG_F16: return 2;
      case XG_F64: case XG_I64: return 8;
      case XG_U8: case XG_I8: return 1;
      default: return 4;
    }
  }
  size_t feat_row_bytes() const { return feat_dim * elem_bytes(feat_dtype); }
};

bool file_exists(const std::string &p) {
  struct stat st;
  return ::stat(p.c_str(), &st) == 0;
}

// file -> memory.  `shared`: an anonymous MAP_SHARED copy that survives fork() and can be registered with
// CUDA by every child (ConverToAnonMmap, engine.cc:91-107); otherwise a private read-only file mapping.
const void *map_file(Dataset &ds, const std::string &path, size_t want_bytes, bool shared) {
  int fd = ::open(path.c_str(), O_RDONLY);
  SAM_CHECK_MSG(fd >= 0, "cannot open " + path);
  struct stat st;
  SAM_CHECK(::fstat(fd, &st) == 0);
  SAM_CHECK_MSG((size_t)st.st_size >= want_bytes, path + " is smaller than meta.txt says");
  if (want_bytes == 0) {
    ::close(fd);
    return nullptr;
  }
  void *file = ::mmap(nullptr, want_bytes, PROT_READ, MAP_PRIVATE, fd, 0);
  SAM_CHECK_MSG(file != MAP_FAILED, "mmap failed for " + path);
  ::close(fd);
  if (!shared) {
    ds.maps.push_back({file, want_bytes});
    return file;
  }
  void *anon = ::mmap(nullptr, want_bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  SAM_CHECK_MSG(anon != MAP_FAILED, "anonymous mmap failed for " + path);
  std::memcpy(anon, file, want_bytes);
  ::munmap(file, want_bytes);
  ds.maps.push_back({anon, want_bytes});
  return anon;
}

int parse_dtype(const std::string &s) {  // DataTypeParseName, common.cc
  if (s == "F32") return XG_F32;
  if (s == "F64") return XG_F64;
  if (s == "F16") return XG_F16;
  if (s == "U8") return XG_U8;
  if (s == "I32") return XG_I32;
  if (s == "I8") return XG_I8;
  if (s == "I64") return XG_I64;
  fatal(__FILE__, __LINE__, "unknown FEAT_DATA_TYPE " + s);
}

void load_dataset(Dataset &ds, const RunConfig &cfg, bool shared) {
  std::string dir = cfg.dataset_path;
  if (dir.empty() || dir.back() != '/') dir.push_back('/');
  std::ifstream meta(dir + "meta.txt");
  SAM_CHECK_MSG(meta.is_open(), "cannot open " + dir + "meta.txt");
  std::unordered_map<std::string, size_t> kv;
  std::string k, v;
  while (meta >> k >> v) {
    if (k == "FEAT_DATA_TYPE") ds.feat_dtype = parse_dtype(v);
    else kv[k] = std::stoull(v);
  }
  for (const char *need : {"NUM_NODE", "NUM_EDGE", "FEAT_DIM", "NUM_CLASS", "NUM_TRAIN_SET", "NUM_TEST_SET", "NUM_VALID_SET"})
    SAM_CHECK_MSG(kv.count(need) > 0, std::string("meta.txt lacks ") + need);
  ds.num_node = kv["NUM_NODE"];
  ds.num_edge = kv["NUM_EDGE"];
  ds.feat_dim = kv["FEAT_DIM"];
  ds.num_class = kv["NUM_CLASS"];
  ds.num_train = kv["NUM_TRAIN_SET"];
  ds.num_test = kv["NUM_TEST_SET"];
  ds.num_valid = kv["NUM_VALID_SET"];
  SAM_CHECK(ds.num_node < 0xFFFFFFFFull && ds.num_edge <= 0xFFFFFFFFull);  // IdType is 32-bit (constant.h:28)
  ds.indptr = static_cast<const uint32_t *>(map_file(ds, dir + "indptr.bin", (ds.num_node + 1) * 4, shared));
  ds.indices = static_cast<const uint32_t *>(map_file(ds, dir + "indices.bin", ds.num_edge * 4, shared));
  if (const char *ef = std::getenv("SAMGRAPH_EMPTY_FEAT")) ds.empty_feat = (unsigned)std::strtoul(ef, nullptr, 10);
  SAM_CHECK_MSG(ds.empty_feat < 32, "SAMGRAPH_EMPTY_FEAT must be below 32");
  if (ds.empty_feat) {
    // Tensor::Empty of 2^k rows in the reference (engine.cc:198-213): contents undefined there; here a fixed byte
    // pattern of the row number so that a run can be checked (byte i of row r = (31 r + i) & 0x3F: small finite
    // values in every dtype).  Anonymous shared mapping: survives fork like the mapped files.
    const size_t bytes = ds.feat_table_rows() * ds.feat_row_bytes();
    void *mem = ::mmap(nullptr, std::max<size_t>(bytes, 4096), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    SAM_CHECK_MSG(mem != MAP_FAILED, "anonymous mmap failed for the mock feature table");
    auto *b = static_cast<unsigned char *>(mem);
    const size_t rb = ds.feat_row_bytes();
    for (size_t r = 0; r < ds.feat_table_rows(); ++r)
      for (size_t i = 0; i < rb; ++i) b[r * rb + i] = (unsigned char)((31 * r + i) & 0x3F);
    ds.maps.push_back({mem, std::max<size_t>(bytes, 4096)});
    ds.feat = mem;
  } else {
    SAM_CHECK_MSG(file_exists(dir + "feat.bin"), "feat.bin is missing: provide it or set SAMGRAPH_EMPTY_FEAT=k (2^k mock rows)");
    ds.feat = map_file(ds, dir + "feat.bin", ds.num_node * ds.feat_row_bytes(), shared);
  }
  SAM_CHECK_MSG(file_exists(dir + "label.bin"), "label.bin is missing");
  ds.label = static_cast<const int64_t *>(map_file(ds, dir + "label.bin", ds.num_node * 8, shared));
  ds.train_set = static_cast<const uint32_t *>(map_file(ds, dir + "train_set.bin", ds.num_train * 4, shared));
  if (cfg.sample_type == XG_WEIGHTED_KHOP || cfg.sample_type == XG_WEIGHTED_KHOP_HASH_DEDUP) {
    ds.prob = static_cast<const float *>(map_file(ds, dir + "prob_table.bin", ds.num_edge * 4, shared));
    ds.alias = static_cast<const uint32_t *>(map_file(ds, dir + "alias_table.bin", ds.num_edge * 4, shared));
  } else if (cfg.sample_type == XG_WEIGHTED_KHOP_PREFIX) {
    ds.prefix = static_cast<const float *>(map_file(ds, dir + "prob_prefix_table.bin", ds.num_edge * 4, shared));
  }
  if (cfg.cache_percentage > 0.0) {
    const char *name = nullptr;
    if (cfg.cache_policy == kCacheByPreSample) {
      // the ranking is produced at sample_init by worker 0 and read by every trainer (dist_engine.cc:455-466): it
      // lives in an anonymous shared mapping created before fork, like the reference's ranking_nodes tensor
      void *mem = ::mmap(nullptr, std::max<size_t>(ds.num_node * 4, 4096), PROT_READ | PROT_WRITE,
                         MAP_SHARED | MAP_ANONYMOUS, -1, 0);
      SAM_CHECK(mem != MAP_FAILED);
      ds.ranking_rw = static_cast<uint32_t *>(mem);
      std::iota(ds.ranking_rw, ds.ranking_rw + ds.num_node, 0u);
      ds.ranking = ds.ranking_rw;
      return;
    }
    switch (cfg.cache_policy) {
      case kCacheByDegree: name = "cache_by_degree.bin"; break;
      case kCacheByHeuristic: name = "cache_by_heuristic.bin"; break;
      case kCacheByDegreeHop: name = "cache_by_degree_hop.bin"; break;
      case kCacheByFakeOptimal: name = "cache_by_fake_optimal.bin"; break;
      case kCacheByRandom: name = "cache_by_random.bin"; break;
      default: fatal(__FILE__, __LINE__, "cache policy needs presampling / dynamic cache: out of scope (DESIGN.md section 8)");
    }
    if (file_exists(dir + name)) {
      ds.ranking = static_cast<const uint32_t *>(map_file(ds, dir + name, ds.num_node * 4, shared));
    } else {
      // no precomputed ranking on disk: nodes by out-degree, descending, stable
      // (what utility/data-process/toolkit/cache/cache_by_degree.cc:29-63 writes into cache_by_degree.bin)
      SAM_CHECK_MSG(cfg.cache_policy == kCacheByDegree, dir + name + " is missing");
      ds.ranking_owned.resize(ds.num_node);
      std::iota(ds.ranking_owned.begin(), ds.ranking_owned.end(), 0u);
      const uint32_t *ip = ds.indptr;
      std::stable_sort(ds.ranking_owned.begin(), ds.ranking_owned.end(),
                       [ip](uint32_t a, uint32_t b) { return ip[a + 1] - ip[a] > ip[b + 1] - ip[b]; });
      ds.ranking = ds.ranking_owned.data();
    }
  }
}

}  // namespace sam

namespace sam {

// ---------------------------------------------------------------------------------------------------
// Profiler, reduced to what the Python side reads back (profiler.h:30-243; the enum numbering is ABI)
constexpr int kNumLogInitItems = 32, kNumLogStepItems = 64, kNumLogEpochItems = 16;
constexpr int kLogL1NumSample = 0, kLogL1NumNode = 1, kLogL1SampleTotalTime = 2, kLogL1SampleTime = 3, kLogL1SendTime = 4,
              kLogL1RecvTime = 5, kLogL1CopyTime = 6,
              kLogL1FeatureBytes = 9, kLogL1LabelBytes = 10, kLogL1IdBytes = 11, kLogL1GraphBytes = 12,
              kLogL2ShuffleTime = 17, kLogL2CoreSampleTime = 20;
constexpr int kLogEpochSampleTime = 0, kLogEpochSampleSendTime = 2, kLogEpochSampleTotalTime = 3, kLogEpochCopyTime = 8, kLogEpochFeatureBytes = 12,
              kLogEpochNumSample = 15;
constexpr int kLogInitL1Common = 0, kLogInitL1Sampler = 1, kLogInitL1Trainer = 2, kLogInitL2LoadDataset = 6,
              kLogInitL2Presample = 8, kLogInitL2BuildCache = 10;

struct Profiler {
  std::mutex mu;
  double init[kNumLogInitItems] = {};
  std::unordered_map<uint64_t, std::vector<double>> step;   // batch key -> items
  std::unordered_map<uint64_t, std::vector<double>> epoch;  // epoch -> items
  std::vector<double> &step_row(uint64_t key) {
    auto &r = step[key];
    if (r.empty()) r.assign(kNumLogStepItems, 0.0);
    return r;
  }
  std::vector<double> &epoch_row(uint64_t e) {
    auto &r = epoch[e];
    if (r.empty()) r.assign(kNumLogEpochItems, 0.0);
    return r;
  }
  void log_step(uint64_t key, int item, double v, bool add) {
    std::lock_guard<std::mutex> lk(mu);
    auto &r = step_row(key);
    r[item] = add ? r[item] + v : v;
  }
  void log_epoch_add(uint64_t e, int item, double v) {
    std::lock_guard<std::mutex> lk(mu);
    epoch_row(e)[item] += v;
  }
  double get_step(uint64_t key, int item) {
    std::lock_guard<std::mutex> lk(mu);
    auto it = step.find(key);
    return it == step.end() ? 0.0 : it->second[item];
  }
  double get_epoch(uint64_t e, int item) {
    std::lock_guard<std::mutex> lk(mu);
    auto it = epoch.find(e);
    return it == epoch.end() ? 0.0 : it->second[item];
  }
};
Profiler g_prof;

// ---------------------------------------------------------------------------------------------------
// DistAlignedShuffler (dist/dist_shuffler_aligned.cc:37-146).  Same generator, same distribution object,
// same swap loop -> the same permutation as the reference for every epoch, in every worker.
struct AlignedShuffler {
  std::vector<uint32_t> data;  // whole train set padded to a multiple of num_worker with its own head
  size_t num_epoch, batch_size, num_local_data, num_local_step, global_step_offset, global_data_offset;
  size_t cur_epoch = 0, cur_local_step = 0;
  bool initialized = false;
  uint32_t *d_local = nullptr;  // this worker's slice on the sampler GPU

  AlignedShuffler(const uint32_t *train, size_t n, size_t num_epoch_, size_t batch, size_t worker, size_t num_worker)
      : num_epoch(num_epoch_), batch_size(batch) {
    SAM_CHECK(batch > 0 && num_worker > 0 && n > 0);
    const size_t padded = (n + num_worker - 1) / num_worker * num_worker;
    data.assign(train, train + n);
    for (size_t i = 0; i < padded - n; ++i) data.push_back(train[i]);  // :52-55 (wrap-around duplicates)
    num_local_data = padded / num_worker;
    num_local_step = (num_local_data + batch - 1) / batch;
    global_step_offset = num_local_step * worker;
    global_data_offset = num_local_data * worker;
    cur_local_step = num_local_step;
    SAM_CUDA(cudaMalloc(&d_local, num_local_data * sizeof(uint32_t)));
  }
  ~AlignedShuffler() { cudaFree(d_local); }

  static void shuffle_epoch(std::vector<uint32_t> &v, size_t epoch) {
    auto g = std::default_random_engine(epoch);  // :86-90: every sampler uses seed = epoch
    for (size_t i = 0; i + 1 < v.size(); ++i) {
      std::uniform_int_distribution<size_t> d(i, v.size() - 1);
      std::swap(v[i], v[d(g)]);
    }
  }
  void reshuffle(cudaStream_t s) {
    if (!initialized) { cur_epoch = 0; initialized = true; }
    else ++cur_epoch;
    cur_local_step = 0;
    if (cur_epoch >= num_epoch) return;
    shuffle_epoch(data, cur_epoch);
    SAM_CUDA(cudaMemcpyAsync(d_local, data.data() + global_data_offset, num_local_data * sizeof(uint32_t),
                             cudaMemcpyHostToDevice, s));
    SAM_CUDA(cudaStreamSynchronize(s));
  }
  // next batch of this worker: device pointer + length; false once every epoch is done
  bool get_batch(cudaStream_t s, const uint32_t **d_seeds, size_t *n, size_t *epoch, size_t *local_step) {
    ++cur_local_step;
    if (cur_local_step >= num_local_step) reshuffle(s);
    if (cur_epoch >= num_epoch) return false;
    const size_t off = cur_local_step * batch_size;
    SAM_CHECK(off < num_local_data);
    *n = std::min(batch_size, num_local_data - off);
    *d_seeds = d_local + off;
    *epoch = cur_epoch;
    *local_step = cur_local_step;
    return true;
  }
};

// ---------------------------------------------------------------------------------------------------
// Cross-process state created by data_init before fork (DistGraph::Create, cuda/dist_graph.cu:566-631): the
// placement, a process-shared barrier and the CUDA IPC handles of every shard.
constexpr int kMaxWorker = XG_MAX_PARTS;
// ---------------------------------------------------------------------------------------------------
// arch5: samplers and trainers are different processes on different GPUs.  The reference moves every
// Task through a host shared-memory ring: D2H of ids + per-layer row/col into a fixed 50 MB slot, memcpy, then H2D on
// the trainer (MemoryQueue, memory_queue.h:65-130, memory_queue.cc:69-141; TransData/ToData/ParseData, task_queue.cc:
// 66-330).  Here only a ~400-byte descriptor crosses host shared memory; the payload is packed once into a "stage"
// buffer in the SAMPLER's HBM (exported over CUDA IPC) and the trainer pulls it with ONE peer copy over NVLink.
// Tickets + per-entry semaphores follow the reference's queue (send_cnt / recv_cnt, sem_list / release_list).
constexpr int kQueueDepth = 16, kStagePerSampler = 4;
struct TaskDesc {
  uint64_t key, epoch, bytes;
  uint32_t num_layer, num_input, num_output, sampler, stage, stage_gen;
  uint64_t num_src[XG_MAX_LAYERS], num_dst[XG_MAX_LAYERS], num_edge[XG_MAX_LAYERS];
  // byte offsets inside the stage buffer; kNone = that array does not exist for this batch
  uint64_t off_input, off_output, off_row[XG_MAX_LAYERS], off_col[XG_MAX_LAYERS], off_data[XG_MAX_LAYERS],
      off_csc[XG_MAX_LAYERS];
  double sample_time;
};
constexpr uint64_t kNone = ~0ull;
struct TaskQueueBlock {
  uint64_t send_cnt, recv_cnt;  // tickets (MQ_MetaData::Put/Get)
  sem_t filled[kQueueDepth], vacant[kQueueDepth];
  TaskDesc desc[kQueueDepth];
  sem_t stage_free[kMaxWorker][kStagePerSampler];
  unsigned char stage_ipc[kMaxWorker][kStagePerSampler][64];
  uint32_t stage_gen[kMaxWorker][kStagePerSampler];  // bumped when a stage buffer is re-allocated (grown)
  int num_sampler, num_trainer;
};

struct SharedBlock {
  pthread_barrier_t barrier;
  int num_worker;
  int num_parts[kMaxWorker];
  int part_ids[kMaxWorker][kMaxWorker];
  int access[kMaxWorker][kMaxWorker];  // access[g][p]: worker that g reads part p from
  unsigned char topo_indptr[kMaxWorker][kMaxWorker][64];   // [owner][part]
  unsigned char topo_indices[kMaxWorker][kMaxWorker][64];
  unsigned char feat[kMaxWorker][kMaxWorker][64];
};

}  // namespace sam

#include <cub/device/device_scan.cuh>

namespace sam {

// ---------------------------------------------------------------------------------------------------
// One finished Task, shared between the pool, the "current batch" and every tensor view handed out
// (TensorPtr captured by torch::from_blob's deleter, torch/adapter.cc:70-76)
struct BatchHolder {
  xg_batch *batch = nullptr;
  xg_batch_view view{};
  uint64_t key = 0;
  std::function<void()> on_release;  // arch5 trainer: the landing buffers behind `view` go back to their pool
  ~BatchHolder();
};
using BatchPtr = std::shared_ptr<BatchHolder>;

// a Task that has been handed to the sampler engine but not waited for yet
struct PendingBatch {
  BatchPtr holder;
  size_t epoch = 0;
  double t_submit = 0, t_shuffle = 0;
};

struct Engine {
  Dataset ds;
  SharedBlock *shared = nullptr;  // arch6 only
  bool dataset_loaded = false, sampler_ready = false, trainer_ready = false, shutdown = false;
  int worker_id = 0, device = 0;
  size_t num_step = 0, num_local_step = 0;
  cudaStream_t stream = nullptr;
  // GGMS handles
  xg_graph graph{};
  xg_feature_store store{};
  std::vector<void *> owned;  // device allocations freed at shutdown
  void *d_label = nullptr;
  const float *d_prob = nullptr, *d_prefix = nullptr;
  const uint32_t *d_alias = nullptr;
  std::unique_ptr<AlignedShuffler> shuffler;
  xg_sampler *sampler = nullptr;
  // GraphPool (graph_pool.cc:31-49) + current batch (operation.cc:366-378)
  std::mutex mu;
  std::condition_variable cv;
  std::deque<BatchPtr> pool;
  BatchPtr current;
  std::vector<std::thread> threads;
  bool producer_done = false;
  // Tasks in flight inside the sampler engine (submitted, not yet waited for): the producers keep kInFlight of them
  // there so that the extract phase of Task k runs beside the sample phase of Task k+1 (pipeline.cu)
  std::deque<PendingBatch> inflight;
  bool shuffler_done = false;
  double last_finish = 0;
  std::mutex slot_mu;
  std::condition_variable slot_cv;  // a batch slot went back to the sampler engine
  uint64_t slot_gen = 0;            // bumped (under slot_mu) by every release: no wake-up can be lost
  // arch5
  TaskQueueBlock *queue = nullptr;
  int role = 0;  // 0 = sampler + trainer in one process (arch0-4, 6), 1 = arch5 sampler, 2 = arch5 trainer
  struct Stage { void *ptr = nullptr; size_t cap = 0; } stages[kStagePerSampler];
  unsigned next_stage = 0;
  struct Landing { void *xfer = nullptr, *feat = nullptr; int64_t *label = nullptr; size_t cap_xfer = 0, cap_feat = 0, cap_label = 0; bool busy = false; };
  std::vector<std::unique_ptr<Landing>> landings;
  struct PeerStage { void *ptr = nullptr; uint32_t gen = 0; };
  PeerStage peer_stage[kMaxWorker][kStagePerSampler];
  cudaStream_t copy_stream = nullptr;  // arch5 trainer: Task k+1's peer copy, under Task k's gather
  // SAMGRAPH_SANITY_CHECK (constant.cc:58, run_config.cc option_sanity_check)
  bool sanity_check = false;
  uint64_t *d_sanity = nullptr;
  uint32_t *d_epoch_map = nullptr;
  size_t sanity_epoch = (size_t)-1;
};
Engine *g_engine = nullptr;

BatchHolder::~BatchHolder() {
  if (batch) xg_batch_release(batch);
  if (on_release) on_release();
  if (batch && g_engine) {
    {
      std::lock_guard<std::mutex> lk(g_engine->slot_mu);
      ++g_engine->slot_gen;
    }
    g_engine->slot_cv.notify_all();
  }
}

Engine &engine() {
  SAM_CHECK_MSG(g_engine != nullptr, "engine is not created: call samgraph_init or samgraph_data_init first");
  return *g_engine;
}

void *dev_alloc(Engine &e, size_t bytes) {
  void *p = nullptr;
  SAM_XG(xg_device_malloc(&p, bytes));
  e.owned.push_back(p);
  return p;
}

// make a host range loadable from kernels (cudaHostRegister of the dataset mappings, dist_engine.cc:219-222,
// 239-240); falls back to a pinned copy where the mapping cannot be registered (private file mappings)
const void *pin_host(const void *p, size_t bytes) {
  if (!p || bytes == 0) return p;
  cudaError_t err = cudaHostRegister(const_cast<void *>(p), bytes, cudaHostRegisterDefault | cudaHostRegisterReadOnly);
  if (err == cudaSuccess || err == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();
    return p;
  }
  cudaGetLastError();
  err = cudaHostRegister(const_cast<void *>(p), bytes, cudaHostRegisterDefault);
  if (err == cudaSuccess) return p;
  cudaGetLastError();
  void *copy = nullptr;
  SAM_CUDA(cudaHostAlloc(&copy, bytes, cudaHostAllocDefault));
  std::memcpy(copy, p, bytes);
  return copy;
}

template <typename T>
const T *to_device(Engine &e, const T *host, size_t count) {
  if (!host || count == 0) return nullptr;
  void *d = dev_alloc(e, count * sizeof(T));
  SAM_CUDA(cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice));
  return static_cast<const T *>(d);
}

void barrier(Engine &e) {
  if (e.shared && e.shared->num_worker > 1) {
    int rc = pthread_barrier_wait(&e.shared->barrier);
    SAM_CHECK(rc == 0 || rc == PTHREAD_BARRIER_SERIAL_THREAD);
  }
}

// ---- topology: DistEngine::SampleDataCopy + DistGraph::GraphLoad (dist_engine.cc:198-244, dist_graph.cu:309-385)
void build_topology(Engine &e) {
  const Dataset &ds = e.ds;
  const RunConfig &cfg = g_cfg;
  std::memset(&e.graph, 0, sizeof(e.graph));
  e.graph.num_node = (uint32_t)ds.num_node;
  const bool dist = cfg.run_arch == kArch6 && cfg.use_dist_graph && e.shared != nullptr;
  if (!dist) {  // whole CSR in this GPU's HBM (DeviceNormalGraph)
    e.graph.num_part = 0;
    e.graph.part_indptr[0] = to_device(e, ds.indptr, ds.num_node + 1);
    e.graph.part_indices[0] = to_device(e, ds.indices, std::max<size_t>(ds.num_edge, 1));
    return;
  }
  const uint32_t P = (uint32_t)e.shared->num_worker;
  const uint32_t ncn = xg_num_cache_node(ds.indptr, (uint32_t)ds.num_node, cfg.dist_graph_percentage);
  const uint32_t *h_indptr = static_cast<const uint32_t *>(pin_host(ds.indptr, (ds.num_node + 1) * 4));
  const uint32_t *h_indices = static_cast<const uint32_t *>(pin_host(ds.indices, ds.num_edge * 4));
  e.graph.num_part = P;
  e.graph.num_cache_node = ncn;
  e.graph.part_indptr[P] = h_indptr;  // slot P: the whole graph on the host (dist_graph.cu:372-383)
  e.graph.part_indices[P] = h_indices;
  const int w = e.worker_id;
  std::vector<const uint32_t *> my_indptr(P, nullptr), my_indices(P, nullptr);
  for (int k = 0; k < e.shared->num_parts[w]; ++k) {
    const uint32_t p = (uint32_t)e.shared->part_ids[w][k];
    const size_t rows = xg_part_num_rows(ncn, p, P);
    auto *d_ptr = static_cast<uint32_t *>(dev_alloc(e, (rows + 1) * sizeof(uint32_t)));
    SAM_CUDA(cudaMemsetAsync(d_ptr, 0, (rows + 1) * sizeof(uint32_t), e.stream));
    size_t edges = 0;
    if (rows) {
      uint32_t *d_deg = nullptr;
      SAM_CUDA(cudaMalloc(&d_deg, (rows + 1) * sizeof(uint32_t)));
      SAM_CUDA(cudaMemsetAsync(d_deg, 0, (rows + 1) * sizeof(uint32_t), e.stream));
      SAM_XG(xg_part_degrees(h_indptr, ncn, p, P, d_deg, e.stream));
      size_t tmp_bytes = 0;
      SAM_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_deg, d_ptr, (int)(rows + 1), e.stream));
      void *tmp = nullptr;
      SAM_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
      SAM_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, d_deg, d_ptr, (int)(rows + 1), e.stream));
      uint32_t last = 0;
      SAM_CUDA(cudaMemcpyAsync(&last, d_ptr + rows, sizeof(uint32_t), cudaMemcpyDeviceToHost, e.stream));
      SAM_CUDA(cudaStreamSynchronize(e.stream));
      cudaFree(tmp);
      cudaFree(d_deg);
      edges = last;
    }
    auto *d_idx = static_cast<uint32_t *>(dev_alloc(e, std::max<size_t>(edges, 1) * sizeof(uint32_t)));
    if (edges) SAM_XG(xg_part_copy_indices(h_indptr, h_indices, ncn, p, P, d_ptr, d_idx, e.stream));
    SAM_CUDA(cudaStreamSynchronize(e.stream));
    my_indptr[p] = d_ptr;
    my_indices[p] = d_idx;
    SAM_XG(xg_ipc_export(d_ptr, e.shared->topo_indptr[w][p]));
    SAM_XG(xg_ipc_export(d_idx, e.shared->topo_indices[w][p]));
  }
  barrier(e);  // every handle is published (DistGraph::_DataIpcShare, dist_graph.cu:274-307)
  for (uint32_t p = 0; p < P; ++p) {
    const int owner = e.shared->access[w][p];
    if (owner == w) {
      SAM_CHECK(my_indptr[p] != nullptr);
      e.graph.part_indptr[p] = my_indptr[p];
      e.graph.part_indices[p] = my_indices[p];
    } else {
      void *a = nullptr, *b = nullptr;
      SAM_XG(xg_ipc_import(e.shared->topo_indptr[owner][p], &a));
      SAM_XG(xg_ipc_import(e.shared->topo_indices[owner][p], &b));
      e.graph.part_indptr[p] = static_cast<const uint32_t *>(a);
      e.graph.part_indices[p] = static_cast<const uint32_t *>(b);
    }
  }
  barrier(e);
}

// ---- features: GPUCacheManager ctor + DistGraph::FeatureLoad (cuda_cache_manager_host.cc:133-254, dist_graph.cu:387-547)
void build_feature_store(Engine &e) {
  const Dataset &ds = e.ds;
  const RunConfig &cfg = g_cfg;
  std::memset(&e.store, 0, sizeof(e.store));
  const size_t row_bytes = ds.feat_row_bytes();
  e.store.row_bytes = row_bytes;
  const size_t V = ds.num_node;
  size_t n_cached = (size_t)(V * cfg.cache_percentage);
  if (n_cached > V) n_cached = V;
  const size_t T = ds.feat_table_rows();  // rows of the host table: V, or 2^k with SAMGRAPH_EMPTY_FEAT=k
  e.store.host_row_mask = ds.feat_row_mask();
  if (n_cached < V) e.store.host = pin_host(ds.feat, T * row_bytes);  // misses: zero-copy UVA loads
  if (n_cached == 0) return;                                          // cache_table == NULL: every row from `host`
  const bool ics22 = cfg.run_arch == kArch6 && cfg.use_ics22_song_solver && e.shared != nullptr;
  const bool parted = (cfg.run_arch == kArch6 && cfg.part_cache && e.shared != nullptr) || ics22;
  const uint32_t P = parted ? (uint32_t)e.shared->num_worker : 1u;
  if (!parted && n_cached == V) {
    // every row cached in one part: lay the part out by node id (where a cached row sits is ours to choose) and
    // skip the direct-address table - no 4-byte random lookup (a 128-byte DRAM line each) in front of every row
    const void *h = pin_host(ds.feat, T * row_bytes);
    void *d_all = dev_alloc(e, T * row_bytes);
    SAM_CUDA(cudaMemcpyAsync(d_all, h, T * row_bytes, cudaMemcpyHostToDevice, e.stream));
    SAM_CUDA(cudaStreamSynchronize(e.stream));
    e.store.host = d_all;  // "host" slot of the handle = the flat table, wherever it lives
    return;
  }
  std::vector<uint32_t> ranks(ds.ranking, ds.ranking + n_cached);
  // two-tier store (B200 extension, config key `ggms_replicate_percentage`, default 0 = the reference layout): the
  // hottest ranks keep the policy order and are replicated per GPU; only the sharded remainder is shuffled
  size_t n_rep = parted ? (size_t)(V * cfg.replicate_percentage) : 0;
  if (n_rep > n_cached) n_rep = n_cached;
  std::sort(ranks.begin(), ranks.begin() + n_rep);  // replicated tier by node id: resolved by rank/select (replica_words)
  if (parted && n_cached > n_rep) {
    if (n_rep && n_cached == V)  // all cached: node-id order (round-robin over parts), covered by the rank/select words
      std::sort(ranks.begin() + n_rep, ranks.end());
    else
      SAM_XG(xg_shuffle_cache_ranks(ranks.data() + n_rep, n_cached - n_rep));  // cuda_cache_manager_host.cc:162-171
  }
  if (!ics22) {
    std::vector<uint32_t> table(V, XG_EMPTY_KEY);
    for (size_t r = 0; r < n_cached; ++r) table[ranks[r]] = (uint32_t)r;
    e.store.cache_table = to_device(e, table.data(), V);
  }
  const void *h_feat = pin_host(ds.feat, T * row_bytes);
  e.store.num_part = P;
  auto gather_rows_to_device = [&](const uint32_t *ids, size_t n) -> void * {
    // CPUExtract + H2D in the reference; here the GPU gathers the rows straight out of the host mapping
    void *d_rows = dev_alloc(e, std::max<size_t>(n * row_bytes, 16));
    if (n) {
      uint32_t *d_ids = nullptr;
      SAM_CUDA(cudaMalloc(&d_ids, n * sizeof(uint32_t)));
      SAM_CUDA(cudaMemcpyAsync(d_ids, ids, n * sizeof(uint32_t), cudaMemcpyHostToDevice, e.stream));
      // CPUMockExtract when the table is a mock (dist_graph.cu:508-515), CPUExtract otherwise
      SAM_XG(xg_mock_extract(d_rows, h_feat, d_ids, n, row_bytes, ds.empty_feat, e.stream));
      SAM_CUDA(cudaStreamSynchronize(e.stream));
      cudaFree(d_ids);
    }
    return d_rows;
  };
  if (ics22) {
    // ICS'22 clique placement (comparison baseline; ICS22SongDistGraph ctor + FeatureLoad, cuda/ics22_song_dist_feature.cc:
    // 169-305, and the compact map of cuda_cache_manager_host.cc:172-222).  The reference solves for every GPU in the
    // parent before the fork; here each worker ranks the nodes by in-degree on its own GPU (same counts -> same order in
    // every worker) and derives only its own view of the placement.
    const uint32_t c = (uint32_t)cfg.clique_size, bw = (uint32_t)cfg.ics22_compact_bitwidth;
    SAM_CHECK_MSG(P % c == 0, "num_worker must be a multiple of clique_size");
    SAM_CHECK_MSG(P <= (1u << bw), "ics22 compact map: too many workers for ics22_compact_bitwidth");
    std::vector<uint32_t> rank_nodes(V), rank_weight(V);
    {
      uint32_t *d_freq = nullptr, *d_rank = nullptr, *d_rfreq = nullptr, *d_chunk = nullptr;
      const size_t chunk = std::min<size_t>(std::max<size_t>(ds.num_edge, 1), (size_t)64 << 20);
      SAM_CUDA(cudaMalloc(&d_freq, V * sizeof(uint32_t)));
      SAM_CUDA(cudaMalloc(&d_rank, V * sizeof(uint32_t)));
      SAM_CUDA(cudaMalloc(&d_rfreq, V * sizeof(uint32_t)));
      SAM_CUDA(cudaMalloc(&d_chunk, chunk * sizeof(uint32_t)));
      SAM_CUDA(cudaMemsetAsync(d_freq, 0, V * sizeof(uint32_t), e.stream));
      for (size_t off = 0; off < ds.num_edge; off += chunk) {  // weight = in-degree (:196-213)
        const size_t n = std::min(chunk, ds.num_edge - off);
        SAM_CUDA(cudaMemcpyAsync(d_chunk, ds.indices + off, n * sizeof(uint32_t), cudaMemcpyHostToDevice, e.stream));
        SAM_XG(xg_presample_accumulate(d_freq, (uint32_t)V, d_chunk, n, nullptr, e.stream));
        SAM_CUDA(cudaStreamSynchronize(e.stream));
      }
      SAM_XG(xg_presample_rank(d_freq, (uint32_t)V, d_rank, d_rfreq, e.stream));  // (weight, id) descending, as :48-56
      SAM_CUDA(cudaMemcpy(rank_nodes.data(), d_rank, V * sizeof(uint32_t), cudaMemcpyDeviceToHost));
      SAM_CUDA(cudaMemcpy(rank_weight.data(), d_rfreq, V * sizeof(uint32_t), cudaMemcpyDeviceToHost));
      cudaFree(d_chunk); cudaFree(d_rfreq); cudaFree(d_rank); cudaFree(d_freq);
    }
    const int w = e.worker_id;
    std::vector<uint32_t> table(V), slots(std::max<size_t>(n_cached, 1));
    uint32_t total = 0;
    SAM_XG(xg_ics22_placement(rank_nodes.data(), rank_weight.data(), (uint32_t)V, (uint32_t)n_cached, P, c,
                              cfg.ics22_song_alpha, (uint32_t)w, bw, table.data(), slots.data(), &total));
    e.store.cache_table = to_device(e, table.data(), V);
    e.store.compact_bitwidth = bw;
    e.store.num_part = P;
    e.store.local_part_mask = 1u << w;
    void *d_mine = gather_rows_to_device(slots.data(), n_cached);
    SAM_XG(xg_ipc_export(d_mine, e.shared->feat[w][w]));
    barrier(e);
    const uint32_t first = (uint32_t)w / c * c;
    for (uint32_t g = first; g < first + c; ++g) {  // only the clique's caches are ever named by the map (:258-274)
      if ((int)g == w) {
        e.store.part[g] = d_mine;
      } else {
        void *a = nullptr;
        SAM_XG(xg_ipc_import(e.shared->feat[g][g], &a));
        e.store.part[g] = a;
      }
    }
    barrier(e);
    if (w == 0)
      std::fprintf(stderr, "[samgraph] ics22 clique placement: %zu rows per GPU, %u distinct nodes per clique of %u (alpha %.3f)\n",
                   n_cached, total, c, cfg.ics22_song_alpha);
    return;
  }
  if (n_rep) {
    e.store.replica = gather_rows_to_device(ranks.data(), n_rep);
    e.store.num_replicated = (uint32_t)n_rep;
    void *words = dev_alloc(e, ((V + 31) / 32) * 8);
    SAM_XG(xg_build_replica_words(e.store.cache_table, (uint32_t)V, (uint32_t)n_rep, words, &e.store.replica_words_cover_all,
                                  e.stream));
    e.store.replica_words = words;
  }
  const int w = e.worker_id;
  std::vector<void *> mine(P, nullptr);
  const int my_parts = parted ? e.shared->num_parts[w] : 1;
  for (int k = 0; k < my_parts; ++k) {
    const uint32_t p = parted ? (uint32_t)e.shared->part_ids[w][k] : 0u;
    std::vector<uint32_t> ids;
    for (size_t r = n_rep + p; r < n_cached; r += P) ids.push_back(ranks[r]);  // rank r' = r - n_rep -> part r' % P, row r' / P
    void *d_part = gather_rows_to_device(ids.data(), ids.size());
    mine[p] = d_part;
    e.store.local_part_mask |= 1u << p;
    if (parted) SAM_XG(xg_ipc_export(d_part, e.shared->feat[w][p]));
  }
  if (!parted) {
    e.store.part[0] = mine[0];
    return;
  }
  barrier(e);
  for (uint32_t p = 0; p < P; ++p) {
    const int owner = e.shared->access[w][p];
    if (owner == w) {
      e.store.part[p] = mine[p];
    } else {
      void *a = nullptr;
      SAM_XG(xg_ipc_import(e.shared->feat[owner][p], &a));
      e.store.part[p] = a;
    }
  }
  barrier(e);
}

}  // namespace sam

namespace sam {

void create_sampler(Engine &e) {
  if (e.sampler) return;
  const RunConfig &cfg = g_cfg;
  const Dataset &ds = e.ds;
  xg_sampler_config sc{};
  sc.num_layer = (uint32_t)cfg.num_layer;
  SAM_CHECK(cfg.num_layer >= 1 && cfg.num_layer <= XG_MAX_LAYERS && cfg.fanout.size() >= cfg.num_layer);
  for (size_t l = 0; l < cfg.num_layer; ++l) sc.fanout[l] = (uint32_t)cfg.fanout[l];
  sc.batch_size = (uint32_t)cfg.batch_size;
  sc.sample_type = cfg.sample_type;
  sc.num_slots = 6;  // current + 2 pooled + in flight + two still referenced by the caller's tensors
  sc.feat_dim = (uint32_t)ds.feat_dim;
  sc.feat_dtype = ds.feat_dtype;
  sc.seed = 0x58474E4Eull;  // reproducible, unlike the reference's wall-clock seed (cuda_random_states.cu:106-108)
  if (const char *s = std::getenv("SAMGRAPH_SEED")) sc.seed = std::strtoull(s, nullptr, 0);
  if (const char *s = std::getenv("SAMGRAPH_SANITY_CHECK")) e.sanity_check = !std::strcmp(s, "ON") || !std::strcmp(s, "1");  // IsEnvSet, common.cc:517-525
  sc.stream = e.stream;
  sc.random_walk_length = (uint32_t)cfg.random_walk_length;
  sc.num_random_walk = (uint32_t)cfg.num_random_walk;
  sc.random_walk_restart_prob = cfg.random_walk_restart_prob;
  sc.prob_table = e.d_prob;
  sc.alias_table = e.d_alias;
  sc.prob_prefix_table = e.d_prefix;
  SAM_XG(xg_sampler_create(&e.sampler, &e.graph, e.trainer_ready ? &e.store : nullptr, e.trainer_ready ? e.d_label : nullptr,
                           &sc, e.device));
}

// RunSampleSubLoopOnce + RunCacheDataCopySubLoopOnce for one Task (dist_loops_arch6.cc:55-199).
// Returns false once every epoch has been produced.
// SAMGRAPH_SANITY_CHECK=1 (constant.cc:58; cuda_shuffler.cc:145-152 -> cuda_sanity_check.cu:28-88): every batch's seeds are
// valid ids and - outside the aligned shuffler's wrap-around padding - unique within the epoch; every list is free of
// kEmptyKey; every edge endpoint is inside its block.  A violation aborts like the reference's device assert.
void sanity_check_batch(Engine &e, const xg_batch_view &v, size_t epoch) {
  const uint32_t V = (uint32_t)e.ds.num_node;
  if (!e.d_sanity) {
    SAM_CUDA(cudaMalloc(&e.d_sanity, sizeof(uint64_t)));
    SAM_CUDA(cudaMalloc(&e.d_epoch_map, (size_t)V * sizeof(uint32_t)));
    e.sanity_epoch = (size_t)-1;
  }
  if (e.sanity_epoch != epoch) {
    SAM_CUDA(cudaMemsetAsync(e.d_epoch_map, 0, (size_t)V * sizeof(uint32_t), e.stream));
    e.sanity_epoch = epoch;
  }
  SAM_CUDA(cudaMemsetAsync(e.d_sanity, 0, sizeof(uint64_t), e.stream));
  // duplicates are legal only as the shuffler's padding (<= num_worker - 1 per epoch, dist_shuffler_aligned.cc:53-55),
  // so the epoch map is only armed when nothing was padded
  const bool padded = (e.ds.num_train % g_cfg.num_worker) != 0;
  SAM_XG(xg_sanity_check_ids(v.output_nodes, v.num_output, V, padded ? nullptr : e.d_epoch_map, e.d_sanity, e.stream));
  SAM_XG(xg_sanity_check_ids(v.input_nodes, v.num_input, V, nullptr, e.d_sanity, e.stream));
  for (uint32_t l = 0; l < v.num_layer; ++l)
    SAM_XG(xg_sanity_check_edges(v.row[l], v.col[l], v.num_edge[l], (uint32_t)v.num_src[l], (uint32_t)v.num_dst[l], e.d_sanity,
                                 e.stream));
  uint64_t bad = 0;
  SAM_CUDA(cudaMemcpyAsync(&bad, e.d_sanity, sizeof(bad), cudaMemcpyDeviceToHost, e.stream));
  SAM_CUDA(cudaStreamSynchronize(e.stream));
  SAM_CHECK_MSG(bad == 0, "SAMGRAPH_SANITY_CHECK: " + std::to_string(bad) + " violation(s) in batch " + std::to_string(v.key));
}

constexpr size_t kInFlight = 2;  // Tasks kept inside the sampler engine: one being sampled + the extract phase of the one before

// DoShuffle + enqueue: takes the next batch of seeds and hands it to the sampler engine WITHOUT waiting for it.
// Returns false once every epoch has been handed out.
bool submit_one(Engine &e) {
  SAM_CHECK_MSG(e.sampler_ready, "sample_init / init has not run");
  create_sampler(e);
  if (e.shuffler_done) return false;
  PendingBatch p;
  p.t_submit = now_s();
  const uint32_t *d_seeds = nullptr;
  size_t n = 0, local_step = 0;
  if (!e.shuffler->get_batch(e.stream, &d_seeds, &n, &p.epoch, &local_step)) {
    e.shuffler_done = true;
    return false;
  }
  p.t_shuffle = now_s() - p.t_submit;
  // Engine::GetBatchKey(epoch, step) with the blocked global step (dist_shuffler_aligned.h:38-40, engine.h:50-54)
  const uint64_t key = p.epoch * e.num_step + e.shuffler->global_step_offset + local_step;
  p.holder = std::make_shared<BatchHolder>();
  p.holder->key = key;
  // every slot may still be referenced by tensors the caller holds (from_blob semantics): block until one dies
  // (BatchHolder's destructor signals slot_cv); 60 s without any release means the caller leaks batches
  const double t_wait0 = now_s();
  for (;;) {
    uint64_t seen;
    {
      std::lock_guard<std::mutex> lk(e.slot_mu);
      seen = e.slot_gen;
    }
    const int rc = xg_sampler_submit(e.sampler, d_seeds, n, 1, key, &p.holder->batch);
    if (rc == XG_OK) break;
    SAM_CHECK_MSG(rc == XG_ERR_CAPACITY, std::string("xg_sampler_submit: ") + xg_last_error());
    SAM_CHECK_MSG(now_s() - t_wait0 < 60.0, "every batch slot is still referenced by caller tensors after 60 s: " +
                                                std::string(xg_last_error()));
    std::unique_lock<std::mutex> lk(e.slot_mu);
    e.slot_cv.wait_for(lk, std::chrono::milliseconds(200), [&] { return e.slot_gen != seen; });
  }
  e.inflight.push_back(std::move(p));
  return true;
}

void send_task(Engine &e, const xg_batch_view &v, size_t epoch, double sample_time);

// Waits for the OLDEST Task in flight, logs it and hands it on (GraphPool::Submit, or the arch5 queue).
void finish_one(Engine &e) {
  SAM_CHECK(!e.inflight.empty());
  PendingBatch p = std::move(e.inflight.front());
  e.inflight.pop_front();
  BatchPtr holder = std::move(p.holder);
  SAM_XG(xg_batch_wait(holder->batch, &holder->view));
  const double t_done = now_s();
  // with Tasks overlapping inside the engine, a Task's share of the epoch is the time since the previous one finished
  const double t_total = t_done - std::max(p.t_submit, e.last_finish);
  e.last_finish = t_done;
  const uint64_t key = holder->key;
  const size_t epoch = p.epoch;
  if (e.sanity_check) sanity_check_batch(e, holder->view, epoch);
  double edges = 0;
  for (uint32_t l = 0; l < holder->view.num_layer; ++l) edges += (double)holder->view.num_edge[l];
  const double feat_bytes = (double)holder->view.num_input * e.ds.feat_row_bytes();
  g_prof.log_step(key, kLogL1NumSample, edges, false);
  g_prof.log_step(key, kLogL1NumNode, (double)holder->view.num_input, false);
  g_prof.log_step(key, kLogL1SampleTime, t_total, false);  // sample and extract are one fused chain: all of it
  g_prof.log_step(key, kLogL1SampleTotalTime, t_total, false);
  g_prof.log_step(key, kLogL2ShuffleTime, p.t_shuffle, false);
  g_prof.log_step(key, kLogL2CoreSampleTime, std::max(0.0, t_total - p.t_shuffle), false);
  g_prof.log_step(key, kLogL1CopyTime, 0.0, false);
  g_prof.log_step(key, kLogL1FeatureBytes, feat_bytes, false);
  g_prof.log_step(key, kLogL1LabelBytes, (double)holder->view.num_output * 8, false);
  g_prof.log_epoch_add(epoch, kLogEpochSampleTime, t_total);
  g_prof.log_epoch_add(epoch, kLogEpochSampleTotalTime, t_total);
  g_prof.log_epoch_add(epoch, kLogEpochNumSample, edges);
  g_prof.log_epoch_add(epoch, kLogEpochFeatureBytes, feat_bytes);
  if (e.role == 1) {  // arch5 sampler: the Task goes to a trainer process (RunArch5LoopsOnce / DoGPUSample + Send)
    send_task(e, holder->view, epoch, t_total);
    return;            // holder dies here: the slot is free again, the payload lives in the stage buffer
  }
  {
    std::lock_guard<std::mutex> lk(e.mu);
    e.pool.push_back(holder);  // GraphPool::Submit
  }
  e.cv.notify_all();
}

// RunSampleSubLoopOnce + RunCacheDataCopySubLoopOnce for one Task (dist_loops_arch6.cc:55-199): one more Task reaches
// the pool.  The sampler engine is kept kInFlight Tasks deep, as the reference's background SampleCopySubLoop threads
// keep its stages busy (dist_loops_arch6.cc:221-243, graph_pool.cc:31-49): Task k+1 is enqueued BEFORE Task k is
// waited for, so its sample phase overlaps Task k's gather - and, for sample_once callers, the caller's training step.
// Batch keys, order and contents do not depend on this (the RNG is keyed by the batch key).
// Returns false once every epoch has been produced.
bool produce_one(Engine &e) {
  while (e.inflight.size() < kInFlight && submit_one(e)) {
  }
  if (e.inflight.empty()) return false;
  finish_one(e);
  return true;
}

// ---------------------------------------------------------------------------------------------------
// arch5 task hand-over (see TaskQueueBlock).  Sampler side = MessageTaskQueue::Send (task_queue.cc:332-367), trainer side
// = MessageTaskQueue::Recv + DoGPUFeatureExtract/DoGPULabelExtract... of RunArch5 extract loops (dist_loops_arch5.cc).
bool sem_wait_or_shutdown(Engine &e, sem_t *sem) {
  // SAMGRAPH_QUEUE_TIMEOUT seconds (default 600; covers a trainer still loading its cache) without the other side posting = it is gone (crashed child, lost Task):
  // fail with a message like any violated CHECK instead of hanging every process of the job
  static const double limit = [] {
    const char *s = std::getenv("SAMGRAPH_QUEUE_TIMEOUT");
    return s ? std::atof(s) : 600.0;
  }();
  const double t_begin = now_s();
  for (;;) {
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    ts.tv_nsec += 50 * 1000 * 1000;
    if (ts.tv_nsec >= 1000000000L) { ts.tv_sec += 1; ts.tv_nsec -= 1000000000L; }
    if (sem_timedwait(sem, &ts) == 0) return true;
    SAM_CHECK_MSG(errno == ETIMEDOUT || errno == EINTR, "sem_timedwait failed");
    SAM_CHECK_MSG(limit <= 0 || now_s() - t_begin < limit, "arch5 task queue: the other side has not posted for SAMGRAPH_QUEUE_TIMEOUT seconds");
    std::lock_guard<std::mutex> lk(e.mu);
    if (e.shutdown) return false;
  }
}

inline uint64_t align256(uint64_t x) { return (x + 255) & ~255ull; }

// Largest packed Task this sampler can produce: the caps xg_sampler_create derives from batch size and fanouts
// (PredictNumNodes, common.cc:488-497, clipped to num_node + batch), every piece 256-byte aligned.
size_t worst_case_task_bytes(const Engine &e) {
  const RunConfig &cfg = g_cfg;
  const size_t B = cfg.batch_size, L = cfg.num_layer;
  size_t limit = B;
  for (size_t i = L; i-- > 0;) limit += limit * cfg.fanout[i];
  if (limit > e.ds.num_node + B) limit = e.ds.num_node + B;
  size_t cur = B, bytes = align256(B * 4);
  for (size_t i = L; i-- > 0;) {
    const size_t cap_edge = cur * cfg.fanout[i];
    bytes += 3 * align256(cap_edge * 4) + align256((cur + 1) * 4);  // row, col, data, csc_indptr
    cur = std::min(cur + cap_edge, limit);
  }
  return bytes + align256(cur * 4) + 4096;
}

void send_task(Engine &e, const xg_batch_view &v, size_t epoch, double sample_time) {
  SAM_CHECK(e.queue != nullptr);
  TaskQueueBlock &q = *e.queue;
  const double t0 = now_s();
  const int s = e.worker_id;
  TaskDesc d;
  std::memset(&d, 0, sizeof(d));
  d.key = v.key; d.epoch = epoch; d.num_layer = v.num_layer; d.num_input = v.num_input; d.num_output = v.num_output;
  d.sampler = (uint32_t)s; d.sample_time = sample_time;
  // layout: input_nodes | output_nodes | per layer row | col | [data] | [csc_indptr], each 256-byte aligned
  struct Piece { const void *src; uint64_t off, bytes; };
  std::vector<Piece> pieces;
  uint64_t at = 0;
  auto place = [&](const void *src, uint64_t count) -> uint64_t {
    if (src == nullptr) return kNone;
    const uint64_t off = at;
    pieces.push_back({src, off, count * sizeof(uint32_t)});
    at = align256(at + count * sizeof(uint32_t));
    return off;
  };
  d.off_input = place(v.input_nodes, v.num_input);
  d.off_output = place(v.output_nodes, v.num_output);
  for (uint32_t l = 0; l < v.num_layer; ++l) {
    d.num_src[l] = v.num_src[l]; d.num_dst[l] = v.num_dst[l]; d.num_edge[l] = v.num_edge[l];
    d.off_row[l] = place(v.row[l], v.num_edge[l]);
    d.off_col[l] = place(v.col[l], v.num_edge[l]);
    d.off_data[l] = place(v.data[l], v.num_edge[l]);
    d.off_csc[l] = place(v.csc_indptr[l], v.num_dst[l] + 1);
  }
  d.bytes = std::max<uint64_t>(at, 256);
  // a stage buffer of this sampler that no trainer is still pulling from
  const unsigned k = e.next_stage++ % kStagePerSampler;
  if (!sem_wait_or_shutdown(e, &q.stage_free[s][k])) return;
  Engine::Stage &st = e.stages[k];
  if (st.ptr == nullptr) {
    // Allocated ONCE, at the worst-case packed size of a Task (the sampler's own buffer caps), and never freed or
    // re-exported while the process lives: trainers keep the IPC mapping of a stage open for as long as they like
    // (peer_stage[][]), and cudaFree of an exported region before every importer has closed it is undefined
    // behaviour.  (r01 grew a stage on demand and freed the old block.)
    st.cap = worst_case_task_bytes(e);
    SAM_XG(xg_device_malloc(&st.ptr, st.cap));  // 2 MiB granular: peer reads of smaller blocks crawl
    SAM_XG(xg_ipc_export(st.ptr, q.stage_ipc[s][k]));
    __atomic_add_fetch(&q.stage_gen[s][k], 1u, __ATOMIC_SEQ_CST);
  }
  SAM_CHECK_MSG(d.bytes <= st.cap, "a packed Task is larger than the worst case computed from batch size and fanouts");
  for (const Piece &p : pieces)
    if (p.bytes)
      SAM_CUDA(cudaMemcpyAsync(static_cast<char *>(st.ptr) + p.off, p.src, p.bytes, cudaMemcpyDeviceToDevice, e.stream));
  SAM_CUDA(cudaStreamSynchronize(e.stream));
  d.stage = k;
  d.stage_gen = __atomic_load_n(&q.stage_gen[s][k], __ATOMIC_SEQ_CST);
  const uint64_t ticket = __atomic_fetch_add(&q.send_cnt, 1ull, __ATOMIC_SEQ_CST);  // MQ_MetaData::Put
  const int slot = (int)(ticket % kQueueDepth);
  if (!sem_wait_or_shutdown(e, &q.vacant[slot])) return;  // ReleaseWait
  q.desc[slot] = d;
  SAM_CHECK(sem_post(&q.filled[slot]) == 0);  // SemPost
  const double t_send = now_s() - t0;
  g_prof.log_step(v.key, kLogL1SendTime, t_send, false);
  g_prof.log_epoch_add(epoch, kLogEpochSampleSendTime, t_send);
}

Engine::Landing *acquire_landing(Engine &e) {
  for (int spins = 0;; ++spins) {
    {
      std::lock_guard<std::mutex> lk(e.mu);
      for (auto &l : e.landings)
        if (!l->busy) { l->busy = true; return l.get(); }
      if (e.landings.size() < 8) {  // current + 2 pooled + 2 in the extract thread + a few held by caller tensors
        e.landings.emplace_back(new Engine::Landing());
        e.landings.back()->busy = true;
        return e.landings.back().get();
      }
      if (e.shutdown) return nullptr;
    }
    SAM_CHECK_MSG(spins < 300000, "every landing buffer is still referenced by caller tensors");
    std::this_thread::sleep_for(std::chrono::microseconds(100));
  }
}

void grow(void **p, size_t *cap, size_t need) {
  if (*cap >= need) return;
  if (*p) SAM_XG(xg_device_free(*p));
  *cap = need + need / 4;
  SAM_XG(xg_device_malloc(p, *cap));
}

// one Task: dequeue -> peer copy of the packed topology -> feature + label gather on THIS (trainer) GPU -> GraphPool.
// Two halves, so that the extract thread can have Task k+1's peer copy in flight (copy stream) while Task k's rows
// are being gathered (engine stream).
struct PendingTask {
  TaskDesc d;
  Engine::Landing *land = nullptr;
  cudaEvent_t copied = nullptr;
  double t0 = 0, t_recv = 0;
  bool prefetched = false;  // begun while the previous Task was still being gathered
};

std::unique_ptr<PendingTask> recv_begin(Engine &e) {
  SAM_CHECK_MSG(e.queue != nullptr && e.trainer_ready, "train_init has not run");
  TaskQueueBlock &q = *e.queue;
  const double t0 = now_s();
  const uint64_t ticket = __atomic_fetch_add(&q.recv_cnt, 1ull, __ATOMIC_SEQ_CST);  // MQ_MetaData::Get
  const int slot = (int)(ticket % kQueueDepth);
  if (!sem_wait_or_shutdown(e, &q.filled[slot])) return nullptr;  // SemWait
  const TaskDesc d = q.desc[slot];
  SAM_CHECK(sem_post(&q.vacant[slot]) == 0);  // the descriptor is copied: the ring entry is free (ReleasePost)
  Engine::PeerStage &ps = e.peer_stage[d.sampler][d.stage];
  if (ps.ptr == nullptr || ps.gen != d.stage_gen) {
    if (ps.ptr) SAM_XG(xg_ipc_close(ps.ptr));
    SAM_XG(xg_ipc_import(q.stage_ipc[d.sampler][d.stage], &ps.ptr));
    ps.gen = d.stage_gen;
  }
  Engine::Landing *land = acquire_landing(e);
  if (!land) return nullptr;
  const size_t row_bytes = e.ds.feat_row_bytes();
  grow(&land->xfer, &land->cap_xfer, d.bytes);
  grow(&land->feat, &land->cap_feat, std::max<size_t>((size_t)d.num_input * row_bytes, 256));
  grow(reinterpret_cast<void **>(&land->label), &land->cap_label, std::max<size_t>((size_t)d.num_output * 8, 256));
  if (!e.copy_stream) SAM_CUDA(cudaStreamCreateWithFlags(&e.copy_stream, cudaStreamNonBlocking));
  std::unique_ptr<PendingTask> p(new PendingTask());
  p->d = d; p->land = land; p->t0 = t0;
  SAM_CUDA(cudaEventCreateWithFlags(&p->copied, cudaEventDisableTiming));
  // the whole packed Task in one copy, sampler HBM -> trainer HBM (NVLink when they are different GPUs)
  SAM_CUDA(cudaMemcpyAsync(land->xfer, ps.ptr, d.bytes, cudaMemcpyDefault, e.copy_stream));
  SAM_CUDA(cudaEventRecord(p->copied, e.copy_stream));
  return p;
}

// Wait for the Task's peer copy and hand its stage buffer back to the sampler.  RULE (deadlock freedom): a trainer must not
// block on the queue - wait for a Task that is still to be produced - while one of its Tasks keeps a stage: the sampler
// that has to produce that Task may be waiting for exactly this stage (it cycles through kStagePerSampler of them, and
// another trainer can drain the ones in between).  So stages are released before every blocking dequeue.
void release_stage(Engine &e, PendingTask &p) {
  if (p.copied == nullptr) return;  // already released
  const double t_wait = now_s();
  SAM_CUDA(cudaEventSynchronize(p.copied));
  SAM_CUDA(cudaEventDestroy(p.copied));
  p.copied = nullptr;
  SAM_CHECK(sem_post(&e.queue->stage_free[p.d.sampler][p.d.stage]) == 0);  // the sampler may overwrite its stage buffer now
  // time this Task kept the trainer waiting (dequeue + copy when nothing overlapped it, ~0 when prefetched)
  p.t_recv = p.prefetched ? now_s() - t_wait : now_s() - p.t0;
}

bool recv_finish(Engine &e, std::unique_ptr<PendingTask> p) {
  if (!p) return false;
  release_stage(e, *p);
  const TaskDesc &d = p->d;
  Engine::Landing *land = p->land;
  const size_t row_bytes = e.ds.feat_row_bytes();
  const double t_recv = p->t_recv;
  auto at = [&](uint64_t off) -> const uint32_t * {
    return off == kNone ? nullptr : reinterpret_cast<const uint32_t *>(static_cast<const char *>(land->xfer) + off);
  };
  auto holder = std::make_shared<BatchHolder>();
  holder->key = d.key;
  xg_batch_view &v = holder->view;
  std::memset(&v, 0, sizeof(v));
  v.key = d.key; v.num_layer = d.num_layer; v.num_input = d.num_input; v.num_output = d.num_output;
  v.input_nodes = at(d.off_input);
  v.output_nodes = at(d.off_output);
  for (uint32_t l = 0; l < d.num_layer; ++l) {
    v.num_src[l] = d.num_src[l]; v.num_dst[l] = d.num_dst[l]; v.num_edge[l] = d.num_edge[l];
    v.row[l] = at(d.off_row[l]); v.col[l] = at(d.off_col[l]); v.data[l] = at(d.off_data[l]); v.csc_indptr[l] = at(d.off_csc[l]);
  }
  Engine *ep = &e;
  holder->on_release = [ep, land] {
    std::lock_guard<std::mutex> lk(ep->mu);
    land->busy = false;
  };
  // DoGPUFeatureExtract / DoGPULabelExtract on the trainer GPU (dist/dist_loops.cc:843-974): one routed gather launch
  const double t1 = now_s();
  if (d.num_input) SAM_XG(xg_extract_ggms(&e.store, v.input_nodes, d.num_input, land->feat, nullptr, e.stream));
  if (d.num_output) SAM_XG(xg_extract(land->label, e.d_label, v.output_nodes, d.num_output, 1, XG_I64, e.stream));
  SAM_CUDA(cudaStreamSynchronize(e.stream));
  v.input_feat = land->feat;
  v.output_label = land->label;
  const double t_copy = now_s() - t1;
  double edges = 0;
  for (uint32_t l = 0; l < d.num_layer; ++l) edges += (double)d.num_edge[l];
  const double feat_bytes = (double)d.num_input * row_bytes;
  g_prof.log_step(d.key, kLogL1NumSample, edges, false);
  g_prof.log_step(d.key, kLogL1NumNode, (double)d.num_input, false);
  g_prof.log_step(d.key, kLogL1SampleTime, d.sample_time, false);
  g_prof.log_step(d.key, kLogL1RecvTime, t_recv, false);
  g_prof.log_step(d.key, kLogL1CopyTime, t_copy, false);
  g_prof.log_step(d.key, kLogL1FeatureBytes, feat_bytes, false);
  g_prof.log_step(d.key, kLogL1LabelBytes, (double)d.num_output * 8, false);
  g_prof.log_epoch_add(d.epoch, kLogEpochSampleTime, d.sample_time);
  g_prof.log_epoch_add(d.epoch, kLogEpochCopyTime, t_recv + t_copy);
  g_prof.log_epoch_add(d.epoch, kLogEpochNumSample, edges);
  g_prof.log_epoch_add(d.epoch, kLogEpochFeatureBytes, feat_bytes);
  {
    std::lock_guard<std::mutex> lk(e.mu);
    e.pool.push_back(holder);
  }
  e.cv.notify_all();
  return true;
}

bool recv_task(Engine &e) { return recv_finish(e, recv_begin(e)); }

void common_device_init(Engine &e, int device) {
  e.device = device;
  SAM_CUDA(cudaSetDevice(device));
  SAM_CUDA(cudaStreamCreateWithFlags(&e.stream, cudaStreamNonBlocking));
}

// dist::PreSampler::DoPreSample + GetRankNode (dist/pre_sampler.cc:76-152, driven from dist_engine.cc:455-466): worker 0
// samples `presample_epoch` epochs of the whole train set (no feature extraction), counts on the GPU how many batches
// need each node, ranks; the other workers wait at the barrier.
void presample_ranking(Engine &e) {
  const RunConfig &cfg = g_cfg;
  Dataset &ds = e.ds;
  if (cfg.cache_policy != kCacheByPreSample || cfg.cache_percentage <= 0.0 || ds.ranking_rw == nullptr) return;
  const double t0 = now_s();
  if (e.worker_id == 0) {
    xg_sampler_config sc{};
    sc.num_layer = (uint32_t)cfg.num_layer;
    for (size_t l = 0; l < cfg.num_layer; ++l) sc.fanout[l] = (uint32_t)cfg.fanout[l];
    sc.batch_size = (uint32_t)cfg.batch_size;
    sc.sample_type = cfg.sample_type;
    sc.num_slots = 2;
    sc.seed = 0x50524553ull;  // "PRES": its own stream of draws, independent of the training batches
    sc.stream = e.stream;
    sc.random_walk_length = (uint32_t)cfg.random_walk_length;
    sc.num_random_walk = (uint32_t)cfg.num_random_walk;
    sc.random_walk_restart_prob = cfg.random_walk_restart_prob;
    sc.prob_table = e.d_prob;
    sc.alias_table = e.d_alias;
    sc.prob_prefix_table = e.d_prefix;
    xg_sampler *ps = nullptr;
    SAM_XG(xg_sampler_create(&ps, &e.graph, nullptr, nullptr, &sc, e.device));
    uint32_t *d_freq = nullptr, *d_rank = nullptr;
    SAM_CUDA(cudaMalloc(&d_freq, ds.num_node * sizeof(uint32_t)));
    SAM_CUDA(cudaMalloc(&d_rank, ds.num_node * sizeof(uint32_t)));
    SAM_CUDA(cudaMemsetAsync(d_freq, 0, ds.num_node * sizeof(uint32_t), e.stream));
    std::vector<uint32_t> order(ds.train_set, ds.train_set + ds.num_train);
    uint64_t key = 0;
    for (int ep = 0; ep < cfg.presample_epoch; ++ep) {
      AlignedShuffler::shuffle_epoch(order, (size_t)ep);
      for (size_t o = 0; o < order.size(); o += cfg.batch_size) {
        const size_t n = std::min(cfg.batch_size, order.size() - o);
        xg_batch *b = nullptr;
        xg_batch_view v;
        SAM_XG(xg_sampler_submit(ps, order.data() + o, n, 0, key++, &b));
        SAM_XG(xg_batch_wait(b, &v));
        SAM_XG(xg_presample_accumulate(d_freq, (uint32_t)ds.num_node, v.input_nodes, v.num_input, nullptr, e.stream));
        SAM_CUDA(cudaStreamSynchronize(e.stream));
        SAM_XG(xg_batch_release(b));
      }
    }
    SAM_XG(xg_presample_rank(d_freq, (uint32_t)ds.num_node, d_rank, nullptr, e.stream));
    SAM_CUDA(cudaMemcpy(ds.ranking_rw, d_rank, ds.num_node * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    cudaFree(d_freq);
    cudaFree(d_rank);
    SAM_XG(xg_sampler_destroy(ps));
  }
  barrier(e);
  g_prof.init[kLogInitL2Presample] = now_s() - t0;
}

void sampler_side_init(Engine &e) {
  const double t0 = now_s();
  build_topology(e);
  if (g_cfg.sample_type == XG_WEIGHTED_KHOP || g_cfg.sample_type == XG_WEIGHTED_KHOP_HASH_DEDUP) {
    e.d_prob = to_device(e, e.ds.prob, e.ds.num_edge);
    e.d_alias = to_device(e, e.ds.alias, e.ds.num_edge);
  } else if (g_cfg.sample_type == XG_WEIGHTED_KHOP_PREFIX) {
    e.d_prefix = to_device(e, e.ds.prefix, e.ds.num_edge);
  }
  e.shuffler.reset(new AlignedShuffler(e.ds.train_set, e.ds.num_train, g_cfg.num_epoch, g_cfg.batch_size,
                                       (size_t)e.worker_id, g_cfg.num_worker));
  SAM_CHECK(e.shuffler->num_local_step == e.num_local_step);
  presample_ranking(e);
  e.sampler_ready = true;
  g_prof.init[kLogInitL1Sampler] = now_s() - t0;
}

void trainer_side_init(Engine &e) {
  const double t0 = now_s();
  build_feature_store(e);
  e.d_label = const_cast<int64_t *>(to_device(e, e.ds.label, e.ds.num_node));  // labels: 8 B x V, kept in HBM
  e.trainer_ready = true;
  g_prof.init[kLogInitL1Trainer] = now_s() - t0;
  g_prof.init[kLogInitL2BuildCache] = now_s() - t0;
}

void create_engine(bool shared_dataset) {
  SAM_CHECK(g_cfg.is_configured);
  SAM_CHECK_MSG(g_engine == nullptr, "engine already created");
  g_engine = new Engine();
  Engine &e = *g_engine;
  const double t0 = now_s();
  load_dataset(e.ds, g_cfg, shared_dataset);
  e.dataset_loaded = true;
  g_prof.init[kLogInitL2LoadDataset] = now_s() - t0;
  // dist_engine.cc:139-148: steps are counted on the padded, per-worker slice
  const size_t W = g_cfg.num_worker;
  e.num_local_step = ((e.ds.num_train + W - 1) / W + g_cfg.batch_size - 1) / g_cfg.batch_size;
  e.num_step = e.num_local_step * W;
}

}  // namespace sam

using namespace sam;

extern "C" {

// operation.cc:49-326
void samgraph_config(const char **keys, const char **vals, const size_t n) {
  RunConfig &c = g_cfg;
  SAM_CHECK(!c.is_configured);
  std::unordered_map<std::string, std::string> m;
  for (size_t i = 0; i < n; ++i) m[keys[i]] = vals[i];
  for (const char *need : {"dataset_path", "_arch", "_sample_type", "batch_size", "num_epoch", "_cache_policy",
                           "cache_percentage", "max_sampling_jobs", "max_copying_jobs", "omp_thread_num", "num_layer",
                           "num_hidden", "lr", "dropout"})
    SAM_CHECK_MSG(m.count(need) > 0, std::string("missing config key ") + need);  // operation.cc:68-81
  c.raw = m;
  c.dataset_path = m["dataset_path"];
  c.run_arch = std::stoi(m["_arch"]);
  c.sample_type = std::stoi(m["_sample_type"]);
  c.batch_size = std::stoull(m["batch_size"]);
  c.num_epoch = std::stoull(m["num_epoch"]);
  c.cache_policy = std::stoi(m["_cache_policy"]);
  c.cache_percentage = std::stod(m["cache_percentage"]);
  if (m.count("presample_epoch")) c.presample_epoch = std::stoi(m["presample_epoch"]);
  if (m.count("ggms_replicate_percentage")) c.replicate_percentage = std::stod(m["ggms_replicate_percentage"]);
  else if (const char *rp = std::getenv("SAMGRAPH_GGMS_REPLICATE")) c.replicate_percentage = std::atof(rp);
  SAM_CHECK(c.replicate_percentage >= 0.0 && c.replicate_percentage <= 1.0);
  c.omp_thread_num = std::stoi(m["omp_thread_num"]);
  c.num_layer = std::stoull(m["num_layer"]);
  switch (c.run_arch) {
    case kArch0: case kArch1: case kArch2: case kArch3: case kArch4: case kArch7:
      SAM_CHECK(m.count("sampler_ctx") && m.count("trainer_ctx"));
      c.sampler_device = parse_ctx(m["sampler_ctx"]);
      c.trainer_device = parse_ctx(m["trainer_ctx"]);
      c.num_worker = 1;
      // arch0's CPU sampler is the oracle, not a product path: every single-process arch samples on a GPU here
      if (c.sampler_device < 0) c.sampler_device = c.trainer_device;
      SAM_CHECK_MSG(c.sampler_device >= 0, "a CUDA context is required (xgnn_b200 has no CPU path)");
      break;
    case kArch5:  // operation.cc:110-119: decoupled sampler / trainer processes
      SAM_CHECK(m.count("num_sample_worker") && m.count("num_train_worker"));
      c.num_worker = std::stoull(m["num_sample_worker"]);  // the train set is split over the SAMPLERS (dist_engine.cc:144-148)
      c.num_train_worker = std::stoull(m["num_train_worker"]);
      SAM_CHECK(c.num_worker >= 1 && c.num_worker <= (size_t)kMaxWorker);
      SAM_CHECK(c.num_train_worker >= 1 && c.num_train_worker <= (size_t)kMaxWorker);
      SAM_CHECK_MSG(!(m.count("have_switcher") && std::stoi(m["have_switcher"]) != 0), "arch5 switcher: out of scope");
      break;
    case kArch6:
      SAM_CHECK(m.count("num_worker"));
      c.num_worker = std::stoull(m["num_worker"]);
      SAM_CHECK(c.num_worker >= 1 && c.num_worker <= (size_t)kMaxWorker);
      break;
    default:
      fatal(__FILE__, __LINE__, "arch " + m["_arch"] + " is a different deployment mode (arch9, UM sampler): out of scope");
  }
  if (c.sample_type != XG_RANDOM_WALK) {  // operation.cc:150-161
    SAM_CHECK(m.count("num_fanout") && m.count("fanout"));
    const size_t nf = std::stoull(m["num_fanout"]);
    std::stringstream ss(m["fanout"]);
    for (size_t i = 0; i < nf; ++i) {
      size_t f = 0;
      ss >> f;
      c.fanout.push_back(f);
    }
  } else {  // operation.cc:162-175
    SAM_CHECK(m.count("random_walk_length") && m.count("random_walk_restart_prob") && m.count("num_random_walk") &&
              m.count("num_neighbor"));
    c.random_walk_length = std::stoull(m["random_walk_length"]);
    c.random_walk_restart_prob = std::stod(m["random_walk_restart_prob"]);
    c.num_random_walk = std::stoull(m["num_random_walk"]);
    c.num_neighbor = std::stoull(m["num_neighbor"]);
    c.fanout.assign(c.num_layer, c.num_neighbor);
  }
  if (m.count("use_dist_graph")) {  // operation.cc:191-203
    c.dist_graph_percentage = std::stod(m["use_dist_graph"]);
    c.use_dist_graph = c.dist_graph_percentage > 0.0;
  }
  if (m.count("part_cache") && m["part_cache"] == "True") {
    SAM_CHECK_MSG(c.run_arch == kArch6, "partition cache can only be used in arch6");  // operation.cc:205-211
    c.part_cache = true;
  }
  if (m.count("clique_size")) c.clique_size = std::stoi(m["clique_size"]);              // operation.cc:213-216
  if (m.count("ics22_song_alpha")) c.ics22_song_alpha = std::stod(m["ics22_song_alpha"]);  // :217-220
  if (m.count("use_ics22_song_solver") && m["use_ics22_song_solver"] == "True") {       // :221-230
    SAM_CHECK_MSG(c.use_dist_graph, "use ics22_song_solver need to open dist_graph");
    SAM_CHECK_MSG(!c.part_cache, "use ics22_song_solver need to close part cache");
    SAM_CHECK_MSG(c.clique_size > 0, "use ics22_song_solver need to set clique size");
    SAM_CHECK_MSG(c.run_arch == kArch6, "ics22_song_solver can only be used in arch6");
    SAM_CHECK_MSG(c.cache_policy == kCacheByDegree, "ics22_song_solver only support degree at present platform");  // cuda_cache_manager_host.cc:173-174
    c.use_ics22_song_solver = true;
  }
  c.gpu_extract = m.count("gpu_extract") && m["gpu_extract"] == "True";
  if (m.count("unified_memory") && m["unified_memory"] == "True")
    fatal(__FILE__, __LINE__, "unified-memory sampling (arch9 / UM) is out of scope");
  c.is_configured = true;
}

void samgraph_init(void) {  // operation.cc:328-334 (single process)
  SAM_CHECK(g_cfg.is_configured);
  SAM_CHECK_MSG(g_cfg.run_arch != kArch6 && g_cfg.run_arch != kArch5,
                "arch5 / arch6 start with samgraph_data_init (then fork, sample_init, train_init)");
  create_engine(false);
  Engine &e = engine();
  common_device_init(e, g_cfg.sampler_device);
  sampler_side_init(e);
  trainer_side_init(e);
}

void samgraph_data_init(void) {  // operation.cc:509-533: parent process, BEFORE fork; no CUDA call in here
  SAM_CHECK(g_cfg.is_configured);
  create_engine(true);
  Engine &e = engine();
  if (g_cfg.run_arch == kArch5) {  // the task queue every sampler and trainer process inherits (MemoryQueue ctor, memory_queue.cc:32-40)
    void *mem = ::mmap(nullptr, sizeof(TaskQueueBlock), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    SAM_CHECK(mem != MAP_FAILED);
    std::memset(mem, 0, sizeof(TaskQueueBlock));
    e.queue = static_cast<TaskQueueBlock *>(mem);
    e.queue->num_sampler = (int)g_cfg.num_worker;
    e.queue->num_trainer = (int)g_cfg.num_train_worker;
    for (int i = 0; i < kQueueDepth; ++i) {
      SAM_CHECK(sem_init(&e.queue->filled[i], 1, 0) == 0);  // sem_list
      SAM_CHECK(sem_init(&e.queue->vacant[i], 1, 1) == 0);  // release_list
    }
    for (int w = 0; w < kMaxWorker; ++w)
      for (int k = 0; k < kStagePerSampler; ++k) SAM_CHECK(sem_init(&e.queue->stage_free[w][k], 1, 1) == 0);
  }
  if (g_cfg.run_arch == kArch6) {
    void *mem = ::mmap(nullptr, sizeof(SharedBlock), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    SAM_CHECK(mem != MAP_FAILED);
    e.shared = new (mem) SharedBlock();
    std::memset(static_cast<void *>(e.shared), 0, sizeof(SharedBlock));
    const int W = (int)g_cfg.num_worker;
    e.shared->num_worker = W;
    pthread_barrierattr_t attr;
    pthread_barrierattr_init(&attr);
    pthread_barrierattr_setpshared(&attr, PTHREAD_PROCESS_SHARED);  // dist_graph.cu:619-630
    SAM_CHECK(pthread_barrier_init(&e.shared->barrier, &attr, (unsigned)W) == 0);
    // placement: PartitionSolver over the link matrix.  SAMGRAPH_TOPO_MATRIX = file with W rows of W GB/s numbers
    // (0 = no P2P path); default = all-to-all (NVSwitch), for which the solver returns the identity placement.
    std::vector<double> bw((size_t)W * W, 770.0);
    for (int i = 0; i < W; ++i) bw[(size_t)i * W + i] = 6500.0;
    if (const char *f = std::getenv("SAMGRAPH_TOPO_MATRIX")) {
      std::ifstream in(f);
      SAM_CHECK_MSG(in.is_open(), std::string("cannot open ") + f);
      for (auto &x : bw) SAM_CHECK_MSG(static_cast<bool>(in >> x), "SAMGRAPH_TOPO_MATRIX is too short");
    }
    std::vector<int> nparts(W), parts((size_t)W * W), access((size_t)W * W);
    SAM_XG(xg_partition_solve(bw.data(), W, nparts.data(), parts.data(), access.data()));
    for (int g = 0; g < W; ++g) {
      e.shared->num_parts[g] = nparts[g];
      for (int p = 0; p < W; ++p) {
        e.shared->part_ids[g][p] = parts[(size_t)g * W + p];
        e.shared->access[g][p] = access[(size_t)g * W + p];
      }
    }
  }
}

void samgraph_sample_init(int worker_id, const char *ctx) {  // operation.cc:535-540 -> DistEngine::SampleInit
  Engine &e = engine();
  SAM_CHECK((g_cfg.run_arch == kArch6 || g_cfg.run_arch == kArch5) && worker_id >= 0 && (size_t)worker_id < g_cfg.num_worker);
  if (g_cfg.run_arch == kArch5) {
    SAM_CHECK_MSG(e.role == 0, "arch5: a process is either a sampler or a trainer");
    SAM_CHECK_MSG(g_cfg.cache_policy != kCacheByPreSample, "arch5: presampled ranking needs the sampler in the trainer process");
    e.role = 1;
  }
  e.worker_id = worker_id;
  common_device_init(e, parse_ctx(ctx));
  sampler_side_init(e);
}

void samgraph_train_init(int worker_id, const char *ctx) {  // operation.cc:549-554 -> DistEngine::TrainInit
  Engine &e = engine();
  if (g_cfg.run_arch == kArch5) {  // a trainer-only process on its own GPU: cache + labels, no topology
    SAM_CHECK_MSG(e.role == 0, "arch5: a process is either a sampler or a trainer");
    SAM_CHECK(worker_id >= 0 && (size_t)worker_id < g_cfg.num_train_worker);
    e.role = 2;
    e.worker_id = worker_id;
    common_device_init(e, parse_ctx(ctx));
    trainer_side_init(e);
    return;
  }
  SAM_CHECK(g_cfg.run_arch == kArch6 && worker_id == e.worker_id);
  SAM_CHECK_MSG(parse_ctx(ctx) == e.device, "arch6 samples and trains on the same GPU (dist_engine.cc:589-700)");
  trainer_side_init(e);
}

void samgraph_um_sample_init(int) { fatal(__FILE__, __LINE__, "samgraph_um_sample_init belongs to arch9 (UM sampler): out of scope"); }
void samgraph_switch_init(int, const char *, double) { fatal(__FILE__, __LINE__, "samgraph_switch_init belongs to arch5's switcher: out of scope"); }

void samgraph_sample_once(void) {  // operation.cc:380
  Engine &e = engine();
  if (e.role == 2) {  // arch5 trainer, not pipelined: one extract-loop iteration (multi_gpu/train_graphsage.py:293-296)
    recv_task(e);
    return;
  }
  if (!produce_one(e)) {
    std::lock_guard<std::mutex> lk(e.mu);
    e.producer_done = true;
  }
}

static void extract_loop(Engine *e, long count) {  // DistEngine::StartExtract (arch5 trainer): `count` Tasks off the queue
  SAM_CUDA(cudaSetDevice(e->device));
  std::unique_ptr<PendingTask> next;
  for (long i = 0; i < count; ++i) {
    if (next) release_stage(*e, *next);  // its copy ran under the previous gather: do not sit on the stage while the pool is full
    {
      std::unique_lock<std::mutex> lk(e->mu);
      e->cv.wait(lk, [&] { return e->pool.size() < 2 || e->shutdown; });
      if (e->shutdown) return;
    }
    if (!next) next = recv_begin(*e);
    std::unique_ptr<PendingTask> cur = std::move(next);
    if (!cur) break;
    release_stage(*e, *cur);  // before the (possibly blocking) dequeue below: see release_stage
    if (i + 1 < count) {      // Task i+1 crosses NVLink while Task i's rows are gathered
      next = recv_begin(*e);
      if (next) next->prefetched = true;
    }
    if (!recv_finish(*e, std::move(cur))) break;
  }
  if (next) {  // shutdown in mid-flight: give the stage and the landing buffer back
    release_stage(*e, *next);
    std::lock_guard<std::mutex> lk(e->mu);
    next->land->busy = false;
  }
}

static void producer_loop(Engine *e, long count) {
  SAM_CUDA(cudaSetDevice(e->device));
  for (long i = 0; count < 0 || i < count; ++i) {
    {  // GraphPool::Full(): bounded look-ahead (graph_pool.h: max_copying_jobs)
      std::unique_lock<std::mutex> lk(e->mu);
      e->cv.wait(lk, [&] { return e->pool.size() < 2 || e->shutdown; });
      if (e->shutdown) return;
    }
    if (!produce_one(*e)) break;
  }
  std::lock_guard<std::mutex> lk(e->mu);
  e->producer_done = true;
  e->cv.notify_all();
}

void samgraph_start(void) {  // operation.cc:336-345: background producer for every epoch
  Engine &e = engine();
  SAM_CHECK(e.sampler_ready && !e.shutdown);
  e.threads.emplace_back(producer_loop, &e, -1L);
}

void samgraph_extract_start(int count) {  // operation.cc:556-559 -> DistEngine::StartExtract
  Engine &e = engine();
  if (e.role == 2) {
    SAM_CHECK(e.trainer_ready && !e.shutdown);
    e.threads.emplace_back(extract_loop, &e, (long)count);
    return;
  }
  SAM_CHECK(e.sampler_ready && !e.shutdown);
  e.threads.emplace_back(producer_loop, &e, (long)count);
}

uint64_t samgraph_get_next_batch(void) {  // operation.cc:366-378
  Engine &e = engine();
  std::unique_lock<std::mutex> lk(e.mu);
  e.current.reset();  // SetGraphBatch(nullptr): the previous batch is dropped unless a tensor still holds it
  e.cv.notify_all();
  e.cv.wait(lk, [&] { return !e.pool.empty() || e.shutdown; });
  SAM_CHECK(!e.pool.empty());
  e.current = e.pool.front();
  e.pool.pop_front();
  e.cv.notify_all();
  return e.current->key;
}

static BatchHolder &current_batch() {
  Engine &e = engine();
  SAM_CHECK_MSG(e.current != nullptr, "no current batch: call samgraph_get_next_batch first");
  return *e.current;
}
size_t samgraph_get_graph_num_src(uint64_t, int g) { return current_batch().view.num_src[g]; }
size_t samgraph_get_graph_num_dst(uint64_t, int g) { return current_batch().view.num_dst[g]; }
size_t samgraph_get_graph_num_edge(uint64_t, int g) { return current_batch().view.num_edge[g]; }

size_t samgraph_num_epoch(void) { return g_cfg.num_epoch; }
size_t samgraph_steps_per_epoch(void) { return engine().num_step; }
size_t samgraph_num_local_step(void) { return engine().num_local_step; }
size_t samgraph_num_class(void) { return engine().ds.num_class; }
size_t samgraph_feat_dim(void) { return engine().ds.feat_dim; }

void samgraph_shutdown(void) {  // operation.cc:397-403
  Engine &e = engine();
  if (e.role == 1 && e.queue) {
    // the payload of every Task this sampler sent lives in ITS HBM until a trainer has pulled it: do not let the
    // process (and with it the IPC-exported stage buffers) go away earlier (the reference's scripts keep samplers
    // alive with a run-end barrier, multi_gpu/train_graphsage.py:170-171)
    for (int k = 0; k < kStagePerSampler; ++k) {
      if (!e.stages[k].ptr) continue;
      timespec ts;
      clock_gettime(CLOCK_REALTIME, &ts);
      ts.tv_sec += 120;
      if (sem_timedwait(&e.queue->stage_free[e.worker_id][k], &ts) != 0)
        std::fprintf(stderr, "[samgraph] sampler %d: stage %d was never pulled by a trainer\n", e.worker_id, k);
    }
  }
  {
    std::lock_guard<std::mutex> lk(e.mu);
    e.shutdown = true;
  }
  e.cv.notify_all();
  for (auto &t : e.threads) t.join();
  e.threads.clear();
  e.inflight.clear();  // Tasks still inside the sampler engine: their slots go back (xg_batch_release waits for them)
  e.pool.clear();
  e.current.reset();
  if (e.stream) cudaStreamSynchronize(e.stream);
}

int samgraph_wait_one_child(void) {  // operation.cc:573-584
  int st = 0;
  pid_t pid = ::waitpid(-1, &st, 0);
  if (pid < 0) return 1;
  if (WIFSIGNALED(st)) return 1;
  return WEXITSTATUS(st) != 0 ? 1 : 0;
}

void samgraph_forward_barrier(void) { barrier(engine()); }

}  // extern "C"

extern "C" {

// ---- Profiler surface (operation.cc:405-500) ---------------------------------------------------------
static uint64_t batch_key(uint64_t epoch, uint64_t step) { return epoch * engine().num_step + step; }  // engine.h:50-54

void samgraph_log_step(uint64_t epoch, uint64_t step, int item, double val) {
  SAM_CHECK(item >= 0 && item < kNumLogStepItems);
  g_prof.log_step(batch_key(epoch, step), item, val, false);
}
void samgraph_log_step_by_key(uint64_t key, int item, double val) {
  SAM_CHECK(item >= 0 && item < kNumLogStepItems);
  g_prof.log_step(key, item, val, false);
}
void samgraph_log_step_add(uint64_t epoch, uint64_t step, int item, double val) {
  SAM_CHECK(item >= 0 && item < kNumLogStepItems);
  g_prof.log_step(batch_key(epoch, step), item, val, true);
}
void samgraph_log_epoch_add(uint64_t epoch, int item, double val) {
  SAM_CHECK(item >= 0 && item < kNumLogEpochItems);
  g_prof.log_epoch_add(epoch, item, val);
}
double samgraph_get_log_init_value(int item) {
  SAM_CHECK(item >= 0 && item < kNumLogInitItems);
  return g_prof.init[item];
}
double samgraph_get_log_step_value(uint64_t epoch, uint64_t step, int item) {
  SAM_CHECK(item >= 0 && item < kNumLogStepItems);
  return g_prof.get_step(batch_key(epoch, step), item);
}
double samgraph_get_log_step_value_by_key(uint64_t key, int item) {
  SAM_CHECK(item >= 0 && item < kNumLogStepItems);
  return g_prof.get_step(key, item);
}
double samgraph_get_log_epoch_value(uint64_t epoch, int item) {
  SAM_CHECK(item >= 0 && item < kNumLogEpochItems);
  return g_prof.get_epoch(epoch, item);
}

void samgraph_report_init(void) {
  std::printf("    [Init Profiler Level 1] load dataset %.4f | sampler %.4f | trainer %.4f\n",
              g_prof.init[kLogInitL2LoadDataset], g_prof.init[kLogInitL1Sampler], g_prof.init[kLogInitL1Trainer]);
  std::fflush(stdout);
}
void samgraph_report_step(uint64_t epoch, uint64_t step) {
  const uint64_t k = batch_key(epoch, step);
  std::printf("    [Step(%llu) Profiler Level 1 E%llu S%llu] sample+extract %.4f | nodes %.0f | samples %.0f\n",
              (unsigned long long)k, (unsigned long long)epoch, (unsigned long long)step, g_prof.get_step(k, kLogL1SampleTime),
              g_prof.get_step(k, kLogL1NumNode), g_prof.get_step(k, kLogL1NumSample));
}
void samgraph_report_step_average(uint64_t epoch, uint64_t step) {
  samgraph_report_step(epoch, step);
  std::fflush(stdout);
}
void samgraph_report_epoch(uint64_t epoch) {
  std::printf("    [Epoch Profiler E%llu] sample+extract %.4f | samples %.0f | feature bytes %.0f\n",
              (unsigned long long)epoch, g_prof.get_epoch(epoch, kLogEpochSampleTime),
              g_prof.get_epoch(epoch, kLogEpochNumSample), g_prof.get_epoch(epoch, kLogEpochFeatureBytes));
}
void samgraph_report_epoch_average(uint64_t epoch) {
  samgraph_report_epoch(epoch);
  std::fflush(stdout);
}
void samgraph_report_node_access(void) {  // Profiler::ReportNodeAccess (profiler.cc): SAMGRAPH_LOG_NODE_ACCESS statistics
  std::fprintf(stderr, "[samgraph] node access statistics: not collected (SAMGRAPH_LOG_NODE_ACCESS is not implemented by xgnn_b200)\n");
}

// Chrome-trace style begin/end marks (profiler.cc TraceStepBegin/End, DumpTrace)
namespace {
struct TraceEvent { uint64_t key; int item; uint64_t begin_us, end_us; };
std::mutex g_trace_mu;
std::vector<TraceEvent> g_trace;
uint64_t now_us() {
  return (uint64_t)std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
}
}  // namespace
void samgraph_trace_step_begin(uint64_t key, int item, uint64_t ts) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  g_trace.push_back({key, item, ts, 0});
}
void samgraph_trace_step_end(uint64_t key, int item, uint64_t ts) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  for (auto it = g_trace.rbegin(); it != g_trace.rend(); ++it)
    if (it->key == key && it->item == item && it->end_us == 0) {
      it->end_us = ts;
      return;
    }
}
void samgraph_trace_step_begin_now(uint64_t key, int item) { samgraph_trace_step_begin(key, item, now_us()); }
void samgraph_trace_step_end_now(uint64_t key, int item) { samgraph_trace_step_end(key, item, now_us()); }
void samgraph_dump_trace(void) {
  std::lock_guard<std::mutex> lk(g_trace_mu);
  std::fprintf(stderr, "[\n");
  for (const auto &t : g_trace)
    std::fprintf(stderr, "{\"name\":\"item%d\",\"ph\":\"X\",\"pid\":0,\"tid\":%d,\"ts\":%llu,\"dur\":%llu,\"args\":{\"key\":%llu}},\n",
                 t.item, t.item, (unsigned long long)t.begin_us,
                 (unsigned long long)(t.end_us > t.begin_us ? t.end_us - t.begin_us : 0), (unsigned long long)t.key);
  std::fprintf(stderr, "]\n");
}

// ---- tensor views (torch/adapter.cc:62-193) -----------------------------------------------------------
void samgraph_get_graph_tensor(uint64_t key, int which, int layer, samgraph_tensor_view *out) {
  SAM_CHECK(out != nullptr);
  Engine &e = engine();
  std::memset(out, 0, sizeof(*out));
  out->ndim = 1;
  if (which == SAMGRAPH_TENSOR_DATASET_FEAT || which == SAMGRAPH_TENSOR_DATASET_LABEL) {
    out->device = -1;
    if (which == SAMGRAPH_TENSOR_DATASET_FEAT) {
      out->data = const_cast<void *>(e.ds.feat);
      out->ndim = 2;
      out->shape[0] = (int64_t)e.ds.feat_table_rows();
      out->shape[1] = (int64_t)e.ds.feat_dim;
      out->dtype = e.ds.feat_dtype;
    } else {
      out->data = const_cast<int64_t *>(e.ds.label);
      out->shape[0] = (int64_t)e.ds.num_node;
      out->dtype = XG_I64;
    }
    return;
  }
  BatchPtr cur;
  {
    std::lock_guard<std::mutex> lk(e.mu);
    cur = e.current;
  }
  SAM_CHECK_MSG(cur != nullptr, "no current batch");
  SAM_CHECK_MSG(key == cur->key, "key does not name the current batch (CHECK_EQ, torch/adapter.cc:68)");
  const xg_batch_view &v = cur->view;
  out->device = e.device;
  out->dtype = XG_I32;
  switch (which) {
    case SAMGRAPH_TENSOR_FEAT:
      SAM_CHECK_MSG(v.input_feat != nullptr, "no features: train_init has not run");
      out->data = const_cast<void *>(v.input_feat);
      out->ndim = 2;
      out->shape[0] = v.num_input;
      out->shape[1] = (int64_t)e.ds.feat_dim;
      out->dtype = e.ds.feat_dtype;
      break;
    case SAMGRAPH_TENSOR_LABEL:
      SAM_CHECK_MSG(v.output_label != nullptr, "no labels: train_init has not run");
      out->data = const_cast<int64_t *>(v.output_label);
      out->shape[0] = v.num_output;
      out->dtype = XG_I64;
      break;
    case SAMGRAPH_TENSOR_ROW: case SAMGRAPH_TENSOR_COL: case SAMGRAPH_TENSOR_DATA: {
      SAM_CHECK(layer >= 0 && (uint32_t)layer < v.num_layer);
      const uint32_t *p = which == SAMGRAPH_TENSOR_ROW ? v.row[layer] : which == SAMGRAPH_TENSOR_COL ? v.col[layer] : v.data[layer];
      SAM_CHECK_MSG(p != nullptr || v.num_edge[layer] == 0, "this sampler produces no edge data");
      out->data = const_cast<uint32_t *>(p);
      out->shape[0] = (int64_t)v.num_edge[layer];
      break;
    }
    case SAMGRAPH_TENSOR_CSC_INDPTR:
      SAM_CHECK(layer >= 0 && (uint32_t)layer < v.num_layer);
      // NULL data = no CSC form (weighted / KHop1 samplers, or a seed layer with duplicate seeds): use ROW/COL
      out->data = const_cast<uint32_t *>(v.csc_indptr[layer]);
      out->shape[0] = v.csc_indptr[layer] ? (int64_t)v.num_dst[layer] + 1 : 0;
      break;
    case SAMGRAPH_TENSOR_INPUT_NODES:
      out->data = const_cast<uint32_t *>(v.input_nodes);
      out->shape[0] = v.num_input;
      break;
    case SAMGRAPH_TENSOR_OUTPUT_NODES:
      out->data = const_cast<uint32_t *>(v.output_nodes);
      out->shape[0] = v.num_output;
      break;
    default:
      fatal(__FILE__, __LINE__, "unknown tensor id");
  }
  out->owner = new BatchPtr(cur);  // the view keeps the batch (and its slot) alive, like from_blob's deleter
}

void samgraph_tensor_release(void *owner) {
  if (owner) delete static_cast<BatchPtr *>(owner);
}

}  // extern "C"

/* host-only hook: one epoch of DistAlignedShuffler::ReShuffle (dist/dist_shuffler_aligned.cc:75-121) applied in
 * place to `data` (already padded); lets the CPU test suite pin the permutation without a GPU */
extern "C" int xg_aligned_shuffle_epoch(uint32_t *data, size_t n, uint64_t epoch) {
  if (!data && n) return XG_ERR_INVALID;
  std::vector<uint32_t> v(data, data + n);
  sam::AlignedShuffler::shuffle_epoch(v, (size_t)epoch);
  std::copy(v.begin(), v.end(), data);
  return XG_OK;
}
