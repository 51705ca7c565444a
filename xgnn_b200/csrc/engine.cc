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
  std::vector<size_t> fanout;
  size_t random_walk_length = 0, num_random_walk = 0, num_neighbor = 0;
  double random_walk_restart_prob = 0.0;
  int sampler_device = 0, trainer_device = 0;
  size_t num_worker = 1;
  bool use_dist_graph = false, part_cache = false, gpu_extract = false;
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
  std::vector<Mapping> maps;
  size_t feat_row_bytes() const { return feat_dim * (feat_dtype == XG_F16 ? 2 : feat_dtype == XG_F64 ? 8 : feat_dtype == XG_U8 ? 1 : 4); }
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
  SAM_CHECK_MSG(file_exists(dir + "feat.bin"), "feat.bin is missing (SAMGRAPH_EMPTY_FEAT mock features are not supported)");
  ds.feat = map_file(ds, dir + "feat.bin", ds.num_node * ds.feat_row_bytes(), shared);
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
constexpr int kLogL1NumSample = 0, kLogL1NumNode = 1, kLogL1SampleTotalTime = 2, kLogL1SampleTime = 3, kLogL1CopyTime = 6,
              kLogL1FeatureBytes = 9, kLogL1LabelBytes = 10, kLogL1IdBytes = 11, kLogL1GraphBytes = 12,
              kLogL2ShuffleTime = 17, kLogL2CoreSampleTime = 20;
constexpr int kLogEpochSampleTime = 0, kLogEpochSampleTotalTime = 3, kLogEpochCopyTime = 8, kLogEpochFeatureBytes = 12,
              kLogEpochNumSample = 15;
constexpr int kLogInitL1Common = 0, kLogInitL1Sampler = 1, kLogInitL1Trainer = 2, kLogInitL2LoadDataset = 6,
              kLogInitL2BuildCache = 10;

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
