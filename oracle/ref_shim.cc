// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Glue that lets the reference's own arch0 CPU leaf functions be compiled *in place*
// from /root/reference (see oracle/Makefile) into oracle/_ref/libxgnn_ref.so and
// called through a flat C interface from tests/ and from bench.py's cpu_baseline /
// --impl reference leg.  Nothing here restates reference logic except the thin
// DoCPUSample/DoFeatureExtract driver loop at the bottom, which only *calls* the
// reference functions in the order samgraph/common/cpu/cpu_loops.cc:55-228 does.
//
// Why a shim: Device::Get(CPU()) in the reference goes through cudaHostAlloc
// (samgraph/common/cpu/cpu_device.cc:36-40); here it is backed by aligned malloc so the
// CPU path builds and runs without a CUDA runtime.
#include <omp.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common/common.h"
#include "common/constant.h"
#include "common/cpu/cpu_function.h"
#include "common/cpu/cpu_hashtable0.h"
#include "common/cpu/cpu_hashtable2.h"
#include "common/device.h"
#include "common/run_config.h"

namespace samgraph {
namespace common {

namespace {
class MallocDevice : public Device {
 public:
  void SetDevice(Context) override {}
  void *AllocDataSpace(Context, size_t nbytes, size_t alignment) override {
    void *p = nullptr;
    if (nbytes == 0) nbytes = alignment;
    if (posix_memalign(&p, alignment < 64 ? 64 : alignment, nbytes) != 0) abort();
    return p;
  }
  void FreeDataSpace(Context, void *ptr) override { free(ptr); }
  size_t WorkspaceActualSize(Context, void *) override { return 0; }
  void CopyDataFromTo(const void *from, size_t from_offset, void *to, size_t to_offset,
                      size_t nbytes, Context, Context, StreamHandle) override {
    memcpy(static_cast<char *>(to) + to_offset,
           static_cast<const char *>(from) + from_offset, nbytes);
  }
  void StreamSync(Context, StreamHandle) override {}
};
MallocDevice g_dev;
}  // namespace

// Three one-line helpers that live in common.cc, a file that cannot be compiled here because it
// drags in the CUDA-backed Tensor/Device machinery.  Declared in common.h / constant.h.
Context CPU(int device_id) { return Context(kCPU, device_id); }
std::string GetEnv(std::string key) {
  const char *v = getenv(key.c_str());
  return v ? std::string(v) : std::string();
}
bool IsEnvSet(std::string key) {
  const std::string v = GetEnv(key);
  return v == "ON" || v == "1";
}

Device *Device::Get(Context) { return &g_dev; }
void *Device::AllocWorkspace(Context ctx, size_t nbytes, double) {
  return AllocDataSpace(ctx, nbytes, kTempAllocaAlignment);
}
void Device::FreeWorkspace(Context ctx, void *ptr, size_t) { FreeDataSpace(ctx, ptr); }
StreamHandle Device::CreateStream(Context) { return nullptr; }
void Device::FreeStream(Context, StreamHandle) {}
void Device::SyncStreamFromTo(Context, StreamHandle, StreamHandle) {}

}  // namespace common
}  // namespace samgraph

using namespace samgraph::common;

// Working buffers live across calls, as in the reference: DoCPUSample / DoFeatureExtract take theirs from the
// device's workspace pool (cpu/cpu_loops.cc:55-228, workspace_pool.cc), which recycles blocks, so a steady-state
// batch neither allocates nor zero-fills.  (r01 allocated fresh zero-filled std::vectors inside the timed region:
// ~2.6 GB of first-touch page faults per batch charged to the reference.)  Grow-only; never shrunk.
namespace {
struct Arch0Pool {
  std::vector<uint32_t> out_src, out_dst, unique, cur, new_src, new_dst;
  std::vector<char> feat, label;
};
Arch0Pool g_pool;
template <typename T>
T *at_least(std::vector<T> &v, size_t n) {
  if (v.size() < n) v.resize(n + n / 8);
  return v.data();
}
}  // namespace

extern "C" {

void ref_set_threads(int n) { RunConfig::omp_thread_num = n; }
int ref_get_threads() { return RunConfig::omp_thread_num; }

// cpu_sampling_khop0.cc:29-83
void ref_cpu_sample_khop0(const uint32_t *indptr, const uint32_t *indices, const uint32_t *input,
                          size_t num_input, uint32_t *out_src, uint32_t *out_dst,
                          size_t *num_out, size_t fanout) {
  cpu::CPUSampleKHop0(indptr, indices, input, num_input, out_src, out_dst, num_out, fanout);
}

// cpu_sampling_khop2.cc:29-76  (mutates indices)
void ref_cpu_sample_khop2(const uint32_t *indptr, uint32_t *indices, const uint32_t *input,
                          size_t num_input, uint32_t *out_src, uint32_t *out_dst,
                          size_t *num_out, size_t fanout) {
  cpu::CPUSampleKHop2(indptr, indices, input, num_input, out_src, out_dst, num_out, fanout);
}

// cpu_extraction.cc:66-90 ; dtype uses the reference's DataType enum numbering (common.h)
void ref_cpu_extract(void *dst, const void *src, const uint32_t *index, size_t num_index,
                     size_t dim, int dtype) {
  cpu::CPUExtract(dst, src, index, num_index, dim, static_cast<DataType>(dtype));
}
// cpu_extraction.cc:47-64 + 92-118: the SAMGRAPH_EMPTY_FEAT mock variant (row index masked to 2^bits rows)
void ref_cpu_mock_extract(void *dst, const void *src, const uint32_t *index, size_t num_index,
                          size_t dim, int dtype, unsigned bits) {
  RunConfig::option_empty_feat = bits;
  cpu::CPUMockExtract(dst, src, index, num_index, dim, static_cast<DataType>(dtype));
  RunConfig::option_empty_feat = 0;
}
int ref_dtype_f32() { return kF32; }
int ref_dtype_f16() { return kF16; }
int ref_dtype_i64() { return kI64; }
int ref_dtype_u8() { return kU8; }
int ref_dtype_i32() { return kI32; }
int ref_dtype_f64() { return kF64; }

// cpu_hashtable0.cc / cpu_hashtable2.cc behind one opaque handle. kind: 0 | 2
void *ref_hashtable_create(int kind, size_t max_items) {
  cpu::CPUHashTable *t = nullptr;
  if (kind == 0) t = new cpu::CPUHashTable0(max_items);
  else if (kind == 2) t = new cpu::CPUHashTable2(max_items);
  return t;
}
void ref_hashtable_destroy(void *h) { delete static_cast<cpu::CPUHashTable *>(h); }
void ref_hashtable_reset(void *h) { static_cast<cpu::CPUHashTable *>(h)->Reset(); }
void ref_hashtable_populate(void *h, const uint32_t *in, size_t n) {
  static_cast<cpu::CPUHashTable *>(h)->Populate(in, n);
}
size_t ref_hashtable_num_items(void *h) { return static_cast<cpu::CPUHashTable *>(h)->NumItems(); }
void ref_hashtable_map_nodes(void *h, uint32_t *out, size_t n) {
  static_cast<cpu::CPUHashTable *>(h)->MapNodes(out, n);
}
void ref_hashtable_map_edges(void *h, const uint32_t *src, const uint32_t *dst, size_t len,
                             uint32_t *new_src, uint32_t *new_dst) {
  static_cast<cpu::CPUHashTable *>(h)->MapEdges(src, dst, len, new_src, new_dst);
}

// ---------------------------------------------------------------------------------------------
// One arch0 minibatch: the call order of DoCPUSample + DoFeatureExtract
// (samgraph/common/cpu/cpu_loops.cc:55-228).  Every arithmetic step is a call into the
// reference; this function only owns buffers.  H2D copies (DoGraphCopy/DoFeatureCopy) are not
// part of the CPU baseline.  Outputs are optional (pass NULL to skip the copy-out).
//   fanouts[num_layers] is consumed last -> first, as in the reference.
//   per layer l: rows[l] = new_dst (local neighbour ids), cols[l] = new_src, counts[3*l..] =
//   {num_src(num_unique), num_dst(num_input), num_edge}.
// Returns total sampled edges; *num_input_nodes receives |input_nodes|.
struct RefBatchOut {
  std::vector<std::vector<uint32_t>> rows, cols;
  std::vector<uint32_t> input_nodes;
  std::vector<char> feat;
  std::vector<char> label;
  std::vector<size_t> counts;
};

void *ref_batch_create() { return new RefBatchOut(); }
void ref_batch_destroy(void *b) { delete static_cast<RefBatchOut *>(b); }
const uint32_t *ref_batch_row(void *b, int l) { return static_cast<RefBatchOut *>(b)->rows[l].data(); }
const uint32_t *ref_batch_col(void *b, int l) { return static_cast<RefBatchOut *>(b)->cols[l].data(); }
const size_t *ref_batch_counts(void *b) { return static_cast<RefBatchOut *>(b)->counts.data(); }
const uint32_t *ref_batch_input_nodes(void *b) { return static_cast<RefBatchOut *>(b)->input_nodes.data(); }
const void *ref_batch_feat(void *b) { return static_cast<RefBatchOut *>(b)->feat.data(); }
const void *ref_batch_label(void *b) { return static_cast<RefBatchOut *>(b)->label.data(); }

size_t ref_arch0_batch(void *hashtable, int sample_type /*0=khop0, 5=khop2*/,
                       const uint32_t *indptr, uint32_t *indices, const uint32_t *seeds,
                       size_t num_seeds, const size_t *fanouts, int num_layers,
                       const void *feat, size_t feat_dim, int feat_dtype, size_t feat_elem_bytes,
                       const void *label /*i64 or NULL*/, void *batch_out /*or NULL*/,
                       size_t *num_input_nodes, double *sample_seconds, double *extract_seconds) {
  auto *ht = static_cast<cpu::CPUHashTable *>(hashtable);
  auto *out = static_cast<RefBatchOut *>(batch_out);
  Arch0Pool &pool = g_pool;
  auto t_begin = std::chrono::steady_clock::now();
  ht->Reset();
  ht->Populate(seeds, num_seeds);
  if (out) {
    out->rows.assign(num_layers, {});
    out->cols.assign(num_layers, {});
    out->counts.assign(3 * num_layers, 0);
  }
  uint32_t *cur = at_least(pool.cur, num_seeds);
  memcpy(cur, seeds, num_seeds * sizeof(uint32_t));
  size_t num_cur = num_seeds;
  size_t total = 0;
  for (int i = num_layers - 1; i >= 0; --i) {
    const size_t fanout = fanouts[i];
    const size_t num_input = num_cur;
    uint32_t *out_src = at_least(pool.out_src, num_input * fanout);
    uint32_t *out_dst = at_least(pool.out_dst, num_input * fanout);
    size_t num_out = 0;
    if (sample_type == 0) {
      cpu::CPUSampleKHop0(indptr, indices, cur, num_input, out_src, out_dst, &num_out, fanout);
    } else {
      cpu::CPUSampleKHop2(indptr, indices, cur, num_input, out_src, out_dst, &num_out, fanout);
    }
    ht->Populate(out_dst, num_out);
    size_t num_unique = ht->NumItems();
    uint32_t *unique = at_least(pool.unique, num_unique);
    ht->MapNodes(unique, num_unique);
    uint32_t *new_src = at_least(pool.new_src, num_out);
    uint32_t *new_dst = at_least(pool.new_dst, num_out);
    ht->MapEdges(out_src, out_dst, num_out, new_src, new_dst);
    if (out) {  // parity tests only: copy the layer out of the pool
      out->rows[i].assign(new_dst, new_dst + num_out);
      out->cols[i].assign(new_src, new_src + num_out);
      out->counts[3 * i + 0] = num_unique;
      out->counts[3 * i + 1] = num_input;
      out->counts[3 * i + 2] = num_out;
    }
    pool.cur.swap(pool.unique);  // the unique list is the next layer's input
    cur = pool.cur.data();
    num_cur = num_unique;
    total += num_out;
  }
  auto t_sampled = std::chrono::steady_clock::now();
  if (num_input_nodes) *num_input_nodes = num_cur;
  char *f = nullptr, *l = nullptr;
  if (feat) {
    f = at_least(pool.feat, num_cur * feat_dim * feat_elem_bytes);
    cpu::CPUExtract(f, feat, cur, num_cur, feat_dim, static_cast<DataType>(feat_dtype));
  }
  if (label) {
    l = at_least(pool.label, num_seeds * 8);
    cpu::CPUExtract(l, label, seeds, num_seeds, 1, kI64);
  }
  auto t_end = std::chrono::steady_clock::now();
  if (out) {
    out->input_nodes.assign(cur, cur + num_cur);
    if (f) out->feat.assign(f, f + num_cur * feat_dim * feat_elem_bytes);
    if (l) out->label.assign(l, l + num_seeds * 8);
  }
  if (sample_seconds) *sample_seconds = std::chrono::duration<double>(t_sampled - t_begin).count();
  if (extract_seconds) *extract_seconds = std::chrono::duration<double>(t_end - t_sampled).count();
  return total;
}

}  // extern "C"
