// GGMS placement solver.
//
// Reference behaviour being replaced: PartitionSolver::solve + solver_recursive + search_access_config
// (samgraph/common/cuda/dist_graph.cu:40-222,728-777; standalone twin
// utility/nvlink_placement_solver/solver.cc) and PartitionSolver::DetectTopo (dist_graph.cu:684-884).
//
// Problem: n GPUs, part i starts on GPU i.  A GPU can only load from itself and from its P2P
// neighbours (bandwidth[g][j] != 0).  Replicate parts until every GPU reaches every part, so that
//   (1) the largest number of parts any GPU stores is minimal, and among those placements
//   (2) max over GPUs g of [ min over source choices of max_j (parts read from j) / bandwidth[g][j] ]
//       is minimal,
// and report for every GPU the parts it stores and the GPU it reads each part from.
//
// The search is the reference's exhaustive DFS; its visiting order and its strict "<" comparisons
// decide which of several equally good placements is returned, so both are kept (the goldens in
// tests/golden/solver_golden.json are the reference solver's own output).  The state is held in
// small fixed arrays and bit masks instead of std::set / std::multiset copies.
// On NVSwitch (B200) every pair is connected: nothing is unreachable, the DFS visits one leaf and
// the answer is "GPU g stores part g and reads part j from GPU j".
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace xg {
namespace {

constexpr int kMaxGpu = XG_MAX_PARTS;

struct Placement {
  uint32_t store[kMaxGpu];     // bit p: GPU stores part p
  int access[kMaxGpu][kMaxGpu];  // access[g][p] = GPU that g reads part p from
};

struct Search {
  int n = 0;
  const double *bw = nullptr;           // n x n
  uint32_t adj[kMaxGpu] = {};           // P2P neighbours of g (self included when bw[g][g] != 0)
  uint32_t store[kMaxGpu] = {};         // parts stored on g
  int reach[kMaxGpu][kMaxGpu] = {};     // reach[g][p] = how many neighbours of g store part p
  int best_parts = INT32_MAX;           // objective (1)
  double best_cost = DBL_MAX;           // objective (2)
  bool found = false;
  Placement best{};

  double bandwidth(int g, int j) const { return bw[g * n + j]; }

  // objective (2) for one GPU: cheapest assignment part -> source GPU; first minimum in
  // lexicographic order of the assignment wins (part 0 outermost, sources ascending)
  double cheapest_access(int g, int *assign_out) const {
    uint32_t sources[kMaxGpu];
    for (int p = 0; p < n; ++p) {
      sources[p] = 0;
      for (int j = 0; j < n; ++j)
        if (((adj[g] >> j) & 1u) && ((store[j] >> p) & 1u)) sources[p] |= 1u << j;
      if (!sources[p]) return DBL_MAX;  // part p unreachable from g: not a placement
    }
    int cur[kMaxGpu], cnt[kMaxGpu] = {};
    double best = DBL_MAX;
    // odometer over the per-part source lists; digit 0 (part 0) is the slowest
    for (int p = 0; p < n; ++p) {
      cur[p] = __builtin_ctz(sources[p]);
      cnt[cur[p]] += 1;
    }
    while (true) {
      double worst = 0.0;
      for (int j = 0; j < n; ++j) {
        if (cnt[j]) {
          const double c = static_cast<double>(cnt[j]) / bandwidth(g, j);
          if (c > worst) worst = c;
        }
      }
      if (worst < best) {
        best = worst;
        std::memcpy(assign_out, cur, sizeof(int) * n);
      }
      int p = n - 1;
      for (; p >= 0; --p) {  // advance the fastest digit that still has a larger source
        const uint32_t higher = sources[p] & ~((2u << cur[p]) - 1u);
        cnt[cur[p]] -= 1;
        if (higher) {
          cur[p] = __builtin_ctz(higher);
          cnt[cur[p]] += 1;
          break;
        }
        cur[p] = __builtin_ctz(sources[p]);
        cnt[cur[p]] += 1;
      }
      if (p < 0) break;
    }
    return best;
  }

  void leaf() {
    int most = 0;
    for (int g = 0; g < n; ++g) most = std::max(most, __builtin_popcount(store[g]));
    if (most > best_parts) return;
    if (most < best_parts) {
      best_parts = most;
      best_cost = DBL_MAX;
    }
    Placement cand{};
    double cost = 0.0;
    for (int g = 0; g < n; ++g) {
      const double c = cheapest_access(g, cand.access[g]);
      if (c > cost) cost = c;
    }
    if (cost < best_cost) {
      best_cost = cost;
      std::memcpy(cand.store, store, sizeof(store));
      best = cand;
      found = true;
    }
  }

  void place(int part, int gpu, int delta) {
    if (delta > 0) store[gpu] |= 1u << part;
    else store[gpu] &= ~(1u << part);
    for (int k = 0; k < n; ++k)
      if ((adj[gpu] >> k) & 1u) reach[k][part] += delta;
  }

  // `missing`: parts GPU g cannot reach yet (fixed when g is first visited), next one at index i
  void visit(int g, const std::vector<int> &missing_in, size_t i) {
    if (g == n) return leaf();
    std::vector<int> missing = missing_in;
    if (missing.empty()) {
      for (int p = 0; p < n; ++p)
        if (reach[g][p] == 0) missing.push_back(p);
      i = 0;
    }
    if (i >= missing.size()) return visit(g + 1, {}, 0);
    const int part = missing[i];
    // candidate hosts: neighbours of g, preferring few stored parts, then many neighbours still
    // lacking the part, then hosts that cannot reach it themselves, then low id; only the hosts
    // tied with the best on the first three criteria are tried
    struct Host { int id, stored, need, has; };
    std::vector<Host> hosts;
    for (int j = 0; j < n; ++j) {
      if (!((adj[g] >> j) & 1u)) continue;
      int need = 0;
      for (int k = 0; k < n; ++k)
        if (((adj[j] >> k) & 1u) && reach[k][part] == 0) ++need;
      hosts.push_back({j, __builtin_popcount(store[j]), need, reach[j][part] ? 1 : 0});
    }
    if (hosts.empty()) return;
    std::sort(hosts.begin(), hosts.end(), [](const Host &a, const Host &b) {
      if (a.stored != b.stored) return a.stored < b.stored;
      if (a.need != b.need) return a.need > b.need;
      if (a.has != b.has) return a.has < b.has;
      return a.id < b.id;
    });
    size_t keep = 1;
    while (keep < hosts.size() && hosts[keep].stored == hosts[0].stored && hosts[keep].need == hosts[0].need &&
           hosts[keep].has == hosts[0].has)
      ++keep;
    for (size_t h = 0; h < keep; ++h) {
      place(part, hosts[h].id, +1);
      visit(g, missing, i + 1);
      place(part, hosts[h].id, -1);
    }
  }
};

}  // namespace
}  // namespace xg

using namespace xg;

extern "C" {

int xg_partition_solve(const double *bandwidth, int num_gpu, int *out_num_parts, int *out_part_ids,
                       int *out_access) {
  XG_REQUIRE(bandwidth && out_num_parts && out_part_ids && out_access, "NULL argument");
  XG_REQUIRE(num_gpu >= 1 && num_gpu <= kMaxGpu, "num_gpu must be in [1, 16]");
  Search s;
  s.n = num_gpu;
  s.bw = bandwidth;
  for (int i = 0; i < num_gpu; ++i) {
    s.store[i] = 1u << i;
    for (int j = 0; j < num_gpu; ++j)
      if (std::fabs(bandwidth[i * num_gpu + j]) > 1e-6) s.adj[i] |= 1u << j;  // dist_graph.cu:745
  }
  for (int i = 0; i < num_gpu; ++i)
    for (int j = 0; j < num_gpu; ++j)
      if ((s.adj[i] >> j) & 1u) s.reach[i][j] = 1;  // part j sits on GPU j
  s.visit(0, {}, 0);
  if (!s.found) {
    set_error("no placement: some GPU has no P2P path to a copy of every part");
    return XG_ERR_UNSUPPORTED;
  }
  for (int g = 0; g < num_gpu; ++g) {
    int k = 0;
    for (int p = 0; p < num_gpu; ++p) {
      out_part_ids[g * num_gpu + p] = -1;
      out_access[g * num_gpu + p] = s.best.access[g][p];
    }
    for (int p = 0; p < num_gpu; ++p)
      if ((s.best.store[g] >> p) & 1u) out_part_ids[g * num_gpu + k++] = p;
    out_num_parts[g] = k;
  }
  return XG_OK;
}

/* PartitionSolver::DetectTopo_child (dist_graph.cu:779-884): P2P capability and a timed 128 MiB copy
 * per ordered pair.  Runs in the calling process (the reference forks to keep the parent free of a
 * CUDA context; callers that fork afterwards should do the same around this call). */
int xg_detect_topology(int num_gpu, double *out_bandwidth) {
  XG_REQUIRE(out_bandwidth != nullptr, "out_bandwidth is NULL");
  int have = 0;
  XG_CUDA(cudaGetDeviceCount(&have));
  XG_REQUIRE(num_gpu >= 1 && num_gpu <= have && num_gpu <= kMaxGpu, "num_gpu exceeds the visible devices");
  const size_t nbytes = size_t(1) << 27;
  int prev = 0;
  XG_CUDA(cudaGetDevice(&prev));
  std::vector<void *> a(num_gpu, nullptr), b(num_gpu, nullptr);
  for (int d = 0; d < num_gpu; ++d) {
    XG_CUDA(cudaSetDevice(d));
    XG_CUDA(cudaMalloc(&a[d], nbytes));
    XG_CUDA(cudaMalloc(&b[d], nbytes));
  }
  for (int d = 0; d < num_gpu; ++d) {
    XG_CUDA(cudaSetDevice(d));
    for (int p = 0; p < num_gpu; ++p) {
      out_bandwidth[d * num_gpu + p] = 0.0;
      int can = (d == p);
      if (d != p) XG_CUDA(cudaDeviceCanAccessPeer(&can, d, p));
      if (!can) continue;
      if (d != p) {
        cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else XG_CUDA(e);
      }
      XG_CUDA(cudaMemcpyAsync(a[d], b[p], nbytes, cudaMemcpyDefault, nullptr));  // warm the path
      XG_CUDA(cudaDeviceSynchronize());
      const auto t0 = std::chrono::steady_clock::now();
      XG_CUDA(cudaMemcpyAsync(a[d], b[p], nbytes, cudaMemcpyDefault, nullptr));
      XG_CUDA(cudaDeviceSynchronize());
      const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      out_bandwidth[d * num_gpu + p] = (d == p ? 2.0 : 1.0) * nbytes / sec / 1e9;
    }
  }
  for (int d = 0; d < num_gpu; ++d) {
    cudaSetDevice(d);
    cudaFree(a[d]);
    cudaFree(b[d]);
  }
  cudaSetDevice(prev);
  return XG_OK;
}

}  // extern "C"
