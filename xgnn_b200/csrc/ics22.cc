// ICS'22 (Song et al.) clique feature placement - the comparison baseline of the reference.
//
// Reference behaviour being replaced: ICS22SongPlacementSolver + ICS22SongDistGraph
// (samgraph/common/cuda/ics22_song_dist_feature.cc:22-167,169-222) and the compact cache map GPUCacheManager builds
// from its result (cuda/cuda_cache_manager_host.cc:172-222).
//
// The policy: every GPU starts with the same `num_cached` hottest nodes (rank = in-degree descending, ties by node
// id descending), one row per rank.  Then, walking the row slots from the COLDEST cached rank upwards, a slot is
// turned from "replicated on all c GPUs of a clique" into "c different nodes, one per GPU": c-1 GPUs give the
// slot to the next c-1 uncached candidates, the remaining GPU keeps the slot's original node and serves it to the
// others.  Candidates go to the currently least-loaded GPUs (load = sum of the candidates' weights so far), the
// most-loaded GPU is the one that keeps the original.  A candidate is only admitted while
//     weight(candidate) >= alpha * weight(node of the slot),
// the walk stops at the first candidate that fails this, when the slots run out, or when every node is cached.
// Every clique gets the same layout relative to its first GPU.
//
// The reference materialises, for EVERY GPU, a device map u32[V], an index map u32[V] and the slot list, in one
// process before the fork (16 x V bytes per GPU).  A worker only ever needs its own view, so this function
// computes exactly that - the compact map `(owner << (32 - b)) | row` for the nodes this worker can reach and
// the list of nodes its own slots hold - in one pass over the rank list; the walk itself (which GPU takes which
// candidate) depends only on the weights, hence is identical in every worker.  Host-only, init-time.
#include <algorithm>
#include <cstdint>
#include <utility>
#include <vector>

#include "common.cuh"

extern "C" int xg_ics22_placement(const uint32_t *h_rank_nodes, const uint32_t *h_rank_weight, uint32_t num_node,
                                  uint32_t num_cached, uint32_t num_dev, uint32_t clique_size, double alpha,
                                  uint32_t worker, uint32_t compact_bitwidth, uint32_t *h_cache_table,
                                  uint32_t *h_slot_nodes, uint32_t *h_total_cached) {
  XG_REQUIRE(h_rank_nodes && h_rank_weight && h_cache_table && h_total_cached, "NULL argument");
  XG_REQUIRE(h_slot_nodes != nullptr || num_cached == 0, "h_slot_nodes is NULL");
  XG_REQUIRE(num_cached <= num_node, "num_cached > num_node");
  XG_REQUIRE(clique_size >= 1 && num_dev >= 1 && num_dev % clique_size == 0, "num_dev must be a multiple of clique_size");
  XG_REQUIRE(num_dev <= XG_MAX_PARTS && worker < num_dev, "worker out of range");
  XG_REQUIRE(compact_bitwidth >= 1 && compact_bitwidth <= 8 && num_dev <= (1u << compact_bitwidth),
             "compact_bitwidth cannot hold the device id");
  const uint32_t shift = 32u - compact_bitwidth;
  // cuda_cache_manager_host.cc:218: the row must fit under the device bits (and never collide with kEmptyKey)
  XG_REQUIRE(((uint64_t)num_cached >> shift) == 0, "num_cached does not fit the compact map");
  auto pack = [shift](uint32_t dev, uint32_t row) { return (dev << shift) | row; };

  std::fill(h_cache_table, h_cache_table + num_node, 0xFFFFFFFFu);
  for (uint32_t r = 0; r < num_cached; ++r) {
    h_cache_table[h_rank_nodes[r]] = pack(worker, r);
    h_slot_nodes[r] = h_rank_nodes[r];
  }
  uint32_t admitted = 0;
  if (clique_size > 1 && num_cached > 0) {
    const uint32_t c = clique_size, takers = c - 1;
    const uint32_t first = worker / c * c, me = worker - first;
    std::vector<double> load(c, 0.0);
    std::vector<std::pair<double, uint32_t>> by_load(c);
    std::vector<uint32_t> order(c);
    const uint32_t num_cand = num_node - num_cached;
    for (; admitted < num_cand; ++admitted) {
      const uint32_t turn = admitted % takers, back = admitted / takers;
      if (turn == 0) {  // a new slot: least-loaded GPUs first, ties by GPU index (ics22_song_dist_feature.cc:91-103)
        for (uint32_t d = 0; d < c; ++d) by_load[d] = {load[d], d};
        std::sort(by_load.begin(), by_load.end());
        for (uint32_t d = 0; d < c; ++d) order[d] = by_load[d].second;
      }
      if (back >= num_cached) break;  // no slot left to split
      const uint32_t slot = num_cached - 1 - back;
      const double w_cand = (double)h_rank_weight[num_cached + admitted];
      if (!(w_cand >= alpha * (double)h_rank_weight[slot])) break;  // :117
      const uint32_t taker = order[turn];
      load[taker] += w_cand;
      h_cache_table[h_rank_nodes[num_cached + admitted]] = pack(first + taker, slot);
      if (taker == me) {
        // my copy of the slot's original node is gone: read it from the GPU that keeps it (:130-134)
        h_cache_table[h_rank_nodes[slot]] = pack(first + order[c - 1], slot);
        h_slot_nodes[slot] = h_rank_nodes[num_cached + admitted];
      }
    }
  }
  *h_total_cached = num_cached + admitted;
  return XG_OK;
}
