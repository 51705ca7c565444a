// Uniform k-hop neighbour sampler (KHop0 / KHop2 / KHop3 contract), SURVEY.md rows a2-a4.
#pragma once
#include "common.cuh"
#include "hashtable.cuh"

namespace xg {

constexpr int kSampleBlock = 256;       // threads per CTA
constexpr int kSampleTile = 256;        // input nodes per tile (one per thread in the metadata phase)

// Optional fusion of the src-side remap into the sampler (used by the minibatch pipeline):
//   - out_src receives LOCAL ids (src_local[i], or i itself when src_local == NULL: the input list of
//     every layer but the seed layer IS the unique list, so position == local id);
//   - csc_indptr receives the block's CSC offsets.
// (r01 also inserted every sampled neighbour into the hash table from inside the sampler.  With the 4-byte-slot table
// a probe reads keys back through their edge positions, i.e. out_dst of OTHER threads of the same launch, which would
// need a device-wide fence per edge; the insert is a kernel of its own again - one more coalesced pass over 17 MB.)
struct SampleFusion {
  const uint32_t *src_local;  // local id of input[i], or NULL for identity
  uint32_t *csc_indptr;       // [num_input + 1] or NULL: edge offset of every input node = the block's CSC indptr
  bool enabled;
};

constexpr uint32_t kSampleSalt = 0x5A4D4752u;  // third Philox counter word; the pipeline adds the layer

#ifdef __CUDACC__
// DeviceNormalGraph / DeviceDistGraph::operator[] + NumEdge (cuda/dist_graph.h:114-180): which CSR holds
// node v and at which row.  The handle is a kernel parameter, so the pointer table is in constant memory.
struct NodeRoute {
  const uint32_t *indptr;
  const uint32_t *indices;
  uint32_t row;
};

__device__ __forceinline__ NodeRoute route(const xg_graph &g, uint32_t v) {
  NodeRoute r;
  if (g.num_part == 0) {
    r.indptr = g.part_indptr[0];
    r.indices = g.part_indices[0];
    r.row = v;
  } else if (v < g.num_cache_node) {
    const uint32_t part = v % g.num_part;
    r.indptr = g.part_indptr[part];
    r.indices = g.part_indices[part];
    r.row = v / g.num_part;
  } else {  // host-resident tail of the graph: slot [num_part] holds the whole CSR
    r.indptr = g.part_indptr[g.num_part];
    r.indices = g.part_indices[g.num_part];
    r.row = v;
  }
  return r;
}
#endif

// Philox key of one minibatch: (seed, batch_key) -> 64 bits
inline uint2 philox_key(uint64_t seed, uint64_t batch_key) {
  const uint64_t k64 = seed ^ (batch_key * 0x9E3779B97F4A7C15ull + 0x7F4A7C15F39CC060ull);
  return make_uint2((uint32_t)k64, (uint32_t)(k64 >> 32));
}

// d_rng_key (device, may be NULL): overrides (seed, batch_key) -- the key of a graph-replayed batch.
// d_num_input (device, may be NULL): actual number of input nodes (<= num_input_max); lets the
// pipeline enqueue a layer before the previous layer's unique count is known on the host.
int launch_sample_uniform(const xg_graph &g, const uint32_t *d_input, size_t num_input_max,
                          const uint32_t *d_num_input, size_t fanout, uint32_t *d_out_src,
                          uint32_t *d_out_dst, uint64_t *d_num_out, uint64_t seed, uint64_t batch_key,
                          const uint2 *d_rng_key, uint32_t salt, const SampleFusion &fusion,
                          const ScanSpace &scan, cudaStream_t s);

}  // namespace xg

namespace xg {

// ---- with-replacement / weighted samplers (sampling_ext.cu) and random walk (randomwalk.cu) ----------
// Scratch of the "extended" samplers; sized by ext_reserve (allocates: call it outside graph capture).
struct ExtWorkspace {
  uint32_t *keys_in = nullptr, *keys_out = nullptr;  // [cap_nodes] input ids (EMPTY-padded) / sorted
  uint32_t *vals_in = nullptr, *vals_out = nullptr;  // [cap_nodes] input positions / in sorted order
  uint32_t *tmp_src = nullptr, *tmp_dst = nullptr;   // [cap_slots] padded draws
  void *sort_temp = nullptr;
  size_t sort_bytes = 0, cap_nodes = 0, cap_slots = 0;
  void *scan_mem = nullptr;  // ticket + tile states for the compaction scan
  size_t scan_tiles = 0;
};
int ext_reserve(ExtWorkspace &ws, size_t num_nodes, size_t slots_per_node, cudaStream_t s);
void ext_release(ExtWorkspace &ws);

struct WeightTables {           // aligned with the flat CSR's `indices`
  const float *prob = nullptr;     // alias method: acceptance probability (prob_table.bin)
  const uint32_t *alias = nullptr; // alias method: alias NODE ID (alias_table.bin)
  const float *prefix = nullptr;   // inverse CDF: per-node running sums (prob_prefix_table.bin)
};

// KHop1 / WeightedKHop / WeightedKHopPrefix / WeightedKHopHashDedup.  out_src receives global ids, or
// (emit_local) the local id of the expanded node: src_local[pos] when given, else its input position.
int launch_sample_ext(int sample_type, const xg_graph &g, const WeightTables &w, const uint32_t *d_input,
                      size_t num_input_max, const uint32_t *d_num_input, size_t fanout, bool emit_local,
                      const uint32_t *src_local, uint32_t *d_out_src, uint32_t *d_out_dst, uint64_t *d_num_out,
                      uint2 rng_key, const uint2 *d_rng_key, uint32_t salt, ExtWorkspace &ws, cudaStream_t s);

constexpr uint32_t kHashDedupTriesPerSlot = 16;  // rejection budget of the hash-dedup sampler: 16 * fanout draws

// uniform in (0, 1] from 32 random bits with 24-bit resolution (curand_uniform's range)
__host__ __device__ inline float uniform_open0(uint32_t r) { return (float)((r >> 8) + 1u) * (1.0f / 16777216.0f); }
// uniform in (0, 1) as double from 32 random bits (stand-in for curand_uniform_double)
__host__ __device__ inline double uniform_double(uint32_t r) { return ((double)r + 0.5) * (1.0 / 4294967296.0); }

int launch_random_walk(const xg_graph &g, const uint32_t *d_input, size_t num_input_max, const uint32_t *d_num_input,
                       uint32_t walk_length, double restart_prob, uint32_t num_walk, uint32_t K, bool emit_local,
                       const uint32_t *src_local, uint32_t *d_out_src, uint32_t *d_out_dst, uint32_t *d_out_data,
                       uint64_t *d_num_out, uint2 rng_key, const uint2 *d_rng_key, uint32_t salt,
                       const ScanSpace &scan, cudaStream_t s);

}  // namespace xg
