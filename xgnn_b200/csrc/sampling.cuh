// Uniform k-hop neighbour sampler (KHop0 / KHop2 / KHop3 contract), SURVEY.md rows a2-a4.
#pragma once
#include "common.cuh"
#include "hashtable.cuh"

namespace xg {

constexpr int kSampleBlock = 256;       // threads per CTA
constexpr int kSampleTile = 256;        // input nodes per tile (one per thread in the metadata phase)

// Optional fusion of the remap work into the sampler (used by the minibatch pipeline):
//   - out_src receives LOCAL ids (src_local[i], or i itself when src_local == NULL: the input list of
//     every layer but the seed layer IS the unique list, so position == local id);
//   - every sampled neighbour is inserted into the hash table with its final edge position, and the
//     slot position is written to pos_out.
struct SampleFusion {
  TableView table;
  uint32_t *pos_out;          // [num_input * fanout]
  const uint32_t *src_local;  // local id of input[i], or NULL for identity
  bool enabled;
};

constexpr uint32_t kSampleSalt = 0x5A4D4752u;  // third Philox counter word; the pipeline adds the layer

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
