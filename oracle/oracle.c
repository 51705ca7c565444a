/*
 * TEST INFRASTRUCTURE ONLY -- the CPU oracle for the sample -> unique/remap -> extract path.
 *
 * A plain-C restatement of the reference's algorithms, one function per reference routine, each
 * citing the reference file:line it follows (paths relative to /root/reference/samgraph/common).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library, and
 * only as the checker.  The product (xgnn_b200/) never links, loads or calls anything in here.
 *
 * Pinning status (see DESIGN.md "Oracle"):
 *   - khop0 (deg <= fanout branch), hashtable0, map_edges, extract: pinned against the reference's
 *     own CPU code compiled in place (oracle/_ref/libxgnn_ref.so) by tests/test_oracle.py (live) and
 *     against committed fixtures generated from it (tests/golden/*.npz).
 *   - hashtable properties: pinned against the reference's gtest vectors
 *     (samgraph/unittest/test_hashmap.cc:98-172).
 *   - khop0 deg > fanout: distribution is the contract.  The repo kernel is tested against the uniform law and against
 *     20 000 runs of the reference's CPU reservoir, whose inclusive RandomID(0, j + 1) law is derived and confirmed
 *     (tests/test_ref_cuda.py::test_khop0_against_the_reference_cpu_reservoir).
 *   - khop1 / weighted (alias, prefix, hash-dedup) / random-walk + top-K / khop3: the restatements below take the random
 *     words as an input (GPU == oracle replay, bit for bit) AND the GPU kernels are pinned against the reference's own CUDA
 *     kernels compiled in place (oracle/_ref/libxgnn_ref_cuda.so, tests/test_ref_cuda.py: exact where the draws do not
 *     matter, two-sample chi-square elsewhere).
 *   - presample ranking, ICS'22 placement (tests/oracle_lib.py restatement): pinned against the reference's own
 *     PreSampler / ICS22SongPlacementSolver compiled in place (oracle/ref_host_shim.cc; goldens + live).
 *   - mock extract (SAMGRAPH_EMPTY_FEAT): pinned against CPUMockExtract.
 *   - routing / partition layout / miss-cache split: no reference test or CPU code exists ("parity unpinned by
 *     the reference's tests"); restated from the cited lines and checked against feat[nodes] / the flat CSR.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stddef.h>
#include <math.h>

#define ORC_EMPTY 0xFFFFFFFFu /* Constant::kEmptyKey, constant.h:75 */

/* ------------------------------------------------------------------------------------------ */
/* Small deterministic RNG for the oracle's random branches (splitmix64).  The reference uses a
 * default-seeded thread_local std::mt19937 (cpu/cpu_random.cc:26-30); only uniformity matters. */
static inline uint64_t orc_mix(uint64_t *s) {
  uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
/* uniform integer in [0, n) without modulo bias */
static inline uint64_t orc_below(uint64_t *s, uint64_t n) {
  uint64_t lim = UINT64_MAX - (UINT64_MAX % n);
  uint64_t r;
  do { r = orc_mix(s); } while (r >= lim);
  return r % n;
}

/* stable removal of ORC_EMPTY entries; returns the kept count (std::remove_if in
 * cpu/cpu_sampling_khop0.cc:70-76) */
static size_t orc_compact(uint32_t *a, size_t n) {
  size_t w = 0;
  for (size_t i = 0; i < n; ++i)
    if (a[i] != ORC_EMPTY) a[w++] = a[i];
  return w;
}

/* ------------------------------------------------------------------------------------------ */
/* KHop0: cpu/cpu_sampling_khop0.cc:29-83 (GPU twin cuda/cuda_sampling_khop0.cu:102-153).
 * For every input node: degree <= fanout -> all neighbours in CSR order; otherwise Algorithm-R
 * reservoir of size fanout.  Slots are node-major / slot-minor, padding removed stably.
 * out_src/out_dst must hold num_input*fanout entries. */
size_t orc_sample_khop0(const uint32_t *indptr, const uint32_t *indices, const uint32_t *input,
                        size_t num_input, size_t fanout, uint32_t *out_src, uint32_t *out_dst,
                        uint64_t seed) {
  for (size_t i = 0; i < num_input; ++i) {
    const uint32_t v = input[i];
    const uint32_t begin = indptr[v];
    const uint32_t deg = indptr[v + 1] - begin;
    uint32_t *src = out_src + i * fanout;
    uint32_t *dst = out_dst + i * fanout;
    if (deg <= fanout) {
      for (size_t j = 0; j < fanout; ++j) {
        src[j] = j < deg ? v : ORC_EMPTY;
        dst[j] = j < deg ? indices[begin + j] : ORC_EMPTY;
      }
    } else {
      uint64_t s = seed ^ (0xD1B54A32D192ED03ull * (i + 1));
      for (size_t j = 0; j < fanout; ++j) {
        src[j] = v;
        dst[j] = indices[begin + j];
      }
      for (size_t j = fanout; j < deg; ++j) {
        uint64_t k = orc_below(&s, j + 1); /* RandomID(0, j+1) is inclusive [0, j+1]: see note */
        if (k < fanout) dst[k] = indices[begin + j];
      }
    }
  }
  size_t n = orc_compact(out_src, num_input * fanout);
  orc_compact(out_dst, num_input * fanout);
  return n;
}
/* Note on the reservoir: the reference draws RandomID(0, j+1) with an *inclusive*
 * std::uniform_int_distribution (cpu/cpu_random.cc:27-29), i.e. k in [0, j+1] -- one value too
 * many, which biases late elements slightly low.  The stated contract (BASELINE.json north_star;
 * SURVEY.md section 9 item 12) is the uniform without-replacement law, so the oracle uses the
 * textbook k in [0, j]. */

/* per-node output count min(deg, fanout): the size law shared by KHop0/2/3. */
void orc_sample_counts(const uint32_t *indptr, const uint32_t *input, size_t num_input,
                       size_t fanout, uint32_t *counts) {
  for (size_t i = 0; i < num_input; ++i) {
    uint32_t deg = indptr[input[i] + 1] - indptr[input[i]];
    counts[i] = deg < fanout ? deg : (uint32_t)fanout;
  }
}

/* ------------------------------------------------------------------------------------------ */
/* Ordered hash table, first-occurrence numbering: cpu/cpu_hashtable0.cc:37-66.  This is also one
 * legal outcome of the GPU table (cuda/cuda_hashtable.cu:744-912), whose winner among duplicate
 * occurrences is racy; it is the only deterministic one. */
typedef struct {
  uint32_t *keys;   /* open addressing, ORC_EMPTY = free */
  uint32_t *vals;   /* local id */
  uint32_t *n2o;    /* local id -> global id */
  size_t cap;       /* power of two */
  size_t n2o_cap;
  size_t num_items;
} orc_table;

static size_t orc_pow2(size_t n) { size_t c = 16; while (c < n) c <<= 1; return c; }
static inline size_t orc_hash(uint32_t k, size_t mask) {
  uint32_t h = k * 0x9E3779B1u;
  return (h ^ (h >> 15)) & mask;
}

orc_table *orc_table_create(size_t max_items) {
  orc_table *t = (orc_table *)calloc(1, sizeof(orc_table));
  t->cap = orc_pow2(max_items * 2 + 16);
  t->n2o_cap = max_items + 16;
  t->keys = (uint32_t *)malloc(t->cap * sizeof(uint32_t));
  t->vals = (uint32_t *)malloc(t->cap * sizeof(uint32_t));
  t->n2o = (uint32_t *)malloc(t->n2o_cap * sizeof(uint32_t));
  memset(t->keys, 0xFF, t->cap * sizeof(uint32_t));
  return t;
}
void orc_table_destroy(orc_table *t) {
  if (!t) return;
  free(t->keys); free(t->vals); free(t->n2o); free(t);
}
/* cpu_hashtable0.cc:61-64 */
void orc_table_reset(orc_table *t) {
  memset(t->keys, 0xFF, t->cap * sizeof(uint32_t));
  t->num_items = 0;
}
size_t orc_table_num_items(const orc_table *t) { return t->num_items; }

static int orc_table_find(const orc_table *t, uint32_t key, uint32_t *val) {
  size_t mask = t->cap - 1, p = orc_hash(key, mask);
  while (t->keys[p] != ORC_EMPTY) {
    if (t->keys[p] == key) { *val = t->vals[p]; return 1; }
    p = (p + 1) & mask;
  }
  return 0;
}
/* cpu_hashtable0.cc:37-47 : sequential insert, a new key gets id = num_items++ */
int orc_table_populate(orc_table *t, const uint32_t *input, size_t n) {
  size_t mask = t->cap - 1;
  for (size_t i = 0; i < n; ++i) {
    uint32_t key = input[i];
    size_t p = orc_hash(key, mask);
    int found = 0;
    while (t->keys[p] != ORC_EMPTY) {
      if (t->keys[p] == key) { found = 1; break; }
      p = (p + 1) & mask;
    }
    if (!found) {
      if (t->num_items >= t->n2o_cap) return -1;
      t->keys[p] = key;
      t->vals[p] = (uint32_t)t->num_items;
      t->n2o[t->num_items++] = key;
    }
  }
  return 0;
}
/* cpu_hashtable0.cc:49-51 */
void orc_table_map_nodes(const orc_table *t, uint32_t *out, size_t n) {
  memcpy(out, t->n2o, n * sizeof(uint32_t));
}
/* cpu_hashtable0.cc:53-66 and cuda/cuda_mapping.cu:50-80 ; returns -1 when a key is absent
 * (the reference CHECK-aborts) */
int orc_table_map_edges(const orc_table *t, const uint32_t *src, const uint32_t *dst, size_t len,
                        uint32_t *new_src, uint32_t *new_dst) {
  for (size_t i = 0; i < len; ++i) {
    if (!orc_table_find(t, src[i], &new_src[i])) return -1;
    if (!orc_table_find(t, dst[i], &new_dst[i])) return -1;
  }
  return 0;
}
/* DeviceOrderedHashTable::SearchO2N(id)->local, cuda/cuda_hashtable.h:57-60; ORC_EMPTY if absent */
void orc_table_lookup(const orc_table *t, const uint32_t *keys, size_t n, uint32_t *local) {
  for (size_t i = 0; i < n; ++i)
    if (!orc_table_find(t, keys[i], &local[i])) local[i] = ORC_EMPTY;
}

/* ------------------------------------------------------------------------------------------ */
/* Row gather: cpu/cpu_extraction.cc:31-45,66-90 and cuda/cuda_extraction.cu:31-49.  A byte copy for
 * every dtype, so it is stated on row_bytes = dim * sizeof(T). */
void orc_extract(void *dst, const void *src, const uint32_t *index, size_t num_index,
                 size_t row_bytes) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < num_index; ++i)
    memcpy((char *)dst + i * row_bytes, (const char *)src + (size_t)index[i] * row_bytes, row_bytes);
}
/* mock variant: cpu/cpu_extraction.cc:47-62 -- the *row index* is masked to 2^k rows */
void orc_mock_extract(void *dst, const void *src, const uint32_t *index, size_t num_index,
                      size_t row_bytes, unsigned empty_feat_bits) {
  size_t mask = ((size_t)1 << empty_feat_bits) - 1;
  for (size_t i = 0; i < num_index; ++i)
    memcpy((char *)dst + i * row_bytes, (const char *)src + ((size_t)index[i] & mask) * row_bytes,
           row_bytes);
}

/* ------------------------------------------------------------------------------------------ */
/* GGMS topology routing: DeviceDistGraph::_GetRealPartId, cuda/dist_graph.h:141-151.
 * part == num_part means "host whole-graph slot". */
void orc_route_node(uint32_t v, uint32_t num_part, uint32_t num_cache_node, uint32_t *part,
                    uint32_t *row) {
  if (v < num_cache_node) { *part = v % num_part; *row = v / num_part; }
  else { *part = num_part; *row = v; }
}
/* num_cache_node: smallest n with indptr[n] >= floor(E * ratio), dist/dist_engine.cc:223-232 */
uint32_t orc_num_cache_node(const uint32_t *indptr, uint32_t num_node, double ratio) {
  uint64_t want = (uint64_t)((double)indptr[num_node] * ratio);
  uint32_t n = 0;
  while (n < num_node && indptr[n] < want) ++n;
  return n;
}
/* partition p's CSR, rebased to 0: DistGraph::_DatasetPartition, cuda/dist_graph.cu:228-272.
 * Sizes: indptr has ceil((n-p)/P)+1 entries.  Call with part_indices==NULL to size. Returns the
 * partition's edge count; *indptr_len receives the indptr length. */
size_t orc_partition_csr(const uint32_t *indptr, const uint32_t *indices, uint32_t num_cache_node,
                         uint32_t part, uint32_t num_part, uint32_t *part_indptr,
                         uint32_t *part_indices, size_t *indptr_len) {
  size_t rows = num_cache_node / num_part + (part < num_cache_node % num_part ? 1 : 0);
  size_t e = 0, r = 0;
  for (uint32_t v = part; v < num_cache_node; v += num_part, ++r) {
    uint32_t deg = indptr[v + 1] - indptr[v];
    if (part_indptr) part_indptr[r] = (uint32_t)e;
    if (part_indices) memcpy(part_indices + e, indices + indptr[v], deg * sizeof(uint32_t));
    e += deg;
  }
  if (part_indptr) part_indptr[rows] = (uint32_t)e;
  if (indptr_len) *indptr_len = rows + 1;
  return e;
}
/* KHop0 over a partitioned topology: same law as orc_sample_khop0 but every neighbour list is
 * reached through the routing above (cuda_sampling_khop0.cu:119-121 with GraphType =
 * DeviceDistGraph).  part_indptr/part_indices have num_part+1 entries, the last being the whole
 * host graph.  Used to prove the partitioned store returns the same lists as the flat CSR. */
size_t orc_sample_khop0_dist(const uint32_t *const *part_indptr, const uint32_t *const *part_indices,
                             uint32_t num_part, uint32_t num_cache_node, const uint32_t *input,
                             size_t num_input, size_t fanout, uint32_t *out_src, uint32_t *out_dst,
                             uint64_t seed) {
  for (size_t i = 0; i < num_input; ++i) {
    uint32_t v = input[i], part, row;
    orc_route_node(v, num_part, num_cache_node, &part, &row);
    const uint32_t begin = part_indptr[part][row];
    const uint32_t deg = part_indptr[part][row + 1] - begin;
    const uint32_t *edges = part_indices[part] + begin;
    uint32_t *src = out_src + i * fanout, *dst = out_dst + i * fanout;
    if (deg <= fanout) {
      for (size_t j = 0; j < fanout; ++j) {
        src[j] = j < deg ? v : ORC_EMPTY;
        dst[j] = j < deg ? edges[j] : ORC_EMPTY;
      }
    } else {
      uint64_t s = seed ^ (0xD1B54A32D192ED03ull * (i + 1));
      for (size_t j = 0; j < fanout; ++j) { src[j] = v; dst[j] = edges[j]; }
      for (size_t j = fanout; j < deg; ++j) {
        uint64_t k = orc_below(&s, j + 1);
        if (k < fanout) dst[k] = edges[j];
      }
    }
  }
  size_t n = orc_compact(out_src, num_input * fanout);
  orc_compact(out_dst, num_input * fanout);
  return n;
}

/* ------------------------------------------------------------------------------------------ */
/* Feature cache: direct-address table node -> cache rank (cuda/cuda_cache_manager_host.cc:197-229)
 * rank_nodes[0..num_cached) is the (already shuffled) cached prefix of the policy's rank list. */
void orc_build_cache_table(const uint32_t *rank_nodes, size_t num_cached, uint32_t *table,
                           size_t num_node) {
  for (size_t i = 0; i < num_node; ++i) table[i] = ORC_EMPTY;
  for (size_t i = 0; i < num_cached; ++i) table[rank_nodes[i]] = (uint32_t)i;
}
/* stable split of the batch into misses (src = node id) and hits (src = cache rank), dst = position
 * in the batch: GPUCacheManager::GetMissCacheIndex, cuda/cuda_cache_manager_device.cu:41-169,355-441 */
void orc_miss_cache_index(const uint32_t *table, const uint32_t *nodes, size_t n,
                          uint32_t *miss_src, uint32_t *miss_dst, size_t *num_miss,
                          uint32_t *cache_src, uint32_t *cache_dst, size_t *num_cache) {
  size_t m = 0, c = 0;
  for (size_t i = 0; i < n; ++i) {
    uint32_t r = table[nodes[i]];
    if (r == ORC_EMPTY) { miss_src[m] = nodes[i]; miss_dst[m++] = (uint32_t)i; }
    else { cache_src[c] = r; cache_dst[c++] = (uint32_t)i; }
  }
  *num_miss = m; *num_cache = c;
}
/* rows of partition p of the feature cache: rank r -> part r%P, row r/P
 * (cuda/dist_graph.cu:493-505, cuda/dist_graph.h:203-207).  Returns number of rows. */
size_t orc_partition_features(const void *feat, size_t row_bytes, const uint32_t *rank_nodes,
                              size_t num_cached, uint32_t part, uint32_t num_part, void *out) {
  size_t r = 0;
  for (size_t i = part; i < num_cached; i += num_part, ++r)
    if (out) memcpy((char *)out + r * row_bytes, (const char *)feat + (size_t)rank_nodes[i] * row_bytes,
                    row_bytes);
  return r;
}
/* miss rows from the host table, hit rows from the partitioned cache:
 * extract_miss_data + combine_cache_data_for_partition, cuda_cache_manager_device.cu:234-299 */
void orc_extract_ggms(void *out, size_t row_bytes, const uint32_t *nodes, size_t n,
                      const uint32_t *table, const void *host_feat, const void *const *part_feat,
                      uint32_t num_part) {
  for (size_t i = 0; i < n; ++i) {
    uint32_t r = table ? table[nodes[i]] : ORC_EMPTY;
    const char *src = r == ORC_EMPTY
                          ? (const char *)host_feat + (size_t)nodes[i] * row_bytes
                          : (const char *)part_feat[r % num_part] + (size_t)(r / num_part) * row_bytes;
    memcpy((char *)out + i * row_bytes, src, row_bytes);
  }
}

/* ------------------------------------------------------------------------------------------ */
/* One minibatch, the call order of DoCPUSample (cpu/cpu_loops.cc:55-192) == dist::DoGPUSample
 * (dist/dist_loops.cc:62-368): table reset, seeds inserted, then layers last -> first:
 * sample, insert out_dst, unique list (previous input is its prefix), remap (src,dst).
 * Layer l outputs: row[l] = local(out_dst) (neighbour), col[l] = local(out_src) (seed side),
 * counts[3l..3l+2] = {num_src = num_unique, num_dst = num_input, num_edge}.
 * Buffers row[l]/col[l] must hold PredictNumNodes-style worst case; pass scratch of the same size.
 * Returns |input_nodes| (written to input_nodes). */
size_t orc_minibatch(const uint32_t *indptr, const uint32_t *indices, const uint32_t *seeds,
                     size_t num_seeds, const size_t *fanouts, int num_layers, uint64_t seed,
                     uint32_t **row, uint32_t **col, size_t *counts, uint32_t *input_nodes,
                     size_t max_nodes) {
  orc_table *t = orc_table_create(max_nodes);
  orc_table_populate(t, seeds, num_seeds);
  uint32_t *cur = (uint32_t *)malloc(max_nodes * sizeof(uint32_t));
  memcpy(cur, seeds, num_seeds * sizeof(uint32_t));
  size_t num_input = num_seeds;
  for (int l = num_layers - 1; l >= 0; --l) {
    size_t f = fanouts[l];
    uint32_t *src = (uint32_t *)malloc((num_input * f + 1) * sizeof(uint32_t));
    uint32_t *dst = (uint32_t *)malloc((num_input * f + 1) * sizeof(uint32_t));
    size_t e = orc_sample_khop0(indptr, indices, cur, num_input, f, src, dst,
                                seed + 0x1000003ull * (uint64_t)l);
    orc_table_populate(t, dst, e);
    size_t u = orc_table_num_items(t);
    orc_table_map_edges(t, src, dst, e, col[l], row[l]);
    counts[3 * l + 0] = u; counts[3 * l + 1] = num_input; counts[3 * l + 2] = e;
    orc_table_map_nodes(t, cur, u);
    num_input = u;
    free(src); free(dst);
  }
  memcpy(input_nodes, cur, num_input * sizeof(uint32_t));
  free(cur);
  orc_table_destroy(t);
  return num_input;
}

/* ========================================================================================== */
/* With-replacement / weighted samplers and random walk.  The reference has NO CPU twin for these
 * (cpu/cpu_sampling_khop1.cc:24-27, cpu_sampling_weighted_khop.cc:24-28, cpu_sampling_random_walk.cc
 * :24-28 are empty stubs), so these functions restate the CUDA files statement by statement, executed
 * sequentially in thread-index order.  Random numbers are an INPUT (`rnd`, four 32-bit words per draw
 * slot) so that a test can feed the very words the kernel under test consumed and compare bit for bit;
 * how a word becomes a choice follows the reference (`curand() % len`, `curand_uniform() < prob`).
 * "parity unpinned by the reference's tests": nothing in the reference pins these outputs. */

static float orc_uniform_open0(uint32_t r) { return (float)((r >> 8) + 1u) * (1.0f / 16777216.0f); }
static double orc_uniform_double(uint32_t r) { return ((double)r + 0.5) * (1.0 / 4294967296.0); }

typedef struct { uint32_t src, dst; size_t idx; } orc_pair;
static int orc_pair_cmp(const void *a, const void *b) {
  const orc_pair *x = (const orc_pair *)a, *y = (const orc_pair *)b;
  if (x->src != y->src) return x->src < y->src ? -1 : 1;
  return x->idx < y->idx ? -1 : (x->idx > y->idx);   /* stable: cub::DeviceRadixSort::SortPairs on src only */
}

/* sort by src (stable) + keep the LAST of every run of equal (src,dst), drop padding:
 * cuda/cuda_sampling_khop1.cu:74-128,170-224 (identical in the weighted variants) */
static size_t orc_sort_dedup_adjacent(uint32_t *tmp_src, uint32_t *tmp_dst, size_t num, uint32_t *out_src,
                                      uint32_t *out_dst) {
  orc_pair *p = (orc_pair *)malloc((num + 1) * sizeof(orc_pair));
  for (size_t t = 0; t < num; ++t) { p[t].src = tmp_src[t]; p[t].dst = tmp_dst[t]; p[t].idx = t; }
  qsort(p, num, sizeof(orc_pair), orc_pair_cmp);
  size_t w = 0;
  for (size_t t = 0; t < num; ++t) {
    int keep;
    if (t + 1 < num) keep = (p[t].src != p[t + 1].src || p[t].dst != p[t + 1].dst) && p[t].src != ORC_EMPTY;
    else keep = p[t].src != ORC_EMPTY;
    if (keep) { out_src[w] = p[t].src; out_dst[w] = p[t].dst; ++w; }
  }
  free(p);
  return w;
}

/* mode 1 = KHop1 (cuda_sampling_khop1.cu:42-72), 2 = alias (cuda_sampling_weighted_khop.cu:41-76),
 * 4 = prefix (cuda_sampling_weighted_khop_prefix.cu:41-92).  rnd[(i*fanout+s)*4 + {0,1}] */
size_t orc_sample_with_replacement(int mode, const uint32_t *indptr, const uint32_t *indices, const float *prob,
                                   const uint32_t *alias, const float *prefix, const uint32_t *input,
                                   size_t num_input, size_t fanout, const uint32_t *rnd, uint32_t *out_src,
                                   uint32_t *out_dst) {
  size_t num = num_input * fanout;
  uint32_t *ts = (uint32_t *)malloc((num + 1) * sizeof(uint32_t));
  uint32_t *td = (uint32_t *)malloc((num + 1) * sizeof(uint32_t));
  for (size_t t = 0; t < num; ++t) {
    const uint32_t rid = input[t / fanout];
    const uint32_t off = indptr[rid], len = indptr[rid + 1] - indptr[rid];
    const uint32_t r0 = rnd[t * 4 + 0], r1 = rnd[t * 4 + 1];
    ts[t] = ORC_EMPTY; td[t] = ORC_EMPTY;
    if (len == 0) continue;
    ts[t] = rid;
    if (mode == 1) {
      td[t] = indices[off + r0 % len];
    } else if (mode == 2) {
      size_t k = r0 % len;
      float r = orc_uniform_open0(r1);
      td[t] = (r < prob[off + k]) ? indices[off + k] : alias[off + k];
    } else {
      const float up = prefix[off + len - 1];
      float x = orc_uniform_open0(r0) * up;
      if (x <= prefix[off]) td[t] = indices[off];
      else {
        size_t lo = off, hi = off + len - 1;
        while (hi - lo >= 2) {
          size_t mid = (lo + hi) >> 1;
          if (prefix[mid] >= x) hi = mid; else lo = mid;
        }
        td[t] = indices[hi];
      }
    }
  }
  size_t n = orc_sort_dedup_adjacent(ts, td, num, out_src, out_dst);
  free(ts); free(td);
  return n;
}

/* cuda_sampling_weighted_khop_hash_dedup.cu:59-116 (+ count/compact :118-200): node-major, no sort.
 * rnd[(i*budget + try)*4 + {0,1}], budget = tries_per_slot * fanout draws at most (the reference has no
 * bound and spins forever when a node has fewer than `fanout` distinct outcomes). */
size_t orc_sample_weighted_hash_dedup(const uint32_t *indptr, const uint32_t *indices, const float *prob,
                                      const uint32_t *alias, const uint32_t *input, size_t num_input,
                                      size_t fanout, size_t tries_per_slot, const uint32_t *rnd,
                                      uint32_t *out_src, uint32_t *out_dst) {
  size_t w = 0, budget = tries_per_slot * fanout;
  uint32_t *seen = (uint32_t *)malloc((fanout + 1) * sizeof(uint32_t));
  for (size_t i = 0; i < num_input; ++i) {
    const uint32_t rid = input[i];
    const uint32_t off = indptr[rid], len = indptr[rid + 1] - indptr[rid];
    if (len <= fanout) {
      for (size_t j = 0; j < len; ++j) { out_src[w] = rid; out_dst[w] = indices[off + j]; ++w; }
      continue;
    }
    size_t have = 0;
    for (size_t tries = 0; have < fanout && tries < budget; ++tries) {
      const uint32_t r0 = rnd[(i * budget + tries) * 4 + 0], r1 = rnd[(i * budget + tries) * 4 + 1];
      size_t k = r0 % len;
      float r = orc_uniform_open0(r1);
      uint32_t res = indices[off + k];
      if (r > prob[off + k]) res = alias[off + k];
      int dup = 0;
      for (size_t q = 0; q < have; ++q) dup |= (seen[q] == res);
      if (dup) continue;
      seen[have++] = res;
      out_src[w] = rid; out_dst[w] = res; ++w;
    }
  }
  free(seen);
  return w;
}

/* cuda_sampling_random_walk.cu:43-112 + FrequencyHashmap::GetTopK cuda_frequency_hashmap.cu:643-841.
 * rnd[((i*num_walk + w)*walk_len + t)*4 + {0,1}].  Visit layout per seed: pos = t*num_walk + w (:66-80).
 * Unique (seed, node) entries are listed in ascending order of the visit that created them (the bucket's
 * `index`; sequential execution => the first visit), then sorted stably, descending, on the 64-bit key
 * ((N - i) << 32 | count) (:737-746) and the first min(K, #unique) per seed are emitted (:500-532). */
typedef struct { uint64_t key; uint32_t dst; size_t ord; } orc_freq;
static int orc_freq_cmp(const void *a, const void *b) {
  const orc_freq *x = (const orc_freq *)a, *y = (const orc_freq *)b;
  if (x->key != y->key) return x->key > y->key ? -1 : 1;
  return x->ord < y->ord ? -1 : (x->ord > y->ord);
}
size_t orc_random_walk_topk(const uint32_t *indptr, const uint32_t *indices, const uint32_t *input,
                            size_t num_input, size_t walk_len, double restart_prob, size_t num_walk, size_t K,
                            const uint32_t *rnd, uint32_t *out_src, uint32_t *out_dst, uint32_t *out_data) {
  const size_t M = walk_len * num_walk;
  uint32_t *vis = (uint32_t *)malloc((M + 1) * sizeof(uint32_t));
  orc_freq *fr = (orc_freq *)malloc((M + 1) * sizeof(orc_freq));
  size_t w_out = 0;
  for (size_t i = 0; i < num_input; ++i) {
    for (size_t w = 0; w < num_walk; ++w) {
      uint32_t node = input[i];
      for (size_t t = 0; t < walk_len; ++t) {
        size_t pos = t * num_walk + w;
        vis[pos] = ORC_EMPTY;
        if (node == ORC_EMPTY) continue;
        const uint32_t off = indptr[node], len = indptr[node + 1] - indptr[node];
        if (len == 0) { node = ORC_EMPTY; continue; }
        const uint32_t *r = rnd + ((i * num_walk + w) * walk_len + t) * 4;
        size_t k = r[0] % len;
        vis[pos] = indices[off + k];
        node = indices[off + k];
        if (orc_uniform_double(r[1]) < restart_prob) node = ORC_EMPTY;
      }
    }
    size_t nu = 0;
    for (size_t m = 0; m < M; ++m) {
      if (vis[m] == ORC_EMPTY) continue;
      size_t q = 0;
      for (; q < nu; ++q) if (fr[q].dst == vis[m]) break;
      if (q == nu) { fr[nu].dst = vis[m]; fr[nu].key = 0; fr[nu].ord = nu; ++nu; }
      fr[q].key += 1;
    }
    for (size_t q = 0; q < nu; ++q) fr[q].key |= ((uint64_t)(num_input - i)) << 32;
    qsort(fr, nu, sizeof(orc_freq), orc_freq_cmp);
    size_t take = nu < K ? nu : K;
    for (size_t q = 0; q < take; ++q) {
      out_src[w_out] = input[i]; out_dst[w_out] = fr[q].dst; out_data[w_out] = (uint32_t)fr[q].key; ++w_out;
    }
  }
  free(vis); free(fr);
  return w_out;
}
