/*
 * samgraph_abi.h -- the reference's own C-ABI (samgraph/common/operation.h:28-116), re-implemented over the
 * B200-native operators.  These are the symbols `samgraph/common/__init__.py:280-362` binds with ctypes; the
 * signatures, argument meaning and error behaviour are the reference's: every function is `void`/value-returning,
 * a violated requirement prints `file:line` and abort()s (logging.h:32-46, logging.cc:69-73).
 *
 * One addition replaces the pybind11 module samgraph/torch/c_lib (torch/adapter.cc:62-206), which cannot cross
 * a C ABI because it returns torch::Tensor: samgraph_get_graph_tensor() describes a batch tensor as
 * {pointer, shape, dtype, device} and samgraph_tensor_retain/release give the caller the same ownership the
 * reference's from_blob deleter has (the buffer lives as long as somebody holds it, adapter.cc:70-76).
 *
 * Scope: the data path of arch6 ("sgnn mode" = XGNN: data_init -> fork -> sample_init + train_init ->
 * sample_once / extract_start -> get_next_batch) and the single-process flow (init -> sample_once/start ->
 * get_next_batch) on one GPU.  Entry points that only exist for the out-of-scope deployment modes
 * (samgraph_um_sample_init: arch9, samgraph_switch_init: arch5 switcher) are exported and abort with a message.
 */
#ifndef SAMGRAPH_ABI_H
#define SAMGRAPH_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* operation.h:30-37 */
void samgraph_config(const char **config_keys, const char **config_values, const size_t num_config_items);
void samgraph_init(void);
void samgraph_start(void);
void samgraph_shutdown(void);
/* operation.h:39-55 */
size_t samgraph_num_epoch(void);
size_t samgraph_steps_per_epoch(void);
size_t samgraph_num_class(void);
size_t samgraph_feat_dim(void);
uint64_t samgraph_get_next_batch(void);
void samgraph_sample_once(void);
size_t samgraph_get_graph_num_src(uint64_t key, int graph_id);
size_t samgraph_get_graph_num_dst(uint64_t key, int graph_id);
size_t samgraph_get_graph_num_edge(uint64_t key, int graph_id);
/* operation.h:57-70: Profiler log items; enum numbering of profiler.h:30-140 (mirrored in
 * samgraph/common/__init__.py:148-276) */
void samgraph_log_step(uint64_t epoch, uint64_t step, int item, double val);
void samgraph_log_step_by_key(uint64_t key, int item, double val);
void samgraph_log_step_add(uint64_t epoch, uint64_t step, int item, double val);
void samgraph_log_epoch_add(uint64_t epoch, int item, double val);
double samgraph_get_log_init_value(int item);
double samgraph_get_log_step_value(uint64_t epoch, uint64_t step, int item);
double samgraph_get_log_step_value_by_key(uint64_t key, int item);
double samgraph_get_log_epoch_value(uint64_t epoch, int item);
/* operation.h:72-94 */
void samgraph_report_init(void);
void samgraph_report_step(uint64_t epoch, uint64_t step);
void samgraph_report_step_average(uint64_t epoch, uint64_t step);
void samgraph_report_epoch(uint64_t epoch);
void samgraph_report_epoch_average(uint64_t epoch);
void samgraph_report_node_access(void);
void samgraph_trace_step_begin(uint64_t key, int item, uint64_t ts);
void samgraph_trace_step_end(uint64_t key, int item, uint64_t ts);
void samgraph_trace_step_begin_now(uint64_t key, int item);
void samgraph_trace_step_end_now(uint64_t key, int item);
void samgraph_dump_trace(void);
void samgraph_forward_barrier(void);
/* operation.h:96-115: multi-GPU (arch6).  data_init in the parent BEFORE fork; sample_init + train_init in each
 * child with its own "cuda:N" context string. */
void samgraph_data_init(void);
void samgraph_sample_init(int worker_id, const char *ctx);
void samgraph_um_sample_init(int num_workers);
void samgraph_train_init(int worker_id, const char *ctx);
void samgraph_extract_start(int count);
void samgraph_switch_init(int worker_id, const char *ctx, double cache_percentage);
size_t samgraph_num_local_step(void);
int samgraph_wait_one_child(void);

/* ---- replaces samgraph/torch/adapter.cc:62-193 (samgraph_torch_get_graph_feat / label / row / col / data /
 * input_nodes / output_nodes and samgraph_torch_get_dataset_feat / label) --------------------------------- */
enum {
  SAMGRAPH_TENSOR_FEAT = 0,         /* [num_input, feat_dim], dataset dtype, trainer device   (adapter.cc:62-78)  */
  SAMGRAPH_TENSOR_LABEL = 1,        /* [num_output] i64, trainer device                       (adapter.cc:80-93)  */
  SAMGRAPH_TENSOR_ROW = 2,          /* [num_edge] i32, layer                                  (adapter.cc:95-107) */
  SAMGRAPH_TENSOR_COL = 3,          /* [num_edge] i32, layer                                  (adapter.cc:109-121)*/
  SAMGRAPH_TENSOR_DATA = 4,         /* [num_edge] i32, layer (random-walk visit counts)       (adapter.cc:123-135)*/
  SAMGRAPH_TENSOR_INPUT_NODES = 5,  /* [num_input] i32, sampler device                        (adapter.cc:165-178)*/
  SAMGRAPH_TENSOR_OUTPUT_NODES = 6, /* [num_output] i32, sampler device                       (adapter.cc:180-193)*/
  SAMGRAPH_TENSOR_DATASET_FEAT = 7, /* [num_node, feat_dim], host                             (adapter.cc:137-150)*/
  SAMGRAPH_TENSOR_DATASET_LABEL = 8,/* [num_node] i64, host                                   (adapter.cc:152-163)*/
  SAMGRAPH_TENSOR_CSC_INDPTR = 9    /* [num_dst + 1] i32, layer; KHop0/2/3 only.  (indptr, ROW) is the block in CSC
                                       form: replaces DGL's COO -> CSC conversion behind adapter.py:98-124;
                                       data == NULL when the layer has none (other samplers, duplicate seeds)     */
};
typedef struct samgraph_tensor_view {
  void *data;
  int64_t shape[2];
  int32_t ndim;
  int32_t dtype;   /* DataType numbering of common.h:38-47 (= XG_F32 ...) */
  int32_t device;  /* CUDA device ordinal, -1 = host memory */
  int32_t reserved;
  void *owner;     /* retained for the caller: pass to samgraph_tensor_release when the view dies (NULL = static) */
} samgraph_tensor_view;
/* `key` must equal the current batch's key (CHECK_EQ, adapter.cc:68); `layer` is ignored where it has no meaning */
void samgraph_get_graph_tensor(uint64_t key, int which, int layer, samgraph_tensor_view *out);
void samgraph_tensor_release(void *owner);

#ifdef __cplusplus
}
#endif
#endif /* SAMGRAPH_ABI_H */
