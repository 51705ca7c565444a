"""Mirror of samgraph/common/__init__.py: enum values, helper tables and the ctypes wrapper class the
framework adapters share (SamGraphBasics, :278-574).  Same names, same argument meaning."""
from __future__ import annotations

import ctypes

from .._clib import lib as _lib

# ---- samgraph/common/__init__.py:46-76 ----
kCPU, kMMAP, kGPU = 0, 1, 2
kKHop0, kKHop1, kWeightedKHop, kRandomWalk, kWeightedKHopPrefix, kKHop2, kWeightedKHopHashDedup, kKHop3 = range(8)
kArch0, kArch1, kArch2, kArch3, kArch4, kArch5, kArch6, kArch7 = range(8)
(kCacheByDegree, kCacheByHeuristic, kCacheByPreSample, kCacheByDegreeHop, kCacheByPreSampleStatic, kCacheByFakeOptimal,
 kDynamicCache, kCacheByRandom) = range(8)


def cpu(device_id=0):
    return 'cpu:{:}'.format(device_id)


def gpu(device_id=0):
    return 'cuda:{:}'.format(device_id)


sample_types = {'khop0': kKHop0, 'khop1': kKHop1, 'khop2': kKHop2, 'khop3': kKHop3, 'random_walk': kRandomWalk,
                'weighted_khop': kWeightedKHop, 'weighted_khop_prefix': kWeightedKHopPrefix,
                'weighted_khop_hash_dedup': kWeightedKHopHashDedup}
builtin_archs = {
    'arch0': {'arch': kArch0, 'sampler_ctx': cpu(), 'trainer_ctx': gpu(0)},
    'arch1': {'arch': kArch1, 'sampler_ctx': gpu(0), 'trainer_ctx': gpu(0)},
    'arch2': {'arch': kArch2, 'sampler_ctx': gpu(0), 'trainer_ctx': gpu(0)},
    'arch3': {'arch': kArch3, 'sampler_ctx': gpu(0), 'trainer_ctx': gpu(1)},
    'arch4': {'arch': kArch4, 'sampler_ctx': gpu(1), 'trainer_ctx': gpu(0)},
    'arch5': {'arch': kArch5}, 'arch6': {'arch': kArch6}, 'arch7': {'arch': kArch7},
}
cache_policies = {'degree': kCacheByDegree, 'heuristic': kCacheByHeuristic, 'pre_sample': kCacheByPreSample,
                  'degree_hop': kCacheByDegreeHop, 'presample_static': kCacheByPreSampleStatic,
                  'fake_optimal': kCacheByFakeOptimal, 'dynamic_cache': kDynamicCache, 'random': kCacheByRandom}

# ---- log items read back by the scripts (profiler.h:30-140 <-> samgraph/common/__init__.py:148-276) ----
kLogInitL1Common, kLogInitL1Sampler, kLogInitL1Trainer = 0, 1, 2
kLogInitL2LoadDataset, kLogInitL2BuildCache = 6, 10
(kLogL1NumSample, kLogL1NumNode, kLogL1SampleTotalTime, kLogL1SampleTime, kLogL1SendTime, kLogL1RecvTime, kLogL1CopyTime,
 kLogL1ConvertTime, kLogL1TrainTime, kLogL1FeatureBytes, kLogL1LabelBytes, kLogL1IdBytes, kLogL1GraphBytes,
 kLogL1MissBytes, kLogL1PrefetchAdvanced, kLogL1GetNeighbourTime, kLogL1SamplerId, kLogL2ShuffleTime, kLogL2LastLayerTime,
 kLogL2LastLayerSize, kLogL2CoreSampleTime, kLogL2IdRemapTime, kLogL2GraphCopyTime, kLogL2IdCopyTime, kLogL2ExtractTime,
 kLogL2FeatCopyTime, kLogL2CacheCopyTime) = range(27)
(kLogEpochSampleTime, KLogEpochSampleGetCacheMissIndexTime, kLogEpochSampleSendTime, kLogEpochSampleTotalTime,
 kLogEpochCoreSampleTime, kLogEpochSampleCooTime, kLogEpochIdRemapTime, kLogEpochShuffleTime, kLogEpochCopyTime,
 kLogEpochConvertTime, kLogEpochTrainTime, kLogEpochTotalTime, kLogEpochFeatureBytes, kLogEpochMissBytes,
 kLogEpochLocalCacheBytes, kLogEpochNumSample) = range(16)


class SamGraphBasics(object):
    """ctypes facade, samgraph/common/__init__.py:278-574"""

    def __init__(self):
        c, L = ctypes, _lib
        self.C_LIB_CTYPES = L
        for name in ("samgraph_num_epoch", "samgraph_steps_per_epoch", "samgraph_num_class", "samgraph_feat_dim",
                     "samgraph_get_graph_num_src", "samgraph_get_graph_num_dst", "samgraph_get_graph_num_edge",
                     "samgraph_num_local_step"):
            getattr(L, name).restype = c.c_size_t
        L.samgraph_get_next_batch.restype = c.c_uint64
        L.samgraph_wait_one_child.restype = c.c_int
        for name in ("samgraph_get_graph_num_src", "samgraph_get_graph_num_dst", "samgraph_get_graph_num_edge"):
            getattr(L, name).argtypes = (c.c_uint64, c.c_int)
        L.samgraph_log_step.argtypes = (c.c_uint64, c.c_uint64, c.c_int, c.c_double)
        L.samgraph_log_step_by_key.argtypes = (c.c_uint64, c.c_int, c.c_double)
        L.samgraph_log_step_add.argtypes = (c.c_uint64, c.c_uint64, c.c_int, c.c_double)
        L.samgraph_log_epoch_add.argtypes = (c.c_uint64, c.c_int, c.c_double)
        L.samgraph_get_log_init_value.argtypes = (c.c_int,)
        L.samgraph_get_log_init_value.restype = c.c_double
        L.samgraph_get_log_step_value.argtypes = (c.c_uint64, c.c_uint64, c.c_int)
        L.samgraph_get_log_step_value.restype = c.c_double
        L.samgraph_get_log_step_value_by_key.argtypes = (c.c_uint64, c.c_int)
        L.samgraph_get_log_step_value_by_key.restype = c.c_double
        L.samgraph_get_log_epoch_value.argtypes = (c.c_uint64, c.c_int)
        L.samgraph_get_log_epoch_value.restype = c.c_double
        for name in ("samgraph_report_step", "samgraph_report_step_average"):
            getattr(L, name).argtypes = (c.c_uint64, c.c_uint64)
        for name in ("samgraph_report_epoch", "samgraph_report_epoch_average"):
            getattr(L, name).argtypes = (c.c_uint64,)
        for name in ("samgraph_trace_step_begin", "samgraph_trace_step_end"):
            getattr(L, name).argtypes = (c.c_uint64, c.c_int, c.c_uint64)
        for name in ("samgraph_trace_step_begin_now", "samgraph_trace_step_end_now"):
            getattr(L, name).argtypes = (c.c_uint64, c.c_int)

    def config(self, run_config: dict):
        keys = [str.encode(str(k)) for k in run_config.keys()]
        vals = [str.encode(' '.join(str(x) for x in v)) if isinstance(v, list) else str.encode(str(v))
                for v in run_config.values()]
        n = len(keys)
        return self.C_LIB_CTYPES.samgraph_config((ctypes.c_char_p * n)(*keys), (ctypes.c_char_p * n)(*vals),
                                                 ctypes.c_size_t(n))

    def init(self):
        return self.C_LIB_CTYPES.samgraph_init()

    def data_init(self):
        return self.C_LIB_CTYPES.samgraph_data_init()

    def sample_init(self, worker_id, ctx):
        return self.C_LIB_CTYPES.samgraph_sample_init(ctypes.c_int(worker_id), ctypes.c_char_p(str.encode(ctx)))

    def um_sample_init(self, num_workers):
        return self.C_LIB_CTYPES.samgraph_um_sample_init(ctypes.c_int(num_workers))

    def train_init(self, worker_id, ctx):
        return self.C_LIB_CTYPES.samgraph_train_init(ctypes.c_int(worker_id), ctypes.c_char_p(str.encode(ctx)))

    def switch_init(self, worker_id, ctx, cache_percentage):
        return self.C_LIB_CTYPES.samgraph_switch_init(ctypes.c_int(worker_id), ctypes.c_char_p(str.encode(ctx)),
                                                      ctypes.c_double(cache_percentage))

    def extract_start(self, count):
        return self.C_LIB_CTYPES.samgraph_extract_start(ctypes.c_int(count))

    def num_local_step(self):
        return self.C_LIB_CTYPES.samgraph_num_local_step()

    def start(self):
        return self.C_LIB_CTYPES.samgraph_start()

    def shutdown(self):
        return self.C_LIB_CTYPES.samgraph_shutdown()

    def num_class(self):
        return self.C_LIB_CTYPES.samgraph_num_class()

    def feat_dim(self):
        return self.C_LIB_CTYPES.samgraph_feat_dim()

    def num_epoch(self):
        return self.C_LIB_CTYPES.samgraph_num_epoch()

    def steps_per_epoch(self):
        return self.C_LIB_CTYPES.samgraph_steps_per_epoch()

    def get_next_batch(self):
        return self.C_LIB_CTYPES.samgraph_get_next_batch()

    def get_graph_num_src(self, key, graph_id):
        return self.C_LIB_CTYPES.samgraph_get_graph_num_src(key, graph_id)

    def get_graph_num_dst(self, key, graph_id):
        return self.C_LIB_CTYPES.samgraph_get_graph_num_dst(key, graph_id)

    def get_graph_num_edge(self, key, graph_id):
        return self.C_LIB_CTYPES.samgraph_get_graph_num_edge(key, graph_id)

    def sample_once(self):
        return self.C_LIB_CTYPES.samgraph_sample_once()

    def log_step(self, epoch, step, item, val):
        return self.C_LIB_CTYPES.samgraph_log_step(epoch, step, item, val)

    def log_step_by_key(self, key, item, val):
        return self.C_LIB_CTYPES.samgraph_log_step_by_key(key, item, val)

    def log_step_add(self, epoch, step, item, val):
        return self.C_LIB_CTYPES.samgraph_log_step_add(epoch, step, item, val)

    def log_epoch_add(self, epoch, item, val):
        return self.C_LIB_CTYPES.samgraph_log_epoch_add(epoch, item, val)

    def get_log_init_value(self, item):
        return self.C_LIB_CTYPES.samgraph_get_log_init_value(item)

    def get_log_step_value(self, epoch, step, item):
        return self.C_LIB_CTYPES.samgraph_get_log_step_value(epoch, step, item)

    def get_log_step_value_by_key(self, key, item):
        return self.C_LIB_CTYPES.samgraph_get_log_step_value_by_key(key, item)

    def get_log_epoch_value(self, epoch, item):
        return self.C_LIB_CTYPES.samgraph_get_log_epoch_value(epoch, item)

    def report_init(self):
        return self.C_LIB_CTYPES.samgraph_report_init()

    def report_step(self, epoch, step):
        return self.C_LIB_CTYPES.samgraph_report_step(epoch, step)

    def report_step_average(self, epoch, step):
        return self.C_LIB_CTYPES.samgraph_report_step_average(epoch, step)

    def report_epoch(self, epoch):
        return self.C_LIB_CTYPES.samgraph_report_epoch(epoch)

    def report_epoch_average(self, epoch):
        return self.C_LIB_CTYPES.samgraph_report_epoch_average(epoch)

    def report_node_access(self):
        return self.C_LIB_CTYPES.samgraph_report_node_access()

    def trace_step_begin(self, key, item, ts):
        return self.C_LIB_CTYPES.samgraph_trace_step_begin(key, item, ts)

    def trace_step_end(self, key, item, ts):
        return self.C_LIB_CTYPES.samgraph_trace_step_end(key, item, ts)

    def trace_step_begin_now(self, key, item):
        return self.C_LIB_CTYPES.samgraph_trace_step_begin_now(key, item)

    def trace_step_end_now(self, key, item):
        return self.C_LIB_CTYPES.samgraph_trace_step_end_now(key, item)

    def dump_trace(self):
        return self.C_LIB_CTYPES.samgraph_dump_trace()

    def forward_barrier(self):
        return self.C_LIB_CTYPES.samgraph_forward_barrier()

    def wait_one_child(self):
        return self.C_LIB_CTYPES.samgraph_wait_one_child()


_basics = SamGraphBasics()
