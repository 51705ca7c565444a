"""ctypes loader for libxgnn_b200.so (the C-ABI declared in include/xgnn_b200.h and
include/samgraph_abi.h).  Mirrors how the reference loads its own C library:
samgraph/common/__init__.py:280-362 (ctypes.CDLL(path, RTLD_GLOBAL) + explicit arg/restypes).

There is NO fallback: if the CUDA library is missing this module raises at import time.
"""
from __future__ import annotations

import ctypes
from ctypes import (POINTER, Structure, c_char_p, c_double, c_int, c_size_t, c_uint32, c_uint64,
                    c_void_p)
from pathlib import Path

XG_MAX_PARTS = 16
XG_EMPTY_KEY = 0xFFFFFFFF

LIB_PATH = Path(__file__).resolve().parent / "lib" / "libxgnn_b200.so"


class XgGraph(Structure):
    _fields_ = [
        ("part_indptr", c_void_p * (XG_MAX_PARTS + 1)),
        ("part_indices", c_void_p * (XG_MAX_PARTS + 1)),
        ("num_part", c_uint32),
        ("num_cache_node", c_uint32),
        ("num_node", c_uint32),
        ("reserved", c_uint32),
    ]


class XgFeatureStore(Structure):
    _fields_ = [
        ("part", c_void_p * XG_MAX_PARTS),
        ("host", c_void_p),
        ("cache_table", c_void_p),
        ("num_part", c_uint32),
        ("local_part_mask", c_uint32),
        ("row_bytes", c_uint64),
        ("replica", c_void_p),
        ("num_replicated", c_uint32),
        ("compact_bitwidth", c_uint32),
        ("replica_words", c_void_p),
        ("replica_words_cover_all", c_uint32),
        ("host_row_mask", c_uint32),
    ]


XG_MAX_LAYERS = 8


class XgSamplerConfig(Structure):
    _fields_ = [
        ("num_layer", c_uint32),
        ("fanout", c_uint32 * XG_MAX_LAYERS),
        ("batch_size", c_uint32),
        ("sample_type", c_int),
        ("num_slots", c_uint32),
        ("flags", c_uint32),
        ("feat_dim", c_uint32),
        ("feat_dtype", c_int),
        ("seed", c_uint64),
        ("max_input_nodes", c_uint64),
        ("stream", c_void_p),
        ("random_walk_length", c_uint32),
        ("num_random_walk", c_uint32),
        ("random_walk_restart_prob", c_double),
        ("prob_table", c_void_p),
        ("alias_table", c_void_p),
        ("prob_prefix_table", c_void_p),
    ]


class XgBatchView(Structure):
    _fields_ = [
        ("key", c_uint64),
        ("num_layer", c_uint32),
        ("num_output", c_uint32),
        ("num_input", c_uint32),
        ("reserved", c_uint32),
        ("row", c_void_p * XG_MAX_LAYERS),
        ("col", c_void_p * XG_MAX_LAYERS),
        ("data", c_void_p * XG_MAX_LAYERS),
        ("num_src", c_uint64 * XG_MAX_LAYERS),
        ("num_dst", c_uint64 * XG_MAX_LAYERS),
        ("num_edge", c_uint64 * XG_MAX_LAYERS),
        ("input_nodes", c_void_p),
        ("output_nodes", c_void_p),
        ("input_feat", c_void_p),
        ("output_label", c_void_p),
        ("csc_indptr", c_void_p * XG_MAX_LAYERS),
    ]


def _load() -> ctypes.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m xgnn_b200.build` "
            "(xgnn_b200 has no CPU fallback)")
    return ctypes.CDLL(str(LIB_PATH), mode=ctypes.RTLD_GLOBAL)


lib = _load()

_u32p = c_void_p  # device pointers are passed as integers


def _sig(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


# ---- core -----------------------------------------------------------------------------------
_sig("xg_last_error", c_char_p)
_sig("xg_version", c_int)
_sig("xg_launch_count", c_uint64)

# ---- hash table -----------------------------------------------------------------------------
_sig("xg_hashtable_create", c_int, POINTER(c_void_p), c_size_t, c_int)
_sig("xg_hashtable_destroy", c_int, c_void_p)
_sig("xg_hashtable_reset", c_int, c_void_p, c_void_p)
_sig("xg_hashtable_fill_with_unique", c_int, c_void_p, _u32p, c_size_t, c_void_p)
_sig("xg_hashtable_fill_with_dup_revised", c_int, c_void_p, _u32p, c_size_t, c_void_p)
_sig("xg_hashtable_fill_with_dup_mutable", c_int, c_void_p, _u32p, c_size_t, c_void_p)
_sig("xg_hashtable_fill_with_duplicates", c_int, c_void_p, _u32p, c_size_t, _u32p, POINTER(c_uint32), c_void_p)
_sig("xg_hashtable_ref_unique", c_int, c_void_p, POINTER(c_void_p), POINTER(c_uint32))
_sig("xg_hashtable_copy_unique", c_int, c_void_p, _u32p, c_void_p)
_sig("xg_hashtable_num_items", c_size_t, c_void_p)
_sig("xg_hashtable_lookup", c_int, c_void_p, _u32p, c_size_t, _u32p, c_void_p)
_sig("xg_map_edges", c_int, c_void_p, _u32p, _u32p, _u32p, _u32p, c_size_t, c_void_p)

# ---- samplers -------------------------------------------------------------------------------
_sig("xg_sample_khop", c_int, POINTER(XgGraph), c_int, _u32p, c_size_t, c_size_t, _u32p, _u32p, c_void_p,
     c_uint64, c_uint64, c_void_p)

_sig("xg_sample_khop1", c_int, POINTER(XgGraph), _u32p, c_size_t, c_size_t, _u32p, _u32p, c_void_p, c_uint64, c_uint64,
     c_void_p)
_sig("xg_sample_weighted_khop", c_int, POINTER(XgGraph), c_int, c_void_p, c_void_p, c_void_p, _u32p, c_size_t, c_size_t,
     _u32p, _u32p, c_void_p, c_uint64, c_uint64, c_void_p)
_sig("xg_sample_random_walk", c_int, POINTER(XgGraph), _u32p, c_size_t, c_size_t, c_double, c_size_t, c_size_t, _u32p,
     _u32p, _u32p, c_void_p, c_uint64, c_uint64, c_void_p)
_sig("xg_philox4x32", None, c_uint64, c_uint64, c_uint32, c_uint32, c_uint32, c_uint32, POINTER(c_uint32))
_sig("xg_sample_salt", c_uint32)

# ---- extraction -----------------------------------------------------------------------------
_sig("xg_extract", c_int, c_void_p, c_void_p, _u32p, c_size_t, c_size_t, c_int, c_void_p)
_sig("xg_extract_ggms", c_int, POINTER(XgFeatureStore), _u32p, c_size_t, c_void_p, c_void_p, c_void_p)
_sig("xg_get_miss_cache_index", c_int, _u32p, _u32p, c_size_t, _u32p, _u32p, POINTER(c_size_t), _u32p, _u32p,
     POINTER(c_size_t), c_void_p)
_sig("xg_get_miss_cache_index_noalloc", c_int, _u32p, _u32p, c_size_t, _u32p, _u32p, _u32p, _u32p, c_void_p, c_void_p)
_sig("xg_mock_extract", c_int, c_void_p, c_void_p, _u32p, c_size_t, c_size_t, c_uint32, c_void_p)
_sig("xg_extract_miss_data", c_int, c_void_p, c_void_p, _u32p, _u32p, c_size_t, c_size_t, c_void_p)
_sig("xg_combine_cache_data", c_int, c_void_p, POINTER(XgFeatureStore), _u32p, _u32p, c_size_t, c_void_p)


# ---- minibatch engine -----------------------------------------------------------------------
_sig("xg_sampler_create", c_int, POINTER(c_void_p), POINTER(XgGraph), POINTER(XgFeatureStore), c_void_p,
     POINTER(XgSamplerConfig), c_int)
_sig("xg_sampler_destroy", c_int, c_void_p)
_sig("xg_sampler_submit", c_int, c_void_p, c_void_p, c_size_t, c_int, c_uint64, POINTER(c_void_p))
_sig("xg_batch_wait", c_int, c_void_p, POINTER(XgBatchView))
_sig("xg_batch_release", c_int, c_void_p)
_sig("xg_build_replica_words", c_int, c_void_p, c_uint32, c_uint32, c_void_p, c_void_p, c_void_p)
_sig("xg_ics22_placement", c_int, c_void_p, c_void_p, c_uint32, c_uint32, c_uint32, c_uint32, c_double, c_uint32, c_uint32,
     c_void_p, c_void_p, c_void_p)
_sig("xg_sanity_check_ids", c_int, c_void_p, c_size_t, c_uint32, c_void_p, c_void_p, c_void_p)
_sig("xg_sanity_check_edges", c_int, c_void_p, c_void_p, c_size_t, c_uint32, c_uint32, c_void_p, c_void_p)
_sig("xg_presample_accumulate", c_int, c_void_p, c_uint32, c_void_p, c_size_t, c_void_p, c_void_p)
_sig("xg_presample_rank", c_int, c_void_p, c_uint32, c_void_p, c_void_p, c_void_p)
_sig("xg_sampler_stream", c_void_p, c_void_p)
_sig("xg_sampler_extract_stream", c_void_p, c_void_p)
_sig("xg_batch_timeline", c_int, c_void_p, POINTER(ctypes.c_float))
_sig("xg_sampler_launches_per_batch", c_uint32, c_void_p)
_sig("xg_batch_stage_ms", c_int, c_void_p, c_int, c_int, POINTER(ctypes.c_float))
_sig("xg_synth_degrees", c_int, c_void_p, c_uint64, c_uint64, c_size_t, c_uint64, c_double, c_uint32, c_void_p)
_sig("xg_synth_indices", c_int, c_void_p, c_void_p, c_uint64, c_uint64, c_size_t, c_uint64, c_uint64, c_void_p)
_sig("xg_synth_features", c_int, c_void_p, c_void_p, c_uint64, c_uint64, c_size_t, c_uint32, c_int, c_void_p)
_sig("xg_shuffle_cache_ranks", c_int, c_void_p, c_size_t)
_sig("xg_num_cache_node", c_uint32, c_void_p, c_uint32, c_double)
_sig("xg_part_num_rows", c_size_t, c_uint32, c_uint32, c_uint32)
_sig("xg_part_degrees", c_int, c_void_p, c_uint32, c_uint32, c_uint32, c_void_p, c_void_p)
_sig("xg_part_copy_indices", c_int, c_void_p, c_void_p, c_uint32, c_uint32, c_uint32, c_void_p, c_void_p, c_void_p)
_sig("xg_device_malloc", c_int, POINTER(c_void_p), c_size_t)
_sig("xg_device_free", c_int, c_void_p)
_sig("xg_ipc_export", c_int, c_void_p, c_void_p)
_sig("xg_ipc_import", c_int, c_void_p, POINTER(c_void_p))
_sig("xg_ipc_close", c_int, c_void_p)
_sig("xg_enable_peer_access", c_int, c_int, c_int)
_sig("xg_aligned_shuffle_epoch", c_int, c_void_p, c_size_t, c_uint64)
_sig("xg_partition_solve", c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p)
_sig("xg_detect_topology", c_int, c_int, c_void_p)
_sig("xg_predict_num_nodes", c_size_t, c_size_t, POINTER(c_uint32), c_size_t)


class XgError(RuntimeError):
    pass


def check(rc: int) -> None:
    if rc != 0:
        raise XgError(f"xgnn_b200 error {rc}: {lib.xg_last_error().decode()}")


def declared_symbols() -> list[str]:
    """Every extern "C" symbol the headers under include/ declare (parsed from the headers)."""
    import re
    names: list[str] = []
    inc = Path(__file__).resolve().parent.parent / "include"
    for h in sorted(inc.glob("*.h")):
        text = re.sub(r"/\*.*?\*/", "", h.read_text(), flags=re.S)
        text = re.sub(r"//[^\n]*", "", text)
        for m in re.finditer(r"\b((?:xg|samgraph)_[A-Za-z0-9_]+)\s*\(", text):
            if m.group(1) not in names:
                names.append(m.group(1))
    return names
