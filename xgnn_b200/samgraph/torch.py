"""Mirror of samgraph/torch/adapter.py (+ the pybind11 module samgraph/torch/adapter.cc it calls).

`import xgnn_b200.samgraph.torch as sam` gives the same module-level functions the training scripts use:
config / init / data_init / sample_init / train_init / sample_once / extract_start / get_next_batch /
get_dgl_blocks / get_graph_feat / ... (adapter.py:28-185).  Batch tensors are zero-copy views of the batch
slot (torch.as_tensor over __cuda_array_interface__); each view holds a retained handle and releases it when
it dies -- the ownership rule of the reference's from_blob deleter (adapter.cc:70-76)."""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .._clib import lib as _lib
from .common import *  # noqa: F401,F403
from .common import _basics

(TENSOR_FEAT, TENSOR_LABEL, TENSOR_ROW, TENSOR_COL, TENSOR_DATA, TENSOR_INPUT_NODES, TENSOR_OUTPUT_NODES,
 TENSOR_DATASET_FEAT, TENSOR_DATASET_LABEL, TENSOR_CSC_INDPTR) = range(10)


class _TensorView(ctypes.Structure):  # samgraph_tensor_view, include/samgraph_abi.h
    _fields_ = [("data", ctypes.c_void_p), ("shape", ctypes.c_int64 * 2), ("ndim", ctypes.c_int32),
                ("dtype", ctypes.c_int32), ("device", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("owner", ctypes.c_void_p)]


_lib.samgraph_get_graph_tensor.argtypes = (ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.POINTER(_TensorView))
_lib.samgraph_get_graph_tensor.restype = None
_lib.samgraph_tensor_release.argtypes = (ctypes.c_void_p,)
_lib.samgraph_tensor_release.restype = None

# DataType numbering of common.h:38-47 -> (torch dtype, array-interface typestr)
_DTYPES = {0: (torch.float32, "<f4"), 1: (torch.float64, "<f8"), 2: (torch.float16, "<f2"), 3: (torch.uint8, "|u1"),
           4: (torch.int32, "<i4"), 5: (torch.int8, "|i1"), 6: (torch.int64, "<i8")}


class _Owned:
    """array-interface carrier; releases the retained batch handle when the last tensor view dies"""

    def __init__(self, view: _TensorView, typestr: str, shape, cuda: bool):
        iface = {"shape": tuple(shape), "typestr": typestr, "data": (view.data, False), "version": 3, "strides": None}
        if cuda:
            self.__cuda_array_interface__ = iface
        else:
            self.__array_interface__ = iface
        self._owner = view.owner

    def __del__(self):
        if self._owner:
            _lib.samgraph_tensor_release(self._owner)
            self._owner = None


def _tensor(key, which, layer=0):
    v = _TensorView()
    _lib.samgraph_get_graph_tensor(key, which, layer, ctypes.byref(v))
    dtype, typestr = _DTYPES[v.dtype]
    shape = [int(v.shape[i]) for i in range(v.ndim)]
    n = int(np.prod(shape)) if shape else 0
    if v.device < 0:
        if n == 0:
            return torch.empty(shape, dtype=dtype)
        return torch.as_tensor(np.asarray(_Owned(v, typestr, shape, False)))
    dev = torch.device("cuda", v.device)
    if n == 0 or not v.data:
        if v.owner:
            _lib.samgraph_tensor_release(v.owner)
        return torch.empty(shape, dtype=dtype, device=dev)
    return torch.as_tensor(_Owned(v, typestr, shape, True), device=dev)


# ---- adapter.py:28-73 ------------------------------------------------------------------------------------
config = _basics.config
init = _basics.init
start = _basics.start
num_class = _basics.num_class
feat_dim = _basics.feat_dim
num_epoch = _basics.num_epoch
steps_per_epoch = _basics.steps_per_epoch
get_next_batch = _basics.get_next_batch
get_graph_num_src = _basics.get_graph_num_src
get_graph_num_dst = _basics.get_graph_num_dst
get_graph_num_edge = _basics.get_graph_num_edge
shutdown = _basics.shutdown
sample_once = _basics.sample_once
log_step = _basics.log_step
log_step_add = _basics.log_step_add
log_epoch_add = _basics.log_epoch_add
get_log_init_value = _basics.get_log_init_value
get_log_step_value = _basics.get_log_step_value
get_log_epoch_value = _basics.get_log_epoch_value
report_init = _basics.report_init
report_step = _basics.report_step
report_step_average = _basics.report_step_average
report_epoch = _basics.report_epoch
report_epoch_average = _basics.report_epoch_average
report_node_access = _basics.report_node_access
trace_step_begin = _basics.trace_step_begin
trace_step_end = _basics.trace_step_end
trace_step_begin_now = _basics.trace_step_begin_now
trace_step_end_now = _basics.trace_step_end_now
dump_trace = _basics.dump_trace
forward_barrier = _basics.forward_barrier
wait_one_child = _basics.wait_one_child
log_step_by_key = _basics.log_step_by_key
get_log_step_value_by_key = _basics.get_log_step_value_by_key
switch_init = _basics.switch_init
data_init = _basics.data_init
sample_init = _basics.sample_init
um_sample_init = _basics.um_sample_init
train_init = _basics.train_init
extract_start = _basics.extract_start
num_local_step = _basics.num_local_step


# ---- adapter.py:75-185 -----------------------------------------------------------------------------------
def get_graph_feat(batch_key):
    batch_feat = _tensor(batch_key, TENSOR_FEAT)
    if batch_feat.dtype != torch.float32:  # fp16 datasets are upcast on the Python side (adapter.py:77-78)
        batch_feat = batch_feat.float()
    return batch_feat


def get_graph_label(batch_key):
    return _tensor(batch_key, TENSOR_LABEL)


def get_graph_row(batch_key, layer_idx):
    return _tensor(batch_key, TENSOR_ROW, layer_idx)


def get_graph_col(batch_key, layer_idx):
    return _tensor(batch_key, TENSOR_COL, layer_idx)


def get_graph_data(batch_key, layer_idx):
    return _tensor(batch_key, TENSOR_DATA, layer_idx)


def _create_dgl_block(data, num_src_nodes, num_dst_nodes):
    """adapter.py:98-103.  DGL is not part of this image; where it is installed the block is built exactly as
    the reference does, otherwise the COO triple is returned so that callers without DGL can still consume it."""
    row, col = data
    try:
        import dgl
        from dgl.heterograph import DGLBlock
    except ImportError:
        return (row, col, num_src_nodes, num_dst_nodes)
    gidx = dgl.heterograph_index.create_unitgraph_from_coo(2, num_src_nodes, num_dst_nodes, row, col,
                                                           ['coo', 'csr', 'csc'])
    return DGLBlock(gidx, (['_N'], ['_N']), ['_E'])


def get_dgl_blocks(batch_key, num_layers, with_feat=True):
    feat = get_graph_feat(batch_key) if with_feat else None
    label = get_graph_label(batch_key) if with_feat else None
    blocks = []
    for i in range(num_layers):
        row, col = get_graph_row(batch_key, i), get_graph_col(batch_key, i)
        blocks.append(_create_dgl_block((row, col), get_graph_num_src(batch_key, i), get_graph_num_dst(batch_key, i)))
    return blocks, feat, label


def get_graph_csc_indptr(batch_key, layer_idx):
    """[num_dst + 1] int32 edge offsets of the layer's destination nodes (KHop0/KHop2/KHop3): with
    get_graph_row() as the indices array the block is already in CSC form."""
    t = _tensor(batch_key, TENSOR_CSC_INDPTR, layer_idx)
    return t if t is not None and t.numel() > 0 else None


def _create_dgl_block_csc(indptr, indices, num_src_nodes, num_dst_nodes):
    """hand DGL the destination-major form directly (dgl.create_block with a 'csc' tuple) so
    that message passing needs no COO -> CSC conversion (kLogL1ConvertTime of the reference's loops).
    Without DGL the CSC triple is returned."""
    try:
        import dgl
    except ImportError:
        return (indptr, indices, num_src_nodes, num_dst_nodes)
    return dgl.create_block(('csc', (indptr, indices, torch.empty(0, dtype=indices.dtype, device=indices.device))),
                            num_src_nodes=num_src_nodes, num_dst_nodes=num_dst_nodes)


def get_dgl_blocks_csc(batch_key, num_layers, with_feat=True):
    """get_dgl_blocks (adapter.py:105-124) with every block built from the sampler's own CSC offsets."""
    feat = get_graph_feat(batch_key) if with_feat else None
    label = get_graph_label(batch_key) if with_feat else None
    blocks = []
    for i in range(num_layers):
        indptr = get_graph_csc_indptr(batch_key, i)
        nsrc, ndst = get_graph_num_src(batch_key, i), get_graph_num_dst(batch_key, i)
        if indptr is None:  # no CSC form for this layer (see include/samgraph_abi.h): the reference's COO block
            blocks.append(_create_dgl_block((get_graph_row(batch_key, i), get_graph_col(batch_key, i)), nsrc, ndst))
        else:
            blocks.append(_create_dgl_block_csc(indptr, get_graph_row(batch_key, i), nsrc, ndst))
    return blocks, feat, label


def get_dgl_blocks_with_weights(batch_key, num_layers, with_feat=True):
    feat = get_graph_feat(batch_key) if with_feat else None
    label = get_graph_label(batch_key) if with_feat else None
    blocks = []
    for i in range(num_layers):
        row, col = get_graph_row(batch_key, i), get_graph_col(batch_key, i)
        weights = get_graph_data(batch_key, i)
        block = _create_dgl_block((row, col), get_graph_num_src(batch_key, i), get_graph_num_dst(batch_key, i))
        if isinstance(block, tuple):
            block = block + (weights,)
        else:
            block.edata['weights'] = weights
        blocks.append(block)
    return blocks, feat, label


def notify_sampler_ready(barrier):
    barrier.wait()


def wait_for_sampler_ready(barrier):
    barrier.wait()


def get_dataset_feat():
    return _tensor(0, TENSOR_DATASET_FEAT)


def get_dataset_label():
    return _tensor(0, TENSOR_DATASET_LABEL)


def get_graph_input_nodes(batch_key):
    return _tensor(batch_key, TENSOR_INPUT_NODES)


def get_graph_output_nodes(batch_key):
    return _tensor(batch_key, TENSOR_OUTPUT_NODES)


def load_subtensor(batch_key, feat, label, device):
    input_nodes = get_graph_input_nodes(batch_key).to(feat.device)
    output_nodes = get_graph_output_nodes(batch_key).to(label.device)
    batch_inputs = torch.index_select(feat, 0, input_nodes.long()).to(device, dtype=torch.float32)
    batch_labels = torch.index_select(label, 0, output_nodes.long()).to(device)
    return batch_inputs, batch_labels
