"""Generates the committed golden fixtures from the REFERENCE's own CPU code (oracle/_ref, built in
place from /root/reference by oracle/Makefile).  Run only where /root/reference exists:

    python tests/golden/make_golden.py

Outputs (small, committed):
  tests/golden/ref_path_small.npz   graph + inputs + outputs of CPUSampleKHop0 (fanout >= max degree, the
                                    deterministic branch), CPUHashTable0 Populate/MapNodes/MapEdges,
                                    CPUExtract for f32/f16/i64, and a 3-layer arch0 minibatch.
  tests/golden/solver_golden.json   utility/nvlink_placement_solver inputs (bandwidth matrices) and the
                                    placements solver.cc prints for them.
"""
import ctypes
import json
import re
import subprocess
import sys
from ctypes import c_size_t, c_void_p, c_int, c_double, POINTER
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))
import oracle_lib as O  # noqa: E402
import synth  # noqa: E402

assert O.ref is not None, "build oracle/_ref first (make -C oracle ref)"
ref = O.ref

V, E = 3000, 24000
indptr, indices = synth.make_graph(V, E, seed=20240611, max_deg=60)
maxdeg = int(np.diff(indptr.astype(np.int64)).max())
rng = np.random.default_rng(7)
inp = rng.choice(V, size=500, replace=False).astype(np.uint32)
FAN = maxdeg  # >= every degree: bit-exact branch
src, dst = O.ref_sample_khop0(indptr, indices, inp, FAN)

# hashtable0: two successive fills with duplicates
t = O.RefTable(0, V + 16)
fill1 = rng.integers(0, V, size=1500).astype(np.uint32)
fill2 = rng.integers(0, V, size=2500).astype(np.uint32)
t.populate(fill1)
uniq1 = t.unique().copy()
t.populate(fill2)
uniq2 = t.unique().copy()
ms, md = t.map_edges(fill2, fill2[::-1].copy())

# extract
f32 = synth.make_features(V, 100, np.float32)
f16 = synth.make_features(V, 24, np.float16)
lab = synth.make_labels(V)
idx = rng.integers(0, V, size=700).astype(np.uint32)
x32 = O.ref_extract(f32, idx)
x16 = O.ref_extract(f16, idx)
xl = O.ref_extract(lab, idx)

# 3-layer minibatch through the arch0 call order with CPUHashTable0, fanouts >= max degree
O._sig(ref, "ref_batch_create", c_void_p)
O._sig(ref, "ref_batch_destroy", None, c_void_p)
O._sig(ref, "ref_batch_row", c_void_p, c_void_p, c_int)
O._sig(ref, "ref_batch_col", c_void_p, c_void_p, c_int)
O._sig(ref, "ref_batch_counts", c_void_p, c_void_p)
O._sig(ref, "ref_batch_input_nodes", c_void_p, c_void_p)
O._sig(ref, "ref_batch_feat", c_void_p, c_void_p)
O._sig(ref, "ref_batch_label", c_void_p, c_void_p)
O._sig(ref, "ref_arch0_batch", c_size_t, c_void_p, c_int, O.u32p, O.u32p, O.u32p, c_size_t, POINTER(c_size_t), c_int,
       c_void_p, c_size_t, c_int, c_size_t, c_void_p, c_void_p, POINTER(c_size_t), POINTER(c_double), POINTER(c_double))

# a sparser graph for the multi-layer case so the frontier stays small
indptr_b, indices_b = synth.make_graph(4000, 9000, seed=99, max_deg=12)
maxdeg_b = int(np.diff(indptr_b.astype(np.int64)).max())
seeds = rng.choice(4000, size=64, replace=False).astype(np.uint32)
fan = (c_size_t * 3)(maxdeg_b, maxdeg_b, maxdeg_b)
featb = synth.make_features(4000, 16, np.float32)
labb = synth.make_labels(4000)
tb = ref.ref_hashtable_create(0, 4000 + 16)
bo = ref.ref_batch_create()
nin = c_size_t()
total = ref.ref_arch0_batch(tb, 0, indptr_b, indices_b.copy(), seeds, len(seeds), fan, 3, featb.ctypes.data, 16,
                            ref.ref_dtype_f32(), 4, labb.ctypes.data, bo, ctypes.byref(nin), None, None)
counts = np.ctypeslib.as_array(ctypes.cast(ref.ref_batch_counts(bo), POINTER(c_size_t)), shape=(9,)).copy()
batch = {}
for l in range(3):
    ne = int(counts[3 * l + 2])
    batch[f"b_row{l}"] = np.ctypeslib.as_array(ctypes.cast(ref.ref_batch_row(bo, l), POINTER(ctypes.c_uint32)), shape=(ne,)).copy()
    batch[f"b_col{l}"] = np.ctypeslib.as_array(ctypes.cast(ref.ref_batch_col(bo, l), POINTER(ctypes.c_uint32)), shape=(ne,)).copy()
batch["b_input_nodes"] = np.ctypeslib.as_array(ctypes.cast(ref.ref_batch_input_nodes(bo), POINTER(ctypes.c_uint32)), shape=(nin.value,)).copy()
batch["b_feat"] = np.ctypeslib.as_array(ctypes.cast(ref.ref_batch_feat(bo), POINTER(ctypes.c_float)), shape=(nin.value, 16)).copy()
batch["b_label"] = np.ctypeslib.as_array(ctypes.cast(ref.ref_batch_label(bo), POINTER(ctypes.c_int64)), shape=(len(seeds),)).copy()
ref.ref_batch_destroy(bo)
ref.ref_hashtable_destroy(tb)

np.savez_compressed(
    HERE / "ref_path_small.npz",
    indptr=indptr, indices=indices, inp=inp, fanout=np.int64(FAN), khop0_src=src, khop0_dst=dst,
    fill1=fill1, fill2=fill2, uniq1=uniq1, uniq2=uniq2, map_src=ms, map_dst=md,
    idx=idx, x32=x32, x16=x16, xl=xl,
    b_indptr=indptr_b, b_indices=indices_b, b_seeds=seeds, b_fanout=np.int64(maxdeg_b), b_counts=counts,
    b_total=np.int64(total), **batch)

# ---- placement solver goldens ---------------------------------------------------------------
sol = {}
for n in (5, 6, 7, 8):
    inp_file = Path(f"/root/reference/utility/nvlink_placement_solver/input_{n}GPU.txt")
    out = subprocess.run([str(HERE.parent.parent / "oracle/_ref/solver.out"), str(inp_file), "2"],
                         capture_output=True, text=True, check=True).stdout
    lines = inp_file.read_text().split("\n")
    ng, nl = map(int, lines[0].split())
    links = [[int(a), int(b), float(c)] for a, b, c in (ln.split() for ln in lines[1:1 + nl])]
    placement = []
    for m in re.finditer(r"gpu (\d+): \[([\d ]+)\], \[([\d ]+)\]", out):
        placement.append({"gpu": int(m.group(1)), "part_ids": list(map(int, m.group(2).split())),
                          "access": list(map(int, m.group(3).split()))})
    assert len(placement) == ng
    sol[str(n)] = {"num_gpu": ng, "links": links, "placement": placement}
(HERE / "solver_golden.json").write_text(json.dumps(sol, indent=1))
print("golden fixtures written:", (HERE / "ref_path_small.npz").stat().st_size, "bytes")
