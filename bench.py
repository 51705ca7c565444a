#!/usr/bin/env python
"""bench.py -- sampled edges/s of the per-minibatch data path (sample -> unique/remap -> extract).

  python bench.py [--gpus N --steps K --warmup W]            our arm (CUDA, through the xg_* C-ABI)
  python bench.py --impl reference [...]                     the reference's own CPU path (arch0)

A step = one minibatch of `--batch` (8000) seeds of a synthetic papers100M-shape graph, GraphSAGE
3-hop uniform fanout 15,10,5 from the seeds (reference notation "5 10 15", consumed last -> first),
features 128 x fp32 gathered for every input node, labels for the seeds.  One JSON line on stdout.
After the headline the same line gains (unless --no-secondary):
  epoch          one REAL epoch loop (first epoch discarded, 3 measured), seeds from the aligned shuffler
  configs[]      short runs of BASELINE.json's other configs (products vs CPU arch0, twitter, friendster random walk
                 and weighted KHop, mag240M hot/cold), each with its own roofline block and placement
  sharded        N > 1: the same workload with the reference's own routing (features rank % P, topology v % P)
  e2e_samgraph   N = 1: papers100M / 16 through sam.config -> init -> sample_once -> get_next_batch -> get_dgl_blocks
See DESIGN.md "Measurement" for what every key means and how the roofline numbers are derived.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FANOUT = [5, 10, 15]  # indexed by layer; the seed-side layer (last) uses 15
METRIC = "sampled_edges_per_s"
UNIT = "edges/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="xgnn", choices=["xgnn", "reference"])
    ap.add_argument("--shape", default="papers100M")
    ap.add_argument("--scale", type=int, default=int(os.environ.get("XG_BENCH_SCALE", "1")),
                    help="divide the shape's node/edge counts (smoke runs only; 1 = the named config)")
    ap.add_argument("--batch", type=int, default=8000)
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--stage-steps", type=int, default=20, help="steps of the per-stage CUDA-event pass")
    ap.add_argument("--sample-type", type=int, default=0,
                    help="SampleType numbering of common.h:50-59: 0 KHop0 (default), 1 KHop1, 3 random walk (W=4, len 3, "
                         "restart 0.5, K=5: sgnn/train_pinsage.py:138-142), 7 KHop3; secondary workloads, not the headline")
    ap.add_argument("--max-input-nodes", type=int, default=0,
                    help="cap on a minibatch's unique nodes (sizes the hash table and the batch buffers); 0 = worst case "
                         "PredictNumNodes (common.cc:488-497)")
    ap.add_argument("--in-flight", type=int, default=2,
                    help="minibatches queued ahead of the consumer (one being sampled + the extract phase of the one before)")
    ap.add_argument("--replicate", type=float, default=float(os.environ.get("XG_BENCH_REPLICATE", "0.75")),
                    help="N > 1 only: fraction of the nodes (hottest first, by degree rank) whose feature rows every GPU "
                         "keeps in its own HBM; the rest is sharded rank %% P (0 = the reference's pure sharding)")
    ap.add_argument("--replicate-topology", type=int, default=int(os.environ.get("XG_BENCH_REPLICATE_TOPOLOGY", "1")),
                    help="N > 1 only: 1 = every GPU keeps the whole CSR (papers100M: 6.9 GB), the sample phase needs no "
                         "NVLink; 0 = the reference's v %% P sharding (always used by the `sharded` block and by config 3)")
    ap.add_argument("--cache-pct", type=float, default=1.0,
                    help="fraction of the nodes (by degree rank) whose rows are cached in HBM; the rest is read from a "
                         "pinned-host mock table (SAMGRAPH_EMPTY_FEAT style, 2^20 rows) over PCIe")
    ap.add_argument("--no-secondary", action="store_true", help="headline only: no epoch / configs / sharded / e2e_samgraph")
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region (B200_PROFILING.md)
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line)

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class AlignedShuffler:
    """DistAlignedShuffler (dist/dist_shuffler_aligned.cc:37-121) through the engine's own routine
    (xg_aligned_shuffle_epoch = the Fisher-Yates pass with std::default_random_engine(epoch) that csrc/engine.cc runs):
    the train set padded to a multiple of the world size, re-shuffled IN PLACE epoch after epoch, this worker's
    contiguous slice cut into batches."""

    def __init__(self, train: np.ndarray, rank: int, world: int, batch: int):
        from xgnn_b200._clib import check, lib
        self._check, self._lib = check, lib
        pad = (-len(train)) % world
        self.data = np.ascontiguousarray(np.concatenate([train, train[:pad]]) if pad else train.copy(), dtype=np.uint32)
        self.rank, self.n_local, self.batch, self.epoch = rank, len(self.data) // world, batch, 0
        self.steps = -(-self.n_local // batch)

    def next_epoch(self) -> list[np.ndarray]:
        self._check(self._lib.xg_aligned_shuffle_epoch(self.data.ctypes.data, len(self.data), self.epoch))
        self.epoch += 1
        mine = self.data[self.rank * self.n_local:(self.rank + 1) * self.n_local].copy()
        return [mine[o:o + self.batch] for o in range(0, self.n_local, self.batch)]


def seed_batches(train, rank, world, batch, count):
    """`count` minibatches of seeds, walking through as many epochs as needed."""
    sh, out = AlignedShuffler(train, rank, world, batch), []
    while len(out) < count:
        out.extend(sh.next_epoch()[: count - len(out)])
    return out


# --------------------------------------------------------------------------------------------------
# the reference's CPU path (arch0: CPUSampleKHop0 + CPUHashTable2 + CPUExtract), timed on the host cores
def mem_available_bytes() -> int:
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 0


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def host_dataset(args, want_features=True):
    """The same synthetic graph in host memory (generated on the GPU when there is one, copied back)."""
    import torch
    from xgnn_b200 import dataset as D
    V, E, dim, dtype, ncls, frac = D.SHAPES[args.shape]
    V, E = V // args.scale, E // args.scale
    c = D.degree_constant(V, E)
    dev = torch.device("cuda", 0)
    indptr, indices, ne = D.gen_csr(0, 1, V, V, c, D.SYNTH_SEED, dev)
    h = {"V": V, "E": ne, "dim": dim, "dtype": dtype,
         "indptr": indptr.cpu().numpy().view(np.uint32), "indices": indices.cpu().numpy().view(np.uint32)}
    del indptr, indices
    row_bytes = dim * torch.empty((), dtype=dtype).element_size()
    h["feat"], h["feat_note"] = None, "features skipped (host memory)"
    if want_features and mem_available_bytes() > V * row_bytes * 1.25 + (8 << 30):
        feat = torch.empty((V, dim), dtype=dtype)
        step = max(1, (1 << 30) // row_bytes)
        for o in range(0, V, step):
            n = min(step, V - o)
            feat[o:o + n].copy_(D.gen_features(n, dim, dtype, dev, begin=o))
        h["feat"], h["feat_note"] = feat.numpy(), "full feature table in host memory"
    h["label"] = (np.arange(V, dtype=np.int64) % ncls)
    h["train"] = D.train_set(V, max(1, int(V * frac)))
    torch.cuda.empty_cache()
    return h


class CpuArm:
    """oracle/_ref/libxgnn_ref.so (kind 'reference': the reference's own .cc files compiled in place) or,
    when that is absent, the single-threaded C restatement oracle/liboracle.so (kind 'port')."""

    def __init__(self, h, threads: int):
        ref_so = ROOT / "oracle" / "_ref" / "libxgnn_ref.so"
        self.h = h
        if ref_so.exists():
            self.kind, self.cores = "reference", threads
            self.lib = lib = ctypes.CDLL(str(ref_so))
            lib.ref_hashtable_create.restype = ctypes.c_void_p
            lib.ref_hashtable_create.argtypes = [ctypes.c_int, ctypes.c_size_t]
            lib.ref_arch0_batch.restype = ctypes.c_size_t
            lib.ref_arch0_batch.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int,
                                            ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_size_t,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_size_t),
                                            ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
            lib.ref_set_threads.argtypes = [ctypes.c_int]
            lib.ref_set_threads(threads)
            self.table = lib.ref_hashtable_create(2, h["V"])  # CPUHashTable2: arch0's default (run_config.cc:56)
            self.f32 = lib.ref_dtype_f32() if h["feat"] is None or h["feat"].dtype == np.float32 else lib.ref_dtype_f16()
        else:
            sys.path.insert(0, str(ROOT / "tests"))
            import oracle_lib  # the C restatement; single thread
            self.kind, self.cores, self.lib, self.O = "port", 1, None, oracle_lib

    def batch(self, seeds: np.ndarray):
        """-> (sampled edges, input nodes, seconds)"""
        h = self.h
        t0 = time.perf_counter()
        if self.lib is not None:
            fan = (ctypes.c_size_t * len(FANOUT))(*FANOUT)
            n_in = ctypes.c_size_t(0)
            ts, te = ctypes.c_double(0), ctypes.c_double(0)
            feat = h["feat"]
            edges = self.lib.ref_arch0_batch(
                self.table, 0, h["indptr"].ctypes.data, h["indices"].ctypes.data, seeds.ctypes.data, len(seeds), fan,
                len(FANOUT), feat.ctypes.data if feat is not None else None, h["dim"], self.f32,
                feat.itemsize if feat is not None else 4, h["label"].ctypes.data, None, ctypes.byref(n_in),
                ctypes.byref(ts), ctypes.byref(te))
            n_input = n_in.value
        else:
            mb = self.O.minibatch(h["indptr"], h["indices"], seeds, FANOUT, seed=1)
            edges, n_input = int(sum(mb["num_edge"])), len(mb["input_nodes"])
            if h["feat"] is not None:
                self.O.extract(h["feat"], mb["input_nodes"])
        return int(edges), int(n_input), time.perf_counter() - t0


def cpu_baseline(args, h, budget_s: float) -> dict:
    arm = CpuArm(h, host_cores())
    batches = seed_batches(h["train"], 0, 1, args.batch, 64)
    arm.batch(batches[0][: max(1, args.batch // 8)])  # touch the pages / spin up the thread pool
    edges = secs = 0.0
    n = 0
    for b in batches[1:]:
        e, _, s = arm.batch(b)
        edges, secs, n = edges + e, secs + s, n + 1
        if secs > budget_s:
            break
    return {"value": edges / secs, "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
            "sample": f"{n} minibatches of {args.batch} seeds, {h['feat_note']}, {secs:.1f} s of CPU work"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation (arch0) on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    h = host_dataset(args)
    arm = CpuArm(h, host_cores())
    batches = seed_batches(h["train"], 0, 1, args.batch, args.steps + args.warmup)
    t_first = arm.batch(batches[0])[2]
    per_step = args.batch
    if t_first * (args.steps + args.warmup) > 240.0:  # bound the whole run to a few minutes
        per_step = max(256, int(args.batch * 240.0 / (t_first * (args.steps + args.warmup))))
    for b in batches[1:args.warmup]:
        arm.batch(b[:per_step])
    edges = secs = 0.0
    for b in batches[args.warmup:]:
        e, _, s = arm.batch(b[:per_step])
        edges, secs = edges + e, secs + s
    val = edges / secs
    sample = f"{args.steps} steps of {per_step} seeds (of batch {args.batch}), {h['feat_note']}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": workload_config(args, h["V"], h["E"], 1),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "native_so": "libxgnn_b200.so is loaded by this arm for DATA GENERATION only (the synthetic graph and features are "
                     "generated on the GPU and copied to the host; the shuffler routine orders the seeds); the timed loop "
                     "runs oracle/_ref/libxgnn_ref.so = the reference's own .cc files, with working buffers pooled across "
                     "batches as its WorkspacePool does",
    }))


def placement_note(args, world, V, row_bytes):
    gb = V * row_bytes / 1e9
    cold = ""
    if args.cache_pct < 1.0:
        cold = (f"; hottest {args.cache_pct:.0%} of the nodes (degree rank) cached in HBM, the other {1 - args.cache_pct:.0%} "
                f"read from pinned host memory over PCIe (mock host table of 2^20 rows, SAMGRAPH_EMPTY_FEAT style)")
    if world == 1:
        if args.cache_pct >= 1.0:
            return ("GGMS: topology + features in 1 GPU's HBM, cache 100% (one part, laid out by node id: no cache-table "
                    "lookup; partial caches and N > 1 go through the direct-address table)")
        return f"GGMS: topology + {gb * args.cache_pct:.1f} GB of cached rows in 1 GPU's HBM (direct-address cache table)" + cold
    rep = args.replicate
    topo = ("topology replicated in every GPU" if args.replicate_topology else f"topology sharded over {world} GPUs (v % P)")
    gbc = gb * args.cache_pct
    if rep <= 0:
        return (f"GGMS: {topo}; features sharded over {world} GPUs (rank % P), {gbc / world:.1f} GB of rows per GPU, "
                f"peers read over NVLink" + cold)
    rep = min(rep, args.cache_pct)
    return (f"GGMS two-tier: {topo}; features: hottest {rep:.0%} of nodes by degree rank replicated in every GPU "
            f"({gb * rep:.1f} GB), next {args.cache_pct - rep:.0%} sharded rank % P ({gb * (args.cache_pct - rep) / world:.1f} GB "
            f"per GPU) and read from peers over NVLink" + cold)


SAMPLER_NAME = {0: "GraphSAGE 3-hop uniform (KHop0) fanout 15,10,5", 7: "GraphSAGE 3-hop uniform (KHop3) fanout 15,10,5",
                1: "3-hop KHop1 (with replacement, sort + adjacent dedup) fanout 15,10,5",
                2: "3-hop weighted KHop (alias tables, with replacement, sort + adjacent dedup) fanout 15,10,5",
                3: "PinSAGE 3-layer random-walk top-K (4 walks x 3 steps, restart 0.5, K=5)"}


def workload_config(args, V, E, world):
    from xgnn_b200 import dataset as D
    dim, dtype = D.SHAPES[args.shape][2], str(D.SHAPES[args.shape][3]).replace("torch.float", "fp")
    kind = SAMPLER_NAME.get(args.sample_type, "3-hop")
    return {"workload": f"{kind} on synthetic {args.shape}-shape graph "
                        f"({V / 1e6:.1f}M nodes, {E / 1e9:.2f}B edges, {dim}-d {dtype}), batch {args.batch}, "
                        f"sample + unique/remap + feature/label extract per minibatch",
            "shape": args.shape, "scale": args.scale, "num_node": V, "num_edge": E, "batch": args.batch,
            "fanout_from_seeds": fanout_of(args)[::-1], "sample_type": args.sample_type,
            "placement": placement_note(args, world, V, dim * (2 if "16" in dtype else 4)),
            "replicate": (args.replicate if world > 1 else 0.0),
            "replicate_topology": bool(args.replicate_topology) if world > 1 else False,
            "cache_pct": args.cache_pct,
            "l2": "inputs larger than L2 (store is GBs; fresh seeds every step)"}


# --------------------------------------------------------------------------------------------------
# our arm
def exchange_parts(st, dist, torch):
    """Multi-GPU: move this rank's shards into their own cudaMalloc blocks, all-gather the CUDA IPC handles
    (init-time only, like DistGraph::_DataIpcShare, dist_graph.cu:274-307), map every peer's shards."""
    from xgnn_b200 import ops
    from xgnn_b200._clib import check, lib
    P, r = st.world, st.rank
    # the replica (two-tier store) and a replicated topology are private: never exported
    names = ["feat_part"] if st.local.get("replicate_topology") else ["indptr", "indices", "feat_part"]
    mine, raw = [], st.local.setdefault("raw", {})
    for nm in names:
        t = st.local[nm]
        if nm in raw:  # generated straight into its own exportable block (build_store(exportable=True)): nothing to copy
            p = ctypes.c_void_p(raw[nm])
        else:
            p = ctypes.c_void_p()
            nbytes = t.numel() * t.element_size()
            check(lib.xg_device_malloc(ctypes.byref(p), nbytes))
            raw[nm] = p.value
            raw_t = torch.as_tensor(ops._DevMem(p.value, (nbytes,), "|u1", None), device="cuda")
            raw_t.copy_(t.contiguous().view(torch.uint8).view(-1))
        torch.cuda.synchronize()
        hnd = (ctypes.c_ubyte * 64)()
        check(lib.xg_ipc_export(p, hnd))
        mine.append(bytes(hnd))
        st.local[nm] = None  # the torch copy is no longer needed
    torch.cuda.empty_cache()
    blob = torch.tensor(list(b"".join(mine)), dtype=torch.uint8, device="cuda")
    allb = [torch.empty_like(blob) for _ in range(P)]
    dist.all_gather(allb, blob)
    ptrs = {nm: [None] * P for nm in names}
    imported = []
    for q in range(P):
        hb = bytes(allb[q].cpu().numpy().tolist())
        for k, nm in enumerate(names):
            if q == r:
                ptrs[nm][q] = raw[nm]
            else:
                out = ctypes.c_void_p()
                check(lib.xg_ipc_import((ctypes.c_ubyte * 64).from_buffer_copy(hb[64 * k:64 * k + 64]), ctypes.byref(out)))
                ptrs[nm][q] = out.value
                imported.append(out.value)
    st.local["imported"] = imported
    return ptrs


def release_store(R):
    """Give a run's memory back (secondary configs follow in the same process): peers first stop reading, then the
    IPC mappings are closed and the exported shards freed."""
    from xgnn_b200._clib import lib
    import torch
    dist, st = R.get("dist"), R.get("st")
    for k in ("sm", "topo", "graph"):
        R.pop(k, None)
    c = R.pop("clocks", None)
    if c is not None and c.proc is not None and c.proc.poll() is None:  # a failed run: do not leave nvidia-smi behind
        c.stop()
    torch.cuda.synchronize()
    if st is None:
        return
    if dist is not None:
        dist.barrier()
    for p in st.local.get("imported", []):
        lib.xg_ipc_close(ctypes.c_void_p(p))
    if dist is not None:
        dist.barrier()
    for p in st.local.get("raw", {}).values():
        lib.xg_device_free(ctypes.c_void_p(p))
    st.local.clear()
    st.graph = st.store = st.label = None
    for k in ("d_flat", "topo", "weights"):
        R.pop(k, None)
    import gc
    gc.collect()
    torch.cuda.empty_cache()


def run_xgnn(args, dist=None, R0=None, keep_clocks=False):
    """R0: dict filled as the run goes (store first), so that a failure half-way can still be cleaned up"""
    import torch
    from xgnn_b200 import dataset as D
    from xgnn_b200 import ops
    from xgnn_b200 import selfcheck

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    V, E, dim, dtype, ncls, frac = D.SHAPES[args.shape]
    t_build = time.perf_counter()
    rep_topo = bool(args.replicate_topology) and world > 1
    st = D.build_store(args.shape, local, rank, world, scale=args.scale, cache_pct=args.cache_pct,
                       replicate_pct=min(args.replicate, args.cache_pct),
                       replica_words=os.environ.get("XG_BENCH_REPLICA_WORDS", "1") != "0",
                       flat_when_full=os.environ.get("XG_BENCH_FORCE_TABLE", "0") == "0",  # 1: the N > 1 lookup path on one GPU
                       replicate_topology=rep_topo, exportable=world > 1)
    if R0 is not None:
        R0.update(st=st, dist=dist)
    ptrs = None
    if world > 1:
        ptrs = exchange_parts(st, dist, torch)
        D.attach(st, ptrs.get("indptr"), ptrs.get("indices"), ptrs["feat_part"])
        if rep_topo:
            num_edge = st.local["num_local_edge"]
        else:
            ne = torch.tensor([st.local["num_local_edge"]], dtype=torch.int64, device="cuda")
            dist.all_reduce(ne)
            num_edge = int(ne.item())
    else:
        D.attach(st)
        num_edge = st.local["num_local_edge"]
    weights = {}
    if args.sample_type == 2:
        # weights 1 / deg(src), i.e. equal inside a list -> every alias-table entry accepts with
        # probability 1 and is its own alias (create_alias_table.cc semantics); the kernel reads both tables per draw
        idx = st.local["indices"]
        weights = {"prob_table": torch.ones(idx.numel(), dtype=torch.float32, device=idx.device), "alias_table": idx.clone()}
    t_build = time.perf_counter() - t_build

    # sample phase: short latency-bound kernels, scheduled first (XG_BENCH_SAMPLE_PRIO=0: default priority)
    stream = torch.cuda.Stream(priority=int(os.environ.get("XG_BENCH_SAMPLE_PRIO", "-1")))
    graph = st.graph
    if args.sample_type == 2 and world == 1:  # the weight tables are aligned with the flat CSR (num_part = 0)
        graph = ops.make_graph(st.local["indptr"], st.local["indices"], num_node=st.num_node)
    sm = ops.MinibatchSampler(graph, fanout_of(args), args.batch, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                              label=st.label, seed=2024, num_slots=args.in_flight + 1, device=local, stream=stream,
                              max_input_nodes=args.max_input_nodes, **sampler_kw(args), **weights)
    total = args.warmup + args.steps
    batches = seed_batches(st.train, rank, world, args.batch, total)
    flat = np.concatenate(batches)
    offs = np.cumsum([0] + [len(b) for b in batches])
    d_flat = torch.from_numpy(flat.view(np.int32)).cuda()           # `value`: seeds already resident in HBM
    h_flat = torch.from_numpy(flat.view(np.int32)).pin_memory()     # `e2e`: seeds in pinned host memory

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def run(first, count, src):
        """submit `count` minibatches, `--in-flight` of them queued ahead; returns (edges, input nodes)"""
        edges = rows = 0
        pend = []
        for i in range(first, first + count):
            s = src[offs[i]:offs[i + 1]]
            pend.append(sm.submit(s, batch_key=i))
            if len(pend) == args.in_flight:
                mb = sm.wait(pend.pop(0))
                edges += sum(mb.num_edge); rows += mb.num_input
                mb.release()
        for hnd in pend:
            mb = sm.wait(hnd)
            edges += sum(mb.num_edge); rows += mb.num_input
            mb.release()
        return edges, rows

    # ---- self-check (untimed), EVERY rank: the full property set of xgnn_b200/selfcheck.py on one minibatch - prefix
    # property, per-node edge counts, CSR-entry equality of every whole list, CSC offsets, feature rows == the
    # generator's closed form (0 ULP) whichever tier served them (local HBM / NVLink peer / pinned host), labels
    topo = selfcheck.Topology.of_store(st, ptrs)
    ok, why = True, ""
    try:
        mb = sm.wait(sm.submit(d_flat[offs[0]:offs[1]], batch_key=0))
        selfcheck.check_minibatch_properties(mb, st, D, batches[0], fanout_of(args), uniform=args.sample_type in (0, 5, 7),
                                             topo=topo, membership_samples=0 if args.sample_type == 3 else 100)
        mb.release()
    except AssertionError as e:
        ok, why = False, str(e)
    oks = [ok]
    if dist is not None:
        t_ok = torch.zeros(world, dtype=torch.int32, device="cuda")
        t_ok[rank] = 1 if ok else 0
        dist.all_reduce(t_ok)
        oks = [bool(x) for x in t_ok.cpu().tolist()]
    assert all(oks), f"self-check failed on rank(s) {[i for i, o in enumerate(oks) if not o]}: {why}"
    # ---- value: device-resident seeds, CUDA events on the sampler's stream ------------------------
    run(0, args.warmup, d_flat)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ops.launch_count()
    ev0.record(stream)
    edges, rows = run(args.warmup, args.steps, d_flat)
    ev1.record(sm.extract_stream)  # every batch ends on the extract stream (its last kernels + the count read-back)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = ops.launch_count() - launches0
    # ---- e2e: host seeds -> submit (H2D inside) -> wait (D2H of the counts), wall clock --------------
    run(0, min(3, args.warmup), h_flat)
    barrier()
    t0 = time.perf_counter()
    e2e_edges, _ = run(args.warmup, args.steps, h_flat)
    barrier()
    e2e_s = time.perf_counter() - t0
    clk = None if keep_clocks else clocks.stop()  # the epoch pass that follows is part of the measurement: keep sampling

    def allmax(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    ms_max, e2e_max = allmax(ms), allmax(e2e_s)
    edges_all, e2e_edges_all, rows_all = allsum(edges), allsum(e2e_edges), allsum(rows)
    return dict(torch=torch, dist=dist, rank=rank, world=world, local=local, st=st, sm=sm, ms=ms_max, e2e_s=e2e_max,
                edges=edges_all, e2e_edges=e2e_edges_all, rows=rows_all, my_edges=edges, my_rows=rows,
                launches=launches, clk=clk, clocks=clocks, num_edge=num_edge, t_build=t_build, batches=batches, d_flat=d_flat,
                offs=offs, stream=stream, self_check=oks, allmax=allmax, allsum=allsum, barrier=barrier, graph=graph,
                weights=weights, topo=topo)


def fanout_of(args):
    return [5, 5, 5] if args.sample_type == 3 else FANOUT  # random walk: K neighbours per layer


def sampler_kw(args):
    kw = {"sample_type": args.sample_type}
    if args.sample_type == 3:
        kw.update(random_walk_length=3, num_random_walk=4, random_walk_restart_prob=0.5)
    return kw


STAGES = {"reset_seeds": 0, "sample": 1, "insert": 2, "number": 3, "input_nodes": 4, "feature": 5, "label": 6}


def stage_pass(R, args):
    """Per-stage device time with CUDA events on the launching stream (a second sampler in
    XG_SAMPLER_STAGE_TIMING mode: same kernels, launched one by one instead of as a graph)."""
    from xgnn_b200 import ops
    from xgnn_b200._clib import check, lib
    torch, st = R["torch"], R["st"]
    sm2 = ops.MinibatchSampler(R["graph"], fanout_of(args), args.batch, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                               label=st.label, seed=2024, num_slots=1, device=R["local"], stream=R["stream"],
                               use_graph=False, stage_timing=True, max_input_nodes=args.max_input_nodes,
                               **sampler_kw(args), **R["weights"])
    L = len(FANOUT)
    acc = {}
    info = {"edges": [0] * L, "inputs": [0] * L, "rows": 0, "seeds": 0, "src_local": 0, "src_peer": 0, "src_host": 0}
    stats = torch.zeros(4, dtype=torch.int64, device="cuda")
    n = min(args.stage_steps, args.steps)
    for i in range(args.warmup, args.warmup + n):
        hnd = sm2.submit(R["d_flat"][R["offs"][i]:R["offs"][i + 1]], batch_key=i)
        mb = sm2.wait(hnd)
        ms = ctypes.c_float(0)
        for name, code in STAGES.items():
            for l in (range(L) if code in (1, 2, 3) else [0]):
                check(lib.xg_batch_stage_ms(hnd, code, l, ctypes.byref(ms)))
                key = f"{name}.L{l}" if code in (1, 2, 3) else name
                acc[key] = acc.get(key, 0.0) + ms.value
        for l in range(L):
            info["edges"][l] += mb.num_edge[l]
            info["inputs"][l] += mb.num_dst[l]
        info["rows"] += mb.num_input
        info["seeds"] += mb.num_output
        if st.store is not None:  # where the rows came from (counted by a separate launch, outside the timed marks)
            stats.zero_()
            ops.ExtractGGMS(st.store, mb.input_nodes, st.dim, st.dtype, stats=stats, out=mb.feat)
            c = stats.cpu().tolist()
            if st.local.get("flat"):  # the one-part node-id layout sits in the handle's `host` slot but lives in HBM
                c[0], c[2] = c[0] + c[2], 0
            info["src_local"] += c[0]; info["src_peer"] += c[1]; info["src_host"] += c[2]
        mb.release()
    del sm2
    return {k: v / n for k, v in acc.items()}, {k: ([x / n for x in v] if isinstance(v, list) else v / n)
                                                 for k, v in info.items()}


NVLINK_GBS = 770.0  # per direction per GPU: the measured peer copy B200_PROFILING.md states (our probe: 822-830)
# what one direction delivers in ROW bytes when every GPU pulls and serves at once (tools/nvlink_req_probe.cu, ncu nvlrx/nvltx bytes,
# profiles/r02n_nvlink_requests.md): a direction carries the rows it receives x 1.125 plus the 128-byte read requests of the
# rows it serves x 0.25 -> ~860 GB/s raw / 1.375 = 621 GB/s measured
NVLINK_ALL_PULL_GBS = 621.0
PCIE_GBS = 55.0     # pinned H2D copy on the box's PCIe gen5 x16 link (profiles/r01g_host_probe.jsonl: 51.5 GB/s gathered = 0.93)


def kernel_sources_digest() -> str:
    """sha256 over the gather kernel's sources: profiles/traffic.json is only quoted when it was captured on this code"""
    import hashlib
    h = hashlib.sha256()
    for f in ("xgnn_b200/csrc/extract.cu", "xgnn_b200/csrc/extract.cuh", "xgnn_b200/csrc/common.cuh"):
        h.update((ROOT / f).read_bytes())
    return h.hexdigest()[:16]


def summarize(R, args, stage_ms, info):
    """roofline block + per-stage table of one run"""
    st, world = R["st"], R["world"]
    peak, peak_src = measured_peaks()
    row_bytes = st.row_bytes
    # algorithmic bytes per launch: extract n * (2 * row_bytes + 4); sample N_in * 8 + E_out * 12;
    # unique + remap E * 16 + U * 4 (the numbering pass does both)
    L = len(FANOUT)
    alg = {"feature": info["rows"] * (2 * row_bytes + 4), "label": info["seeds"] * (2 * 8 + 4)}
    for l in range(L):
        alg[f"sample.L{l}"] = info["inputs"][l] * 8 + info["edges"][l] * 12
        # unique + remap (insert + number passes together): E * 16 + U * 4; quoted on the numbering pass
        alg[f"number.L{l}"] = info["edges"][l] * 16 + (info["inputs"][l - 1] - info["inputs"][l] if l > 0 else
                                                      info["rows"] - info["inputs"][0]) * 4
    stages = {}
    for k, ms in stage_ms.items():
        stages[k] = {"ms": round(ms, 4)}
        if k in alg and ms > 0:
            stages[k]["alg_GBps"] = round(alg[k] / ms / 1e6, 1)
    top = max((k for k in stage_ms if k in ("feature",) or k.startswith("sample")), key=lambda k: stage_ms[k])
    achieved = alg[top] / stage_ms[top] / 1e6
    bound, peak_used, frac = "hbm", peak, achieved / peak
    roof_note = f"{top}: all rows local HBM"
    split = None
    if top == "feature" and (info["src_peer"] > 0 or info["src_host"] > 0):
        # T_roof = max((R_local + W_all) / BW_hbm, R_peer / BW_nvlink_in, R_host / BW_pcie); frac = T_roof / T
        r_local, r_peer, r_host = (info[k] * row_bytes for k in ("src_local", "src_peer", "src_host"))
        t = {"hbm": (r_local + info["rows"] * (row_bytes + 4)) / (peak * 1e6), "nvlink": r_peer / (NVLINK_GBS * 1e6),
             "pcie": r_host / (PCIE_GBS * 1e6)}
        bound = max(t, key=t.get)
        frac = t[bound] / stage_ms[top]
        if bound == "nvlink":
            peak_used, achieved = NVLINK_GBS, r_peer / stage_ms[top] / 1e6
        elif bound == "pcie":
            peak_used, achieved = PCIE_GBS, r_host / stage_ms[top] / 1e6
        rows = max(info["rows"], 1)
        split = {"local": round(info["src_local"] / rows, 4), "peer": round(info["src_peer"] / rows, 4),
                 "host": round(info["src_host"] / rows, 4)}
        roof_note = (f"feature gather: {split['local']:.1%} of rows from local HBM, {split['peer']:.1%} from NVLink peers, "
                     f"{split['host']:.1%} from pinned host; roofline time max(HBM {t['hbm']:.3f} ms, NVLink-in "
                     f"{t['nvlink']:.3f} ms, PCIe {t['pcie']:.3f} ms) vs measured {stage_ms[top]:.3f} ms; peaks: HBM {peak:.0f} "
                     f"GB/s measured copy, NVLink {NVLINK_GBS:.0f} GB/s per direction, PCIe {PCIE_GBS:.0f} GB/s")
    roof = {"bound": bound, "kernel": top, "achieved": round(achieved, 1), "peak": peak_used, "unit": "GB/s",
            "frac": round(frac, 4), "traffic": None, "peak_source": peak_src, "note": roof_note,
            "alg_bytes_per_launch": int(alg[top]), "launch_ms": round(stage_ms[top], 4)}
    if split:
        roof["row_sources"] = split
        if world > 1 and info["src_peer"] > 0:  # informative: the same max() with the NVLink rate of an all-pull run
            t_all = max(t["hbm"], r_peer / (NVLINK_ALL_PULL_GBS * 1e6), t["pcie"])
            roof["all_gpus_pull"] = {"nvlink_GBps": NVLINK_ALL_PULL_GBS, "roofline_ms": round(t_all, 4),
                                     "frac": round(t_all / stage_ms[top], 4),
                                     "note": "NVLink term at the measured row rate when all GPUs pull and serve at once (request + header "
                                             "bytes share each direction: profiles/r02n_nvlink_requests.md, a 2-GPU probe - a value a little above 1 means this "
                                             "run's links did better than the probe's); `frac` above uses 770 GB/s"}
    # whole-step view: every stage's algorithmic bytes over the step time, against the HBM copy peak
    step_ms = R["ms"] / args.steps
    alg_step = sum(v for k, v in alg.items() if k in stage_ms)
    roof["step"] = {"alg_bytes": int(alg_step), "ms": round(step_ms, 4), "frac_of_hbm_peak": round(alg_step / step_ms / 1e6 / peak, 4),
                    "gather_roofline_ms_over_step_ms": round((alg["feature"] / (peak * 1e6)) / step_ms, 4) if bound == "hbm" else None}
    return roof, stages


def epoch_pass(R, args, epochs=4):
    """SURVEY 8(d) / example/samgraph/common_config.py:174-175, sgnn/train_graphsage.py:334-370: whole epochs, every
    step of them, seeds from the aligned shuffler in its own order, the first epoch discarded as warm-up, the others
    timed with CUDA events (begin on the sample stream, end on the extract stream), max over ranks."""
    torch, sm, st = R["torch"], R["sm"], R["st"]
    sh = AlignedShuffler(st.train, R["rank"], R["world"], args.batch)
    times, edges_ep = [], []
    key = 1 << 20
    for ep in range(epochs):
        batches = sh.next_epoch()
        flat = torch.from_numpy(np.concatenate(batches).view(np.int32)).cuda()
        offs = np.cumsum([0] + [len(b) for b in batches])
        R["barrier"]()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(R["stream"])
        edges, pend = 0, []
        for i in range(len(batches)):
            pend.append(sm.submit(flat[offs[i]:offs[i + 1]], batch_key=key))
            key += 1
            if len(pend) == args.in_flight:
                mb = sm.wait(pend.pop(0)); edges += sum(mb.num_edge); mb.release()
        for hnd in pend:
            mb = sm.wait(hnd); edges += sum(mb.num_edge); mb.release()
        ev1.record(sm.extract_stream)
        R["barrier"]()
        times.append(R["allmax"](ev0.elapsed_time(ev1)) / 1e3)
        edges_ep.append(R["allsum"](edges))
    use_t, use_e = times[1:], edges_ep[1:]
    return {"measured": True, "steps_per_epoch": sh.steps, "epochs_timed": len(use_t), "first_epoch_discarded_s": round(times[0], 5),
            "sample_extract_time_s": round(sum(use_t) / len(use_t), 5), "epoch_times_s": [round(t, 5) for t in use_t],
            "sampled_edges_per_s": sum(use_e) / sum(use_t),
            "seeds": "DistAlignedShuffler order (xg_aligned_shuffle_epoch, the engine's routine), every step of every epoch"}


def one_config(args, dist, *, with_e2e=True, epochs=0):
    """build -> self-check -> timed run -> stage pass -> (epoch pass) -> release; returns the JSON fragment"""
    R = {}
    try:
        R.update(run_xgnn(args, dist, R, keep_clocks=epochs > 0))
        return _one_config(R, args, with_e2e, epochs)
    finally:  # also after a failure half-way (out of memory ...): the next config needs the memory back
        release_store(R)


def _one_config(R, args, with_e2e, epochs):
    st, world = R["st"], R["world"]
    stage_ms, info = stage_pass(R, args)
    roof, stages = summarize(R, args, stage_ms, info)
    frag = {"value": R["edges"] / (R["ms"] / 1e3), "unit": UNIT, "ms_per_step": R["ms"] / args.steps, "steps": args.steps,
            "config": workload_config(args, st.num_node, R["num_edge"], world),
            "roofline": roof, "stages": stages,
            "input_nodes_per_batch": info["rows"], "edges_per_batch": sum(info["edges"]),
            "feature_gather_GBps": round(info["rows"] * st.row_bytes / stage_ms["feature"] / 1e6, 1),
            "self_check": {"ranks_ok": R["self_check"], "what": "per rank, untimed: xgnn_b200/selfcheck.py property set on one "
                           "minibatch (prefix, per-node counts, CSR-entry equality of whole lists, CSC offsets, 0-ULP feature "
                           "rows vs the generator, labels)"},
            "build_s": round(R["t_build"], 1)}
    if with_e2e:
        frag["e2e"] = {"value": R["e2e_edges"] / R["e2e_s"], "unit": UNIT, "h2d_bytes_per_step": args.batch * 4 + 24,
                       "d2h_bytes_per_step": 120, "ms_per_step": 1e3 * R["e2e_s"] / args.steps,
                       "note": f"host seeds -> xg_sampler_submit (H2D inside) -> xg_batch_wait (D2H counts); {args.in_flight} batches in flight"}
    if epochs:
        frag["epoch"] = epoch_pass(R, args, epochs)
        R["clk"] = R["clocks"].stop()  # nvidia-smi sampled every 100 ms from the timed steps through the measured epochs
        frag["epoch"].update(feature_gather_GBps=frag["feature_gather_GBps"], input_nodes_per_batch=info["rows"],
                             edges_per_batch=sum(info["edges"]))
    frag["_launches"], frag["_clocks"] = int(R["launches"]), R["clk"]
    return frag


def variant(args, **kw):
    a = argparse.Namespace(**vars(args))
    for k, v in kw.items():
        setattr(a, k, v)
    return a


def secondary_configs(args, dist, world):
    """BASELINE.json's other configs, short runs (not bench lines: each carries its own value / roofline / placement)"""
    quick = dict(steps=30, warmup=5, stage_steps=5, max_input_nodes=0)
    if world == 1:
        specs = [
            ("cfg1 products / KHop0 (CPU arch0 beside it)", dict(shape="products", scale=1, sample_type=0, cache_pct=1.0), True),
            ("cfg3 twitter / KHop0, all local", dict(shape="twitter", scale=1, sample_type=0, cache_pct=1.0), False),
            ("cfg4 friendster / random walk top-K", dict(shape="friendster", scale=1, sample_type=3, cache_pct=1.0), False),
            ("cfg4 friendster / weighted KHop (alias)", dict(shape="friendster", scale=1, sample_type=2, cache_pct=1.0), False),
            ("cfg5 mag240M-homo / hot rows in HBM, cold rows in pinned host", dict(shape="mag240M-homo", scale=1, sample_type=0,
                                                                                   cache_pct=0.2, steps=12, warmup=3, stage_steps=3,
                                                                                   in_flight=1), False),
        ]
    else:
        per_gpu_gb = 375.0 / world
        full = min(1.0, 80.0 / per_gpu_gb)  # at most 80 GB of cached rows per GPU: batch slots (13 GB each) must fit beside them
        specs = [
            ("cfg3 twitter / KHop0, topology v % P over NVLink peers, features two-tier",
             dict(shape="twitter", scale=1, sample_type=0, cache_pct=1.0, replicate_topology=0), False),
            (f"cfg5 mag240M-homo / {full:.0%} of the rows sharded over the GPUs' HBM",
             dict(shape="mag240M-homo", scale=1, sample_type=0, cache_pct=full, replicate=0.0, replicate_topology=0, steps=12,
                  warmup=3, stage_steps=3, in_flight=1), False),
            ("cfg5 mag240M-homo / half of that in HBM, cold rows in pinned host: HBM + NVLink + PCIe in one launch",
             dict(shape="mag240M-homo", scale=1, sample_type=0, cache_pct=full / 2, replicate=0.0, replicate_topology=0, steps=12,
                  warmup=3, stage_steps=3, in_flight=1), False),
        ]
    out = []
    for name, kw, with_cpu in specs:
        a = variant(args, **{**quick, **kw})
        t0 = time.perf_counter()
        try:
            frag = one_config(a, dist, with_e2e=False)
            frag.pop("_launches", None); frag.pop("_clocks", None)
            frag["name"] = name
            if with_cpu and world == 1:
                frag["cpu_baseline"] = cpu_baseline(a, host_dataset(a), 8.0)
                frag["cpu_baseline"]["note"] = "the reference's arch0 CPU sampler + extractor on this box's host cores, same graph"
        except Exception as e:  # a secondary config must never cost the headline line
            frag = {"name": name, "failed": f"{type(e).__name__}: {e}"[:300]}
        frag["wall_s"] = round(time.perf_counter() - t0, 1)
        out.append(frag)
    return out


def samgraph_e2e(args):
    """The reference-facing API on one GPU: papers100M / 16 dumped in the reference's on-disk format, then
    sam.config -> init -> sample_once -> get_next_batch -> get_dgl_blocks (tools/samgraph_e2e.py), with the
    operator-level number of the same scaled shape beside it."""
    try:
        r = subprocess.run([sys.executable, str(ROOT / "tools" / "samgraph_e2e.py"), "--scale", "16", "--epochs", "4"],
                           capture_output=True, text=True, timeout=300)
        d = json.loads(r.stdout.strip().splitlines()[-1])
        op = one_config(variant(args, scale=16, steps=60, warmup=10, stage_steps=3, sample_type=7, cache_pct=1.0), None,
                        with_e2e=False)
        return {"value": d["sampled_edges_per_s"], "unit": UNIT, "ms_per_step": d["ms_per_step"], "shape": "papers100M / 16",
                "api": "xgnn_b200.samgraph.torch: config -> init -> [sample_once -> get_next_batch -> get_dgl_blocks] x steps, "
                       "first epoch discarded", "steps_per_epoch": d["steps_per_epoch"], "epochs": len(d["epochs"]),
                "operator_level_value": op["value"], "ratio_to_operator_level": d["sampled_edges_per_s"] / op["value"]}
    except Exception as e:
        return {"failed": f"{type(e).__name__}: {e}"[:300]}


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
    head = one_config(args, dist, with_e2e=True, epochs=0 if args.no_secondary else 4)
    out = {
        "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic", "config": head["config"], "e2e": head["e2e"],
        "gpu_launches": head.pop("_launches"), "clocks": head.pop("_clocks"), "roofline": head["roofline"],
        "stages": head["stages"],
        "epoch": head.get("epoch", {"measured": False, "steps_per_epoch": None}),
        "build_s": head["build_s"], "self_check": head["self_check"],
    }
    if world == 1:  # the dominant kernel's DRAM traffic: only a capture of THIS code counts
        tfile = ROOT / "profiles" / "traffic.json"
        try:
            t = json.loads(tfile.read_text())
            if t.get("sources_sha256_16") == kernel_sources_digest() and out["roofline"]["kernel"] == "feature":
                out["roofline"]["traffic"] = t["feature"]
                out["roofline"]["traffic_source"] = t.get("source")
        except Exception:
            pass
    if not args.no_secondary:
        if world > 1:
            # the reference's own routing (cuda/dist_graph.h:141-151,203-207; evaluation/figure10/run.sh:24-28): features
            # rank % P, topology v % P, nothing replicated
            try:
                sh = one_config(variant(args, replicate=0.0, replicate_topology=0, steps=min(args.steps, 40), stage_steps=5),
                                dist, with_e2e=False)
                sh.pop("_launches", None); sh.pop("_clocks", None)
                out["sharded"] = sh
            except Exception as e:
                out["sharded"] = {"failed": f"{type(e).__name__}: {e}"[:300]}
        out["configs"] = secondary_configs(args, dist, world)
        if world == 1 and rank == 0:
            out["e2e_samgraph"] = samgraph_e2e(args)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            out["cpu_baseline"] = cpu_baseline(args, host_dataset(args), args.cpu_seconds)
        except Exception as e:  # never lose the GPU line to a host-side problem
            out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                                   "sample": f"failed: {type(e).__name__}: {e}"}
    if dist is not None:
        dist.barrier()
    if rank == 0:
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
