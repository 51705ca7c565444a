#!/usr/bin/env python
"""bench.py -- sampled edges/s of the per-minibatch data path (sample -> unique/remap -> extract).

  python bench.py [--gpus N --steps K --warmup W]            our arm (CUDA, through the xg_* C-ABI)
  python bench.py --impl reference [...]                     the reference's own CPU path (arch0)

A step = one minibatch of `--batch` (8000) seeds of a synthetic papers100M-shape graph, GraphSAGE
3-hop uniform fanout 15,10,5 from the seeds (reference notation "5 10 15", consumed last -> first),
features 128 x fp32 gathered for every input node, labels for the seeds.  One JSON line on stdout.
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
    ap.add_argument("--replicate", type=float, default=float(os.environ.get("XG_BENCH_REPLICATE", "0.6")),
                    help="N > 1 only: fraction of the nodes (hottest first, by degree rank) whose feature rows every GPU "
                         "keeps in its own HBM; the rest is sharded rank %% P (0 = the reference's pure sharding)")
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


def epoch_order(train: np.ndarray, epoch: int, rank: int, world: int) -> np.ndarray:
    """This worker's contiguous slice of the globally shuffled, padded train set
    (DistAlignedShuffler, dist/dist_shuffler_aligned.cc:37-121; bench-side: numpy permutation)."""
    pad = (-len(train)) % world
    data = np.concatenate([train, train[:pad]]) if pad else train
    perm = np.random.default_rng(1000 + epoch).permutation(len(data))
    n_local = len(data) // world
    return np.ascontiguousarray(data[perm][rank * n_local:(rank + 1) * n_local])


def seed_batches(train, rank, world, batch, count):
    """`count` minibatches of seeds, walking through as many epochs as needed."""
    out, epoch = [], 0
    while len(out) < count:
        order = epoch_order(train, epoch, rank, world)
        for o in range(0, len(order), batch):
            out.append(order[o:o + batch])
            if len(out) == count:
                break
        epoch += 1
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
    }))


def placement_note(args, world, V, row_bytes):
    if world == 1:
        return ("GGMS: topology + features in 1 GPU's HBM, cache 100% (one part, laid out by node id: no cache-table "
                "lookup; partial caches and N > 1 go through the direct-address table)")
    rep = args.replicate
    gb = V * row_bytes / 1e9
    if rep <= 0:
        return (f"GGMS: topology + features sharded over {world} GPUs (v % P / rank % P), cache 100%, "
                f"{gb / world:.1f} GB of rows per GPU, peers read over NVLink")
    return (f"GGMS two-tier: topology sharded over {world} GPUs (v % P); features: hottest {rep:.0%} of nodes by degree rank "
            f"replicated in every GPU ({gb * rep:.1f} GB), remaining {1 - rep:.0%} sharded rank % P "
            f"({gb * (1 - rep) / world:.1f} GB per GPU) and read from peers over NVLink; cache 100%")


def workload_config(args, V, E, world):
    from xgnn_b200 import dataset as D
    dim, dtype = D.SHAPES[args.shape][2], str(D.SHAPES[args.shape][3]).replace("torch.float", "fp")
    kind = {0: "GraphSAGE 3-hop uniform fanout 15,10,5", 7: "GraphSAGE 3-hop uniform (KHop3) fanout 15,10,5",
            1: "3-hop KHop1 (with replacement, sort + adjacent dedup) fanout 15,10,5",
            3: "PinSAGE 3-layer random-walk top-K (4 walks x 3 steps, restart 0.5, K=5)"}.get(args.sample_type, "3-hop")
    return {"workload": f"{kind} on synthetic {args.shape}-shape graph "
                        f"({V / 1e6:.1f}M nodes, {E / 1e9:.2f}B edges, {dim}-d {dtype}), batch {args.batch}, "
                        f"sample + unique/remap + feature/label extract per minibatch",
            "shape": args.shape, "scale": args.scale, "num_node": V, "num_edge": E, "batch": args.batch,
            "fanout_from_seeds": fanout_of(args)[::-1], "sample_type": args.sample_type,
            "placement": placement_note(args, world, V, dim * (2 if "16" in dtype else 4)),
            "replicate": (args.replicate if world > 1 else 0.0),
            "l2": "inputs larger than L2 (store is GBs; fresh seeds every step)"}


# --------------------------------------------------------------------------------------------------
# our arm
def exchange_parts(st, dist, torch):
    """Multi-GPU: move this rank's shards into their own cudaMalloc blocks, all-gather the CUDA IPC handles
    (init-time only, like DistGraph::_DataIpcShare, dist_graph.cu:274-307), map every peer's shards."""
    from xgnn_b200._clib import check, lib
    P, r = st.world, st.rank
    names = ["indptr", "indices", "feat_part"]  # the replica (two-tier store) is private: never exported
    mine, raw = [], {}
    for nm in names:
        t = st.local[nm]
        p = ctypes.c_void_p()
        nbytes = t.numel() * t.element_size()
        check(lib.xg_device_malloc(ctypes.byref(p), nbytes))
        from xgnn_b200 import ops
        raw_t = torch.as_tensor(ops._DevMem(p.value, (nbytes,), "|u1", None), device="cuda")
        raw_t.copy_(t.contiguous().view(torch.uint8).view(-1))
        torch.cuda.synchronize()
        hnd = (ctypes.c_ubyte * 64)()
        check(lib.xg_ipc_export(p, hnd))
        raw[nm] = p.value
        mine.append(bytes(hnd))
        st.local[nm] = None  # the torch copy is no longer needed
    torch.cuda.empty_cache()
    blob = torch.tensor(list(b"".join(mine)), dtype=torch.uint8, device="cuda")
    allb = [torch.empty_like(blob) for _ in range(P)]
    dist.all_gather(allb, blob)
    ptrs = {nm: [None] * P for nm in names}
    for q in range(P):
        hb = bytes(allb[q].cpu().numpy().tolist())
        for k, nm in enumerate(names):
            if q == r:
                ptrs[nm][q] = raw[nm]
            else:
                out = ctypes.c_void_p()
                check(lib.xg_ipc_import((ctypes.c_ubyte * 64).from_buffer_copy(hb[64 * k:64 * k + 64]), ctypes.byref(out)))
                ptrs[nm][q] = out.value
    st.local["raw"] = raw
    return ptrs


def run_xgnn(args):
    import torch
    from xgnn_b200 import dataset as D
    from xgnn_b200 import ops
    from xgnn_b200._clib import lib

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    V, E, dim, dtype, ncls, frac = D.SHAPES[args.shape]
    t_build = time.perf_counter()
    st = D.build_store(args.shape, local, rank, world, scale=args.scale, replicate_pct=args.replicate,
                       replica_words=os.environ.get("XG_BENCH_REPLICA_WORDS", "1") != "0",
                       flat_when_full=os.environ.get("XG_BENCH_FORCE_TABLE", "0") == "0")  # 1: the N > 1 lookup path on one GPU
    if world > 1:
        ptrs = exchange_parts(st, dist, torch)
        D.attach(st, ptrs["indptr"], ptrs["indices"], ptrs["feat_part"])
        ne = torch.tensor([st.local["num_local_edge"]], dtype=torch.int64, device="cuda")
        dist.all_reduce(ne)
        num_edge = int(ne.item())
    else:
        D.attach(st)
        num_edge = st.local["num_local_edge"]
    t_build = time.perf_counter() - t_build

    # sample phase: short latency-bound kernels, scheduled first (XG_BENCH_SAMPLE_PRIO=0: default priority)
    stream = torch.cuda.Stream(priority=int(os.environ.get("XG_BENCH_SAMPLE_PRIO", "-1")))
    sm = ops.MinibatchSampler(st.graph, fanout_of(args), args.batch, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                              label=st.label, seed=2024, num_slots=args.in_flight + 1, device=local, stream=stream,
                              max_input_nodes=args.max_input_nodes, **sampler_kw(args))
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

    def run(first, count, src, on_dev):
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

    # ---- self-check (untimed): one minibatch's rows against the generator's closed form, labels, prefix property -
    mb = sm.wait(sm.submit(d_flat[offs[0]:offs[1]], batch_key=0))
    want = D.gen_features(mb.num_input, st.dim, st.dtype, torch.device("cuda", local), row_ids=mb.input_nodes)
    bits = torch.int32 if st.dtype == torch.float32 else torch.int16
    ok = bool(torch.equal(mb.feat.view(bits), want.view(bits)))
    ok = ok and bool(torch.equal(mb.input_nodes[: mb.num_output], d_flat[offs[0]:offs[1]]))
    ok = ok and bool(torch.equal(mb.label, (d_flat[offs[0]:offs[1]].to(torch.int64) & 0xFFFFFFFF) % st.num_class))
    del want
    mb.release()
    if dist is not None:
        t_ok = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        ok = bool(t_ok.item())
    assert ok, "self-check failed: gathered rows / labels / seed prefix differ from the generator"
    # ---- value: device-resident seeds, CUDA events on the sampler's stream ------------------------
    run(0, args.warmup, d_flat, True)
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ops.launch_count()
    ev0.record(stream)
    edges, rows = run(args.warmup, args.steps, d_flat, True)
    ev1.record(sm.extract_stream)  # every batch ends on the extract stream (its last kernels + the count read-back)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = ops.launch_count() - launches0
    # ---- e2e: host seeds -> submit (H2D inside) -> wait (D2H of the counts), wall clock --------------
    run(0, min(3, args.warmup), h_flat, False)
    barrier()
    t0 = time.perf_counter()
    e2e_edges, _ = run(args.warmup, args.steps, h_flat, False)
    barrier()
    e2e_s = time.perf_counter() - t0
    clk = clocks.stop()

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
                launches=launches, clk=clk, num_edge=num_edge, t_build=t_build, batches=batches, d_flat=d_flat,
                offs=offs, stream=stream)


def fanout_of(args):
    return [5, 5, 5] if args.sample_type == 3 else FANOUT  # random walk: K neighbours per layer


def sampler_kw(args):
    kw = {"sample_type": args.sample_type}
    if args.sample_type == 3:
        kw.update(random_walk_length=3, num_random_walk=4, random_walk_restart_prob=0.5)
    return kw


STAGES = {"reset_seeds": 0, "sample": 1, "number": 2, "remap": 3, "input_nodes": 4, "feature": 5, "label": 6}


def stage_pass(R, args):
    """Per-stage device time with CUDA events on the launching stream (a second sampler in
    XG_SAMPLER_STAGE_TIMING mode: same kernels, launched one by one instead of as a graph)."""
    from xgnn_b200 import ops
    from xgnn_b200._clib import check, lib
    torch, st = R["torch"], R["st"]
    sm2 = ops.MinibatchSampler(st.graph, fanout_of(args), args.batch, store=st.store, feat_dim=st.dim, feat_dtype=st.dtype,
                               label=st.label, seed=2024, num_slots=1, device=R["local"], stream=R["stream"],
                               use_graph=False, stage_timing=True, max_input_nodes=args.max_input_nodes,
                               **sampler_kw(args))
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
            info["src_local"] += c[0]; info["src_peer"] += c[1]; info["src_host"] += c[2]
        mb.release()
    del sm2
    return {k: v / n for k, v in acc.items()}, {k: ([x / n for x in v] if isinstance(v, list) else v / n)
                                                 for k, v in info.items()}


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    R = run_xgnn(args)
    torch, st, world = R["torch"], R["st"], R["world"]
    stage_ms, info = stage_pass(R, args)
    peak, peak_src = measured_peaks()
    row_bytes = st.row_bytes
    # algorithmic bytes per launch: extract n * (2 * row_bytes + 4);
    # sample N_in * 8 + E_out * 12; unique+remap E * 16 + U * 4 (number + remap stages together)
    L = len(FANOUT)
    alg = {"feature": info["rows"] * (2 * row_bytes + 4), "label": info["seeds"] * (2 * 8 + 4)}
    for l in range(L):
        alg[f"sample.L{l}"] = info["inputs"][l] * 8 + info["edges"][l] * 12
    stages = {}
    for k, ms in stage_ms.items():
        stages[k] = {"ms": round(ms, 4)}
        if k in alg and ms > 0:
            stages[k]["alg_GBps"] = round(alg[k] / ms / 1e6, 1)
    top = max((k for k in stage_ms if k in alg), key=lambda k: stage_ms[k])
    achieved = alg[top] / stage_ms[top] / 1e6
    bound, peak_used, frac = "hbm", peak, achieved / peak
    if world > 1 and top == "feature":
        # T_roof = max((R_local + W_all) / BW_hbm, R_peer / BW_nvlink_in); fraction = T_roof / T_measured
        nvl = 770.0  # GB/s per direction per GPU: the measured peer copy B200_PROFILING.md states (our probe: 822-830)
        r_local, r_peer = info["src_local"] * row_bytes, info["src_peer"] * row_bytes
        t_hbm = (r_local + info["rows"] * (row_bytes + 4)) / (peak * 1e6)
        t_nvl = r_peer / (nvl * 1e6)
        t_roof = max(t_hbm, t_nvl)
        frac = t_roof / stage_ms[top]
        if t_nvl >= t_hbm:
            bound, peak_used, achieved = "nvlink", nvl, r_peer / stage_ms[top] / 1e6
        roof_note = (f"feature gather: {info['src_local'] / max(info['rows'], 1):.1%} of rows from local HBM, "
                     f"{info['src_peer'] / max(info['rows'], 1):.1%} from NVLink peers; roofline time "
                     f"max(HBM {t_hbm:.3f} ms, NVLink-in {t_nvl:.3f} ms) vs measured {stage_ms[top]:.3f} ms; "
                     f"peaks: HBM {peak:.0f} GB/s measured copy, NVLink {nvl:.0f} GB/s per direction")
    else:
        roof_note = f"{top}: all rows local HBM"
    traffic = None
    tfile = ROOT / "profiles" / "traffic.json"
    if tfile.exists():
        try:
            traffic = json.loads(tfile.read_text()).get(top.split(".")[0])
        except Exception:
            traffic = None
    value = R["edges"] / (R["ms"] / 1e3)
    e2e = R["e2e_edges"] / R["e2e_s"]
    steps_per_epoch = -(-(len(st.train) // world) // args.batch)
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": R["ms"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic", "config": workload_config(args, st.num_node, R["num_edge"], world),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": args.batch * 4 + 24, "d2h_bytes_per_step": 120,
                "ms_per_step": 1e3 * R["e2e_s"] / args.steps,
                "note": f"host seeds -> xg_sampler_submit (H2D inside) -> xg_batch_wait (D2H counts); {args.in_flight} batches in flight"},
        "gpu_launches": int(R["launches"]),
        "clocks": R["clk"],
        "roofline": {"bound": bound, "kernel": top, "achieved": round(achieved, 1), "peak": peak_used, "unit": "GB/s",
                     "frac": round(frac, 4), "traffic": traffic if world == 1 else None, "peak_source": peak_src, "note": roof_note,
                     "alg_bytes_per_launch": int(alg[top]), "launch_ms": round(stage_ms[top], 4)},
        "stages": stages,
        "epoch": {"steps_per_epoch": steps_per_epoch, "sample_extract_time_s": steps_per_epoch * R["ms"] / args.steps / 1e3,
                  "feature_gather_GBps": round(info["rows"] * row_bytes / stage_ms["feature"] / 1e6, 1),
                  "input_nodes_per_batch": info["rows"], "edges_per_batch": sum(info["edges"])},
        "build_s": round(R["t_build"], 1),
        "self_check": "one minibatch per rank before timing: feature rows == generator closed form (0 ULP), labels, seed prefix",
    }
    if R["rank"] == 0 and world == 1 and not args.no_cpu_baseline:
        del R["sm"]
        try:
            out["cpu_baseline"] = cpu_baseline(args, host_dataset(args), args.cpu_seconds)
        except Exception as e:  # never lose the GPU line to a host-side problem
            out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                                   "sample": f"failed: {type(e).__name__}: {e}"}
    if R["dist"] is not None:
        R["dist"].barrier()
    if R["rank"] == 0:
        print(json.dumps(out))
    if R["dist"] is not None:
        R["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
