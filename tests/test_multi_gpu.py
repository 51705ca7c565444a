"""Cross-GPU parity on REAL peers (skipped below two devices; run with `gpurun --gpus 2 -- pytest tests/test_multi_gpu.py`).

The driver's GPU test box has one GPU, where the arch6 / arch5 flows put every process on the same device: CUDA IPC is
exercised there but NVLink peer loads and the arch5 peer copy are not.  Here
  * operator level, one process, two devices: a feature store and a topology whose odd parts live on device 1 are
    gathered / sampled from device 0 through P2P loads inside the kernels - byte-exact against the oracle;
  * arch6: two workers on two devices (topology + feature shards read across NVLink through CUDA IPC);
  * arch5: sampler process(es) on cuda:0, trainer process(es) on cuda:1 - the packed Task crosses as one peer copy;
    includes the 2 samplers x 2 trainers case whose deadlock fix (stages released before every blocking dequeue)
    could not be re-run on two GPUs in round 1.
Every batch of the samgraph flows is compared with the CPU oracle bit for bit by tests/samgraph_driver.py.
"""
import numpy as np
import pytest

import oracle_lib as O
import synth
from test_samgraph_api import _drive, dataset  # noqa: F401  (module-scoped dataset fixture)

torch = pytest.importorskip("torch")
pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")]


def _on(a, device):
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(a).to(f"cuda:{device}")


def _peer_access():
    from xgnn_b200._clib import check, lib
    check(lib.xg_enable_peer_access(0, 1))
    check(lib.xg_enable_peer_access(1, 0))


@pytest.mark.parametrize("pct", [1.0, 0.4])
def test_feature_gather_reads_peer_hbm_inside_the_kernel(pct):
    from xgnn_b200 import ops
    _peer_access()
    torch.cuda.set_device(0)
    P, V, dim = 2, 50000, 128
    feat = synth.make_features(V, dim)
    indptr, _ = synth.make_graph(V, 400000, seed=9)
    rank = synth.degree_rank(indptr)
    ncached = int(V * pct)
    table = O.build_cache_table(rank, ncached, V)
    parts = [O.partition_features(feat, rank, ncached, p, P) for p in range(P)]
    d_parts = [_on(parts[0], 0), _on(parts[1], 1)]  # part 1 in the PEER's HBM
    h_feat = torch.from_numpy(feat).pin_memory()
    store = ops.make_feature_store(dim * 4, h_feat, _on(table, 0), d_parts, local_parts=(0,))
    nodes = np.random.default_rng(3).integers(0, V, size=200001).astype(np.uint32)
    stats = torch.zeros(4, dtype=torch.int64, device="cuda:0")
    out = ops.ExtractGGMS(store, _on(nodes, 0), dim, torch.float32, stats=stats)
    torch.cuda.synchronize()
    want = O.extract_ggms(feat, nodes, table, parts)
    assert out.cpu().numpy().tobytes() == want.tobytes()
    st = stats.cpu().numpy()
    assert st[1] > 0 and st[0] > 0 and (st[2] > 0) == (pct < 1.0), "local, peer (and host) tiers must all be live"


@pytest.mark.parametrize("sample_type", [0, 3])
def test_sampling_over_a_topology_partitioned_across_two_gpus(sample_type):
    """DeviceDistGraph routing (cuda/dist_graph.h:141-151) with part 1 on the peer: same edges as the flat CSR"""
    from xgnn_b200 import ops
    _peer_access()
    torch.cuda.set_device(0)
    V, E, P = 20000, 200000, 2
    indptr, indices = synth.make_graph(V, E, seed=12, max_deg=50)
    ncn = O.num_cache_node(indptr, 1.0)
    parts = [O.partition_csr(indptr, indices, ncn, p, P) for p in range(P)]
    keep = [(_on(pi, p), _on(px if len(px) else np.zeros(1, np.uint32), p)) for p, (pi, px) in enumerate(parts)]
    flat = ops.make_graph(_on(indptr, 0), _on(indices, 0))
    dist = ops.make_graph(part_indptr=[k[0] for k in keep], part_indices=[k[1] for k in keep], num_cache_node=ncn,
                          num_node=V, host_indptr=_on(indptr, 0), host_indices=_on(indices, 0))
    inp = _on(np.random.default_rng(1).permutation(V)[:5000].astype(np.uint32), 0)
    if sample_type == 0:
        a, b = ops.GPUSampleKHop(flat, inp, 7, seed=5, batch_key=2), ops.GPUSampleKHop(dist, inp, 7, seed=5, batch_key=2)
    else:
        a = ops.GPUSampleRandomWalk(flat, inp, 3, 0.5, 4, 5, seed=5, batch_key=2)
        b = ops.GPUSampleRandomWalk(dist, inp, 3, 0.5, 4, 5, seed=5, batch_key=2)
    torch.cuda.synchronize()
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    assert a[0].numel() > 0


@pytest.mark.parametrize("sample_type,workers", [(0, 2), (7, 2), (3, 2)])
def test_arch6_two_workers_on_two_devices(dataset, sample_type, workers):
    _drive("arch6", dataset, sample_type, workers, env={"XG_TEST_NUM_GPU": "2"})


def test_arch6_two_tier_store_on_two_devices(dataset):
    _drive("arch6", dataset, 0, 2, env={"XG_TEST_NUM_GPU": "2", "SAMGRAPH_GGMS_REPLICATE": "0.3"})


@pytest.mark.parametrize("sample_type,samplers,trainers", [(0, 1, 1), (0, 2, 2), (7, 2, 1), (3, 1, 2)])
def test_arch5_sampler_and_trainer_on_different_devices(dataset, sample_type, samplers, trainers):
    """with XG_TEST_NUM_GPU=2 and one sampler the driver puts the sampler on cuda:0 and the trainer on cuda:1; with two of
    each, sampler w sits on cuda:(w % 2) and trainer t on cuda:((2 + t) % 2): every Task still crosses processes, and
    half of them cross NVLink"""
    _drive("arch5", dataset, sample_type, samplers, trainers, env={"XG_TEST_NUM_GPU": "2", "SAMGRAPH_QUEUE_TIMEOUT": "60"})
