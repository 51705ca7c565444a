"""The reference-facing boundary: samgraph_* C-ABI (include/samgraph_abi.h) through xgnn_b200.samgraph.torch.

CPU part: shuffler order against an independent restatement of libstdc++'s engine + distribution, config
errors abort like the reference's CHECK, symbols.  GPU part: tests/samgraph_driver.py in its own process per
scenario (the engine is a process singleton), every batch compared with the CPU oracle."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import synth

ROOT = Path(__file__).resolve().parent.parent
DRIVER = str(ROOT / "tests" / "samgraph_driver.py")


@pytest.fixture(scope="module")
def dataset(tmp_path_factory):
    d = tmp_path_factory.mktemp("ds")
    V, E, dim = 3000, 24000, 20
    indptr, indices = synth.make_graph(V, E, seed=5, max_deg=24)
    train = synth.train_set(V, 0.2)[:501]  # odd: the 2-worker split pads with a wrapped-around duplicate
    synth.write_dataset(str(d), indptr, indices, synth.make_features(V, dim), synth.make_labels(V), train,
                        weights=synth.edge_weights(indptr, indices))
    return str(d)


# ------------------------------------------------------------------------------------------------ CPU
@pytest.mark.parametrize("n", [2, 17, 1000])
def test_aligned_shuffle_matches_libstdcxx_restatement(n):
    from xgnn_b200._clib import check, lib
    a = (np.arange(n, dtype=np.uint32) * 3 + 1)
    b = a.copy()
    for epoch in range(4):  # the array is re-shuffled in place epoch after epoch (dist_shuffler_aligned.cc:75-121)
        check(lib.xg_aligned_shuffle_epoch(b.ctypes.data, n, epoch))
        a = synth.minstd_shuffle_epoch(a, epoch)
        assert np.array_equal(a, b)
    assert sorted(b) == sorted(np.arange(n) * 3 + 1)


def _run_py(code, env=None):
    return subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(ROOT),
                          env={**os.environ, **(env or {})}, timeout=120)


def test_config_missing_key_aborts_like_reference_check():
    r = _run_py("import xgnn_b200.samgraph.torch as sam\nsam.config({'dataset_path': '/nope'})")
    assert r.returncode != 0 and "Check failed" in r.stderr and "missing config key" in r.stderr


def test_config_twice_and_out_of_scope_arch_abort(dataset):
    base = ("import xgnn_b200.samgraph.torch as sam\n"
            "cfg={'dataset_path': %r, '_arch': %%d, '_sample_type': 0, 'batch_size': 8, 'num_epoch': 1, '_cache_policy': 0,"
            "'cache_percentage': 0, 'max_sampling_jobs': 1, 'max_copying_jobs': 1, 'omp_thread_num': 1, 'num_layer': 1,"
            "'num_hidden': 1, 'lr': 0.1, 'dropout': 0.5, 'sampler_ctx': 'cuda:0', 'trainer_ctx': 'cuda:0',"
            "'fanout': [3], 'num_fanout': 1, 'num_sample_worker': 1, 'num_train_worker': 1}\n" % dataset)
    r = _run_py(base % 3 + "sam.config(cfg)\nsam.config(cfg)")
    assert r.returncode != 0 and "is_configured" in r.stderr
    r = _run_py(base % 9 + "sam.config(cfg)")  # arch9 = unified-memory sampler
    assert r.returncode != 0 and "out of scope" in r.stderr
    r = _run_py(base % 5 + "cfg['have_switcher'] = 1\nsam.config(cfg)")  # arch5 itself is built, its switcher is not
    assert r.returncode != 0 and "out of scope" in r.stderr
    r = _run_py(base % 5 + "sam.config(cfg)\nsam.data_init()\nprint(sam.steps_per_epoch(), sam.num_local_step())")
    assert r.returncode == 0, r.stderr[-2000:]
    r = _run_py(base % 3 + "sam.config(cfg)\nprint('configured')")
    assert r.returncode == 0 and "configured" in r.stdout


def test_data_init_loads_dataset_and_places_parts_without_cuda(dataset):
    """samgraph_data_init runs in the parent before fork: dataset + placement, no GPU needed"""
    code = ("import xgnn_b200.samgraph.torch as sam\n"
            "sam.config({'dataset_path': %r, '_arch': 6, '_sample_type': 0, 'batch_size': 50, 'num_epoch': 3, '_cache_policy': 0,"
            "'cache_percentage': 0.5, 'max_sampling_jobs': 1, 'max_copying_jobs': 1, 'omp_thread_num': 1, 'num_layer': 2,"
            "'num_hidden': 1, 'lr': 0.1, 'dropout': 0.5, 'num_worker': 4, 'fanout': [3, 4], 'num_fanout': 2,"
            "'use_dist_graph': 1.0, 'part_cache': 'True', 'gpu_extract': 'True'})\n"
            "sam.data_init()\n"
            "print(sam.num_class(), sam.feat_dim(), sam.num_epoch(), sam.num_local_step(), sam.steps_per_epoch())\n"
            "print(tuple(sam.get_dataset_feat().shape), int(sam.get_dataset_label()[46]))\n" % dataset)
    r = _run_py(code, env={"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stderr
    # 501 train ids over 4 workers -> 126 each -> 3 local steps of 50 -> 12 global steps (dist_engine.cc:139-148)
    assert r.stdout.split("\n")[0] == "47 20 3 3 12"
    assert r.stdout.split("\n")[1] == "(3000, 20) 46"


# ------------------------------------------------------------------------------------------------ GPU
def _drive(*args, timeout=300, env=None):
    r = subprocess.run([sys.executable, DRIVER, *map(str, args)], capture_output=True, text=True, cwd=str(ROOT),
                       timeout=timeout, env={**os.environ, **(env or {})})
    assert r.returncode == 0 and "DRIVER OK" in r.stdout, (r.stdout[-2000:], r.stderr[-4000:])


@pytest.mark.gpu
@pytest.mark.parametrize("sample_type", [0, 7, 1, 2, 4, 6, 3])
def test_single_process_flow_every_sampler(dataset, sample_type):
    _drive("single", dataset, sample_type)


@pytest.mark.gpu
def test_single_process_background_producer(dataset):
    _drive("single", dataset, 0, "bg")


@pytest.mark.gpu
@pytest.mark.parametrize("sample_type,workers", [(0, 2), (3, 2), (2, 2), (7, 3)])
def test_arch6_fork_per_worker_ggms_over_ipc(dataset, sample_type, workers):
    """data_init -> fork -> sample_init + train_init per worker; topology and feature shards are read across
    processes through CUDA IPC (both workers may share one GPU on a single-GPU box)"""
    import torch
    _drive("arch6", dataset, sample_type, workers, env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count()))})


@pytest.mark.gpu
@pytest.mark.parametrize("sample_type,workers,rep", [(0, 2, 0.2), (7, 3, 0.5)])
def test_arch6_two_tier_feature_store(dataset, sample_type, workers, rep):
    """Same arch6 flow with the hot tier replicated per worker (`ggms_replicate_percentage`, here through its
    environment twin): every batch's features must still equal CPUExtract's rows bit for bit."""
    import torch
    _drive("arch6", dataset, sample_type, workers,
           env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count())), "SAMGRAPH_GGMS_REPLICATE": str(rep)})


@pytest.mark.gpu
@pytest.mark.parametrize("mode,args", [("single", (0,)), ("arch6", (7, 2))])
def test_presample_cache_policy_flows(dataset, mode, args):
    """kCacheByPreSample: worker 0 ranks the nodes by presampled access frequency at sample_init (on the GPU), every
    trainer builds its cache from that ranking; batches must still carry CPUExtract's rows bit for bit."""
    import torch
    _drive(mode, dataset, *args, env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count())), "XG_TEST_PRESAMPLE": "1"})


@pytest.mark.gpu
@pytest.mark.parametrize("mode,args", [("single", (0,)), ("arch6", (0, 2))])
def test_sanity_check_env_flows(dataset, mode, args):
    """SAMGRAPH_SANITY_CHECK=1 (constant.cc:58; cuda_sanity_check.cu:28-88): every batch is checked on the device for
    kEmptyKey, out-of-range ids / edge endpoints and seeds repeated inside an epoch; a clean run must pass untouched."""
    import torch
    _drive(mode, dataset, *args, env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count())), "SAMGRAPH_SANITY_CHECK": "1"})


@pytest.mark.gpu
@pytest.mark.parametrize("workers,clique,alpha", [(2, 2, 0.0), (4, 2, 0.6), (3, 3, 0.3)])
def test_arch6_ics22_clique_placement(dataset, workers, clique, alpha):
    """`use_ics22_song_solver` (operation.cc:213-230): every worker ranks by in-degree on its GPU, solves its own view of
    the ICS'22 clique placement, shares its cache over IPC inside the clique and gathers through the compact
    (owner, row) map; every batch's features must still be CPUExtract's rows bit for bit."""
    import torch
    _drive("arch6", dataset, 0, workers,
           env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count())), "XG_TEST_ICS22": f"{clique},{alpha}"})


@pytest.mark.gpu
@pytest.mark.parametrize("mode,args,bits", [("single", (0,), 7), ("arch6", (0, 2), 5), ("single", (7,), 12)])
def test_empty_feat_mock_features(dataset, mode, args, bits):
    """SAMGRAPH_EMPTY_FEAT=k (run_config.cc:138, engine.cc:198-213): the feature table has 2^k rows and every read -
    cache build (CPUMockExtract), cache miss (extract_miss_data's mask), flat extract (GPUMockExtract) - masks the
    node id to it.  feat.bin is not opened (the reference's evaluation scripts use this to run shapes without one)."""
    import torch
    _drive(mode, dataset, *args,
           env={"XG_TEST_NUM_GPU": str(max(1, torch.cuda.device_count())), "SAMGRAPH_EMPTY_FEAT": str(bits)})
