"""GGMS placement solver (xg_partition_solve) against the reference solver's own output.

Goldens: tests/golden/solver_golden.json = what utility/nvlink_placement_solver/solver.cc (compiled as
is into oracle/_ref/solver.out) prints for the four topologies the reference ships (5/6/7/8 GPUs of a
DGX-1 hybrid cube mesh); the 8-GPU one equals the hand-written "perfect solution" in
cuda/dist_graph.cu:569-586.  Host-only code: runs without a GPU.
"""
import json
from pathlib import Path

import numpy as np
import pytest

from xgnn_b200._clib import XgError, check, lib

GOLD = json.loads((Path(__file__).parent / "golden" / "solver_golden.json").read_text())


def solve(bw: np.ndarray):
    n = bw.shape[0]
    bw = np.ascontiguousarray(bw, dtype=np.float64)
    nparts = np.zeros(n, np.int32)
    parts = np.zeros((n, n), np.int32)
    access = np.zeros((n, n), np.int32)
    check(lib.xg_partition_solve(bw.ctypes.data, n, nparts.ctypes.data, parts.ctypes.data, access.ctypes.data))
    return [list(map(int, parts[g, :nparts[g]])) for g in range(n)], access.tolist()


def matrix(case):
    n = case["num_gpu"]
    bw = np.zeros((n, n))
    for a, b, w in case["links"]:
        bw[int(a), int(b)] = w
        bw[int(b), int(a)] = w
    return bw


@pytest.mark.parametrize("n", ["5", "6", "7", "8"])
def test_reference_topologies_bit_exact(n):
    case = GOLD[n]
    parts, access = solve(matrix(case))
    for g, want in enumerate(case["placement"]):
        assert want["gpu"] == g
        assert parts[g] == want["part_ids"], (g, parts[g], want["part_ids"])
        assert access[g] == want["access"], (g, access[g], want["access"])


def test_8gpu_equals_hand_written_placement_in_reference_source():
    # cuda/dist_graph.cu:569-586
    want_parts = [[0, 4], [1, 5], [2, 6], [3, 7], [3, 4], [2, 5], [1, 6], [0, 7]]
    want_access = [[0, 6, 2, 3, 0, 1, 6, 3], [7, 1, 2, 3, 0, 1, 2, 7], [0, 1, 2, 3, 4, 1, 2, 3], [0, 1, 2, 3, 0, 5, 2, 3],
                   [7, 6, 5, 4, 4, 5, 2, 7], [7, 6, 5, 4, 4, 5, 6, 3], [0, 6, 5, 4, 0, 5, 6, 7], [7, 1, 5, 4, 4, 1, 6, 7]]
    parts, access = solve(matrix(GOLD["8"]))
    assert parts == want_parts and access == want_access


@pytest.mark.parametrize("n", [1, 2, 4, 8, 16])
def test_nvswitch_all_to_all_is_identity(n):
    bw = np.full((n, n), 770.0)
    np.fill_diagonal(bw, 6500.0)
    parts, access = solve(bw)
    assert parts == [[g] for g in range(n)]
    assert access == [list(range(n)) for _ in range(n)]


def test_every_gpu_reaches_every_part_on_random_topologies():
    rng = np.random.default_rng(3)
    for _ in range(20):
        n = int(rng.integers(3, 7))
        bw = np.zeros((n, n))
        for i in range(n):  # a ring guarantees a solution exists; extra random links
            bw[i, (i + 1) % n] = bw[(i + 1) % n, i] = 25.0
        for _ in range(n):
            i, j = rng.integers(0, n, 2)
            if i != j:
                bw[i, j] = bw[j, i] = float(rng.choice([25.0, 50.0]))
        np.fill_diagonal(bw, 700.0)
        parts, access = solve(bw)
        for g in range(n):
            for p in range(n):
                src = access[g][p]
                assert bw[g, src] != 0 and p in parts[src]
        assert all(g in parts[g] for g in range(n))


def test_islands_replicate_everything():
    """No P2P at all: each GPU can only host what it needs itself -> every part on every GPU."""
    bw = np.zeros((3, 3))
    np.fill_diagonal(bw, 700.0)
    parts, access = solve(bw)
    assert parts == [[0, 1, 2]] * 3
    assert access == [[g, g, g] for g in range(3)]


def test_bad_arguments():
    with pytest.raises(XgError):
        check(lib.xg_partition_solve(None, 2, None, None, None))
    bw = np.ones((17, 17))
    out = np.zeros((17, 17), np.int32)
    with pytest.raises(XgError):
        check(lib.xg_partition_solve(bw.ctypes.data, 17, out.ctypes.data, out.ctypes.data, out.ctypes.data))
