"""bench.py's roofline arithmetic on known numbers - no GPU: the split roofline of a gather with local, peer and
host rows, the all-pull NVLink variant, and the whole-step fraction.  Inputs are the r02p 8-GPU line's own counts."""
import argparse
import types

import pytest

import bench


def _run(world, info, stage_ms, steps_ms):
    st = types.SimpleNamespace(row_bytes=512)
    R = {"st": st, "world": world, "ms": steps_ms * 10}
    args = argparse.Namespace(steps=10)
    return bench.summarize(R, args, stage_ms, info)


def _info(rows, local, peer, host):
    L = len(bench.FANOUT)
    return {"edges": [4_600_000, 800_000, 100_000][:L], "inputs": [900_000, 100_000, 8000][:L], "rows": rows, "seeds": 8000,
            "src_local": local, "src_peer": peer, "src_host": host}


def _stages(feature_ms):
    s = {"reset_seeds": 0.04, "input_nodes": 0.006, "feature": feature_ms, "label": 0.008}
    for l in range(len(bench.FANOUT)):
        s[f"sample.L{l}"], s[f"insert.L{l}"], s[f"number.L{l}"] = 0.1, 0.1, 0.09
    return s


def test_all_local_gather_is_measured_against_the_hbm_copy_peak():
    peak, _ = bench.measured_peaks()
    rows = 4_918_000
    roof, stages = _run(1, _info(rows, rows, 0, 0), _stages(0.846), 1.153)
    want = rows * (2 * 512 + 4) / 0.846 / 1e6
    assert roof["bound"] == "hbm" and roof["kernel"] == "feature"
    assert roof["achieved"] == pytest.approx(want, rel=1e-3) and roof["frac"] == pytest.approx(want / peak, abs=2e-4)
    assert "all_gpus_pull" not in roof and "row_sources" not in roof
    assert roof["step"]["ms"] == pytest.approx(1.153) and 0 < roof["step"]["frac_of_hbm_peak"] < roof["frac"]
    assert stages["feature"]["alg_GBps"] == pytest.approx(want, rel=1e-3)


def test_split_roofline_with_peer_rows_and_the_all_pull_variant():
    peak, _ = bench.measured_peaks()
    rows = 4_900_000
    peer = int(rows * 0.206)
    roof, _ = _run(8, _info(rows, rows - peer, peer, 0), _stages(0.997), 1.306)
    t_hbm = ((rows - peer) * 512 + rows * 516) / (peak * 1e6)
    t_nvl = peer * 512 / (bench.NVLINK_GBS * 1e6)
    assert roof["bound"] == ("hbm" if t_hbm >= t_nvl else "nvlink")
    assert roof["frac"] == pytest.approx(max(t_hbm, t_nvl) / 0.997, abs=2e-4)
    assert roof["row_sources"]["peer"] == pytest.approx(0.206, abs=1e-3)
    ap = roof["all_gpus_pull"]  # same max() with the NVLink term at the measured all-pull rate
    t_all = max(t_hbm, peer * 512 / (bench.NVLINK_ALL_PULL_GBS * 1e6))
    assert ap["roofline_ms"] == pytest.approx(t_all, abs=1e-4) and ap["frac"] == pytest.approx(t_all / 0.997, abs=2e-4)
    assert ap["frac"] > roof["frac"]  # a slower link is a closer roofline


def test_host_rows_make_the_gather_pcie_bound():
    rows = 4_000_000
    host = int(rows * 0.56)
    roof, _ = _run(1, _info(rows, rows - host, 0, host), _stages(22.0), 23.0)
    assert roof["bound"] == "pcie" and roof["peak"] == bench.PCIE_GBS
    assert roof["frac"] == pytest.approx(host * 512 / (bench.PCIE_GBS * 1e6) / 22.0, abs=2e-4)
    assert "all_gpus_pull" not in roof  # one GPU, no peer rows
