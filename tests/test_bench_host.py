"""Host logic of bench.py that needs no GPU: the workloads of SURVEY 8d at 1/2/4/8 ranks, the config dict both arms
share (the driver compares them), and the reference arm's behaviour under a multi-rank launch."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


@pytest.mark.parametrize("cfg", ["c2", "c3", "c4", "c5"])
@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_workloads_decompose(cfg, world):
    lat, G, nlinks, scaling, wl = bench.workload(cfg, world)
    assert int(np.prod(G)) == world
    local = [lat[d] // G[d] for d in range(4)]
    assert all(lat[d] % G[d] == 0 and local[d] % 2 == 0 for d in range(4))
    if nlinks == 8:
        assert all(x >= 4 for x in local)                 # three-hop shifts (qexb200_gauge_create refuses less)
    if cfg == "c2":
        assert local == [32, 32, 32, 32] and scaling == "weak"     # weak series: 32^4 per GPU
    else:
        assert scaling == "strong" and lat == bench.workload(cfg, 1)[0]
    assert "x".join(map(str, lat)) in wl


def test_alt8_is_the_other_8_gpu_series():
    lat, G, _, _, _ = bench.workload("c2", 8, alt8=True)
    assert G == (1, 2, 2, 2) and lat == (32, 64, 64, 64)
    lat, G, _, _, _ = bench.workload("c2", 8)
    assert G == (1, 1, 2, 4) and lat == (32, 32, 64, 128)      # split the slowest dimensions


def test_config_dict_is_the_same_in_both_arms():
    """Nothing measured, nothing arm-specific: the driver's same_config check compares these dicts."""
    for cfg in ("c2", "c3", "c5"):
        for world in (1, 8):
            lat, G, nlinks, _, wl = bench.workload(cfg, world)
            a = bench.config_dict(wl, lat, G, nlinks)
            b = bench.config_dict(wl, lat, G, nlinks)
            assert a == b and set(a) == {"workload", "lattice", "rankgeom", "mass", "r2req", "l2"}
            json.dumps(a)


def test_reference_arm_ranks_other_than_zero_exit_quietly():
    """Under torchrun (N > 1) rank 0 alone runs the CPU arm; the others exit 0 without work or output."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_reference_arm_has_no_line_for_the_trajectory_config():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c4"],
                       capture_output=True, text=True, timeout=120)
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0 and line["impl"] == "reference" and "unavailable" in line


def test_failed_checks_turn_into_a_non_zero_exit():
    """rc 0 of bench.py must MEAN parity: a failed preflight, a missing preflight or a true residual above r2req each
    give a reason (run_ours prints the line, then exits 3 when the list is not empty)."""
    good = {"ok": True, "dslash_rel_l2": 3.7e-16, "cg_iters": [161, 161], "lattice": [8, 8, 8, 16], "rankgeom": [1, 1, 1, 2],
            "fat_naik": {"cg_iters": [160, 160]}}
    assert bench.verdict_of_checks(good, [8.8e-9, 9.98e-9]) == []
    assert bench.verdict_of_checks(None, [8.8e-9], no_parity=True) == []          # --no-parity is an explicit choice
    assert len(bench.verdict_of_checks(None, [8.8e-9])) == 1
    assert len(bench.verdict_of_checks(dict(good, ok=False, dslash_rel_l2=0.3), [8.8e-9])) == 1
    assert len(bench.verdict_of_checks(good, [8.8e-9, 1.2e-8])) == 1
    assert len(bench.verdict_of_checks(good, [float("nan")])) == 1
    s = bench.parity_summary(good)
    assert s["ok"] and s["cg_iters"] == [161, 161] and s["cg_iters_fat_naik"] == [160, 160] and s["rankgeom"] == [1, 1, 1, 2]
    assert bench.parity_summary(None) is None
