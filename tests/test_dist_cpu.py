"""world_size-2/4 gloo runs on CPU of the host-side multi-rank logic: rank numbering, halo plan pairing,
message order (tests/multi_worker.py --cpu)."""
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def launch(nproc, extra, timeout=300, env_extra=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(HERE, "multi_worker.py")] + extra
    env = dict(os.environ, OMP_NUM_THREADS="2")
    env.update(env_extra or {})
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)


@pytest.mark.parametrize("nproc,lat,rg,nlinks", [
    (2, "8,8,8,16", "1,1,1,2", 4),
    (2, "8,8,8,16", "1,1,1,2", 8),
    (4, "8,8,16,16", "1,1,2,2", 8),
    (4, "16,8,8,8", "4,1,1,1", 4),
])
def test_halo_plan_pairs_up_gloo(nproc, lat, rg, nlinks):
    r = launch(nproc, ["--cpu", "--lat", lat, "--rankgeom", rg, "--nlinks", str(nlinks)])
    assert r.returncode == 0 and "MULTI_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("nproc,lat,rg,nlinks", [
    (2, "8,8,8,16", "1,1,1,2", 4),
    (2, "8,8,12,8", "1,1,2,1", 8),
    (4, "8,8,16,16", "1,1,2,2", 8),
    (4, "16,8,8,8", "4,1,1,1", 4),
])
def test_multi_process_sweep_over_gloo(nproc, lat, rg, nlinks):
    """world_size 2 and 4 on CPU, one process per rank: the site code of the CUDA sweep (csrc/sweep_site.cuh under g++)
    packs, torch.distributed (gloo) carries the halos with the library's plan, interior + boundary run on the received
    buffers, every rank checks its block of the oracle's Dslash (tests/multi_worker.py cpu_sweep_check)."""
    import test_sweep_emu
    test_sweep_emu.build_emu()             # build tests/emu/_emu_sweep.so once, before the ranks start
    r = launch(nproc, ["--cpu", "--cpu-sweep", "--lat", lat, "--rankgeom", rg, "--nlinks", str(nlinks)])
    assert r.returncode == 0 and "MULTI_OK" in r.stdout and "sweep_rel_l2" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
