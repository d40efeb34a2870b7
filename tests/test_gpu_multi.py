"""Multi-GPU parity (one process per GPU, NCCL halo exchange): every rank checks its block of the
Dslash / CG / solve results against the oracle's global result.  Needs >= 2 GPUs on the box."""
import pytest

from test_dist_cpu import launch

pytestmark = pytest.mark.gpu


def only_nproc(nproc):
    """QEXB200_TEST_NPROC=8 runs just the 8-rank cases (GPU time on an N-GPU box is charged N-fold)."""
    import os
    want = os.environ.get("QEXB200_TEST_NPROC")
    if want and int(want) != nproc:
        pytest.skip("QEXB200_TEST_NPROC=%s" % want)


def run_case(nproc, lat, rg, inner, nlinks, transport, recon=18):
    only_nproc(nproc)
    if ngpus() < nproc:
        pytest.skip("needs %d GPUs" % nproc)
    r = launch(nproc, ["--lat", lat, "--rankgeom", rg, "--inner", inner, "--nlinks", str(nlinks),
                       "--transport", transport, "--recon", str(recon)], timeout=600)
    assert r.returncode == 0 and "MULTI_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def ngpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("nproc,lat,rg,inner,nlinks", [
    (2, "8,8,8,16", "1,1,1,2", "1,1,1,1", 4),
    (2, "8,8,8,16", "1,1,1,2", "1,2,2,2", 8),
    (2, "8,8,12,8", "1,1,2,1", "1,1,1,1", 8),       # local extent 6 in the split direction: 3-hop both ways
    (2, "8,8,8,8", "1,1,1,2", "1,1,1,1", 8),        # local extent 4 < 2*depth: a site hops off-rank both ways
    (4, "8,8,16,16", "1,1,2,2", "1,1,1,1", 8),
    (4, "16,8,8,8", "4,1,1,1", "1,1,1,1", 4),       # x0 split; neighbours are distinct ranks
    (8, "8,16,16,16", "1,2,2,2", "1,1,1,1", 8),
])
@pytest.mark.parametrize("transport", ["p2p", "nccl"])
def test_multi_gpu_parity(nproc, lat, rg, inner, nlinks, transport):
    """transport p2p: NVLink stores into the neighbours' windows + in-kernel cross-rank sums;
    nccl: ncclSend/ncclRecv + ncclAllReduce."""
    run_case(nproc, lat, rg, inner, nlinks, transport)


@pytest.mark.parametrize("nproc,lat,rg,nlinks,transport", [
    (2, "8,8,8,16", "1,1,1,2", 4, "p2p"),
    (2, "8,8,12,8", "1,1,2,1", 8, "p2p"),
    (2, "8,8,8,16", "1,1,1,2", 8, "nccl"),
    (4, "8,8,16,16", "1,1,2,2", 8, "p2p"),
])
def test_multi_gpu_parity_recon12(nproc, lat, rg, nlinks, transport):
    """Same checks with the 12-real link storage (sign bits of off-rank hops come from the packer)."""
    run_case(nproc, lat, rg, "1,1,1,1", nlinks, transport, recon=12)


@pytest.mark.parametrize("nproc,lat,rg,nlinks", [
    (2, "8,8,8,16", "1,1,1,2", 4),
    (2, "8,8,12,8", "1,1,2,1", 8),
    (4, "8,8,16,16", "1,1,2,2", 8),
    (8, "8,16,16,16", "1,2,2,2", 4),
])
def test_multi_gpu_parity_recon14(nproc, lat, rg, nlinks):
    """14-real storage (rows 0,1 + one complex factor per link) on a split lattice: U(3) links with a site-dependent
    determinant phase -- what the nHYP solves of BASELINE configs[3] read."""
    run_case(nproc, lat, rg, "1,1,1,1", nlinks, "p2p", recon=14)


@pytest.mark.parametrize("nproc,lat,rg,nlinks,transport", [
    (4, "8,16,16,8", "1,2,2,1", 4, "p2p"),          # x1 split: every fourth tile touches a face
    (4, "8,16,16,8", "1,2,2,1", 8, "p2p"),
    (8, "8,8,16,32", "1,1,2,4", 8, "p2p"),          # the rank geometry SURVEY 8e recommends for 8 GPUs
    (8, "8,8,16,32", "1,1,2,4", 4, "nccl"),
])
def test_multi_gpu_parity_more_geometries(nproc, lat, rg, nlinks, transport):
    run_case(nproc, lat, rg, "1,1,1,1", nlinks, transport)


@pytest.mark.parametrize("nproc,lat,rg,inner", [
    (2, "8,8,8,16", "1,1,1,2", "1,1,1,1"),
    (2, "8,8,8,8", "1,1,1,2", "1,2,2,1"),        # local extent 4: the 3-hop and the staples leave the rank both ways
    (2, "8,8,12,8", "1,1,2,1", "1,1,1,1"),
    (4, "8,8,16,16", "1,1,2,2", "1,1,1,1"),      # the lower staple's corner x - nu + mu crosses two ranks
    (4, "16,8,8,8", "4,1,1,1", "1,1,1,1"),
    (8, "8,16,16,16", "1,2,2,2", "1,1,1,1"),
])
def test_multi_gpu_links_and_force(nproc, lat, rg, inner):
    """nHYP / HISQ / fat7+Naik smearing, plaq, stagDeriv / stagD2deriv on a lattice split over ranks: link-shaped
    fields live in the IPC-mapped arena and off-rank neighbours are read over NVLink (csrc/smear.cu, force.cu)."""
    only_nproc(nproc)
    if ngpus() < nproc:
        pytest.skip("needs %d GPUs" % nproc)
    # the nHYP force keeps 52 fields and works in 28 more: a larger arena than the default 64
    r = launch(nproc, ["--lat", lat, "--rankgeom", rg, "--inner", inner, "--links"], timeout=600,
               env_extra={"QEXB200_ARENA_FIELDS": "128"})
    assert r.returncode == 0 and "MULTI_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
