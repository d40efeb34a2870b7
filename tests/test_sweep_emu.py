"""The multi-rank Dslash sweep on the CPU: the per-site functions of qex_b200/csrc/sweep_site.cuh -- the code the CUDA
kernels k_pack / k_sweep_mr wrap (pack_item, stencil_local, stencil_ghost, site_epilogue) -- compiled with g++ and run
over a lattice split into 2, 4 or 8 emulated ranks, with the library's own halo plan (qexb200_plan_halo) and block ->
tile table (qexb200_plan_tiles), against the oracle's global stagD2.  No GPU: this is the N > 1 path's index
arithmetic, buffer layout and ghost substitution under the driver's CPU suite (the multi-GPU pytest cases need 2-8
GPUs and are skipped on a 1-GPU box)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle
from util import make_inputs, rel_l2

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "emu", "_emu_sweep.so")


class EmuHalo(C.Structure):
    _fields_ = [("mu", C.c_int), ("depth", C.c_int), ("faceh", C.c_int), ("rank_plus", C.c_int), ("rank_minus", C.c_int),
                ("off_f", C.c_longlong), ("off_b", C.c_longlong)]


class EmuRecon(C.Structure):
    _fields_ = [("u12", C.POINTER(C.c_double) * 2), ("sgn", C.POINTER(C.c_ushort)), ("inv", C.c_double * 2)]


def build_emu():
    """g++ of the harness + the library's host-device header (rebuilt when either changed); returns the .so path."""
    src = os.path.join(HERE, "emu", "emu_sweep.cpp")
    hdrs = [os.path.join(HERE, "..", "qex_b200", "csrc", h) for h in ("sweep_site.cuh", "linkops.cuh")]
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(p) for p in [src] + hdrs):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", "-w",
                               "-I/usr/local/cuda/include", src, "-o", SO])
    return SO


@pytest.fixture(scope="module")
def emu():
    lib = C.CDLL(build_emu())
    dpp, ip = C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_int)
    dp_ = C.POINTER(C.c_double)
    lib.emu_sweep_mr.argtypes = [C.c_int, C.c_int, C.c_int, ip, C.c_int, C.c_int, dpp, C.POINTER(EmuRecon), dpp, dpp, C.c_double,
                                 C.c_double, C.c_double, C.c_double, C.POINTER(EmuHalo), C.c_int, dpp,
                                 C.POINTER(C.POINTER(C.c_uint)), C.c_int]
    lib.emu_sweep_nccl.argtypes = [C.c_int, C.c_int, C.c_int, ip, C.c_int, C.c_int, dpp, C.POINTER(EmuRecon), dpp, dpp, C.c_double,
                                   C.c_double, C.c_double, C.c_double, C.POINTER(EmuHalo), C.c_int, dpp, dpp]
    lib.emu_compress.argtypes = [C.c_int, ip, dp_, dp_, dp_, C.POINTER(C.c_ushort), C.c_int, C.c_int, dp_]
    lib.emu_site.argtypes = [C.c_int, C.c_int, ip, ip]
    return lib


def rank_coords(rank, G):
    rc = [0] * 4
    for d in (3, 2, 1, 0):          # dimension 0 slowest (qlayout.nim:94-113)
        rc[d] = rank % G[d]
        rank //= G[d]
    return rc


def block_index(L, G, rank):
    """Global oracle index of every local site of `rank`, in the device order (parity, idx)."""
    X = [L[d] // G[d] for d in range(4)]
    rc = rank_coords(rank, G)
    Vh = int(np.prod(X)) // 2
    idx = np.arange(Vh)
    xh = X[0] // 2
    out = []
    for par in (0, 1):
        x1 = (idx // xh) % X[1]; x2 = (idx // (xh * X[1])) % X[2]; x3 = idx // (xh * X[1] * X[2])
        x0 = 2 * (idx % xh) + ((par + x1 + x2 + x3) & 1)
        g = [x0 + rc[0] * X[0], x1 + rc[1] * X[1], x2 + rc[2] * X[2], x3 + rc[3] * X[3]]
        lex = ((g[3] * L[2] + g[2]) * L[1] + g[1]) * L[0] + g[0]
        out.append(((g[0] + g[1] + g[2] + g[3]) & 1) * (int(np.prod(L)) // 2) + lex // 2)
    return np.concatenate(out)


def vec_to_dev(v):      # [2 Vh, 3, 2] in (parity, idx) order -> double2[2][3][Vh]
    n = v.shape[0] // 2
    return np.ascontiguousarray(v.reshape(2, n, 3, 2).transpose(0, 2, 1, 3))


def vec_from_dev(d):
    n = d.shape[2]
    return np.ascontiguousarray(d.transpose(0, 2, 1, 3)).reshape(2 * n, 3, 2)


def links_to_dev(g):    # [nlinks, 2 Vh, 3, 3, 2] -> double2[nlinks][2][9][Vh]
    nl, n = g.shape[0], g.shape[1] // 2
    return np.ascontiguousarray(g.reshape(nl, 2, n, 9, 2).transpose(0, 1, 3, 2, 4))


def test_device_site_order_is_the_oracles(emu):
    """idx -> coordinates of the kernels (site_from_idx) = the oracle's site order, on a ragged lattice."""
    L = (4, 6, 8, 10)
    lat = oracle.Lattice(L)
    co = lat.coords
    X = (C.c_int * 4)(*L)
    out = (C.c_int * 5)()
    for i in np.random.default_rng(0).integers(0, lat.vol, 400):
        par, idx = divmod(int(i), lat.half)
        emu.emu_site(idx, par, X, out)
        assert tuple(out[:4]) == tuple(int(co[k][i]) for k in range(4))
    assert np.array_equal(block_index(L, (1, 1, 1, 1), 0), np.arange(lat.vol))


CASES = [
    ((8, 8, 8, 16), (1, 1, 1, 2), 4),
    ((8, 8, 8, 16), (1, 1, 1, 2), 8),
    ((8, 8, 12, 8), (1, 1, 2, 1), 8),       # local extent 6 in the split direction: three-hop both ways
    ((8, 8, 8, 8), (1, 1, 1, 2), 8),        # local extent 4 < 2 * depth: a site hops off-rank both ways
    ((8, 8, 16, 16), (1, 1, 2, 2), 8),      # corners
    ((16, 8, 8, 8), (4, 1, 1, 1), 4),       # x0 split (the checkerboarded direction); neighbours are distinct ranks
    ((16, 8, 8, 8), (2, 1, 1, 1), 8),
    ((8, 16, 16, 8), (1, 2, 2, 1), 4),      # x1 split: every fourth tile touches a face
    ((8, 8, 16, 32), (1, 1, 2, 4), 8),      # the 8-GPU rank grid of SURVEY 8e
    ((8, 16, 16, 16), (1, 2, 2, 2), 4),     # the other 8-GPU rank grid
    ((4, 6, 8, 20), (1, 1, 1, 2), 4),       # ragged local lattice 4.6.8.10: partial last tile
]


@pytest.mark.parametrize("L,G,nlinks", CASES)
def test_multi_rank_sweep_on_the_host_matches_the_oracle(emu, L, G, nlinks):
    lat, g, (x, r0) = make_inputs(L, nlinks)
    run_sweeps(emu, L, G, nlinks, lat, g, x, r0)


def compress(emu, dl, X, nlinks, modes):
    """compress_links (layout.cu) per rank on the host: returns (EmuRecon array, keep-alive list, rc bits) after the
    library's own acceptance test (deviation <= 2e-14, |s|^2 spread <= 1e-13 for the 12-real form)."""
    nl, Vh = nlinks // 4, int(np.prod(X)) // 2
    rec = (EmuRecon * len(dl))()
    keep, rcs = [], set()
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    for r, links in enumerate(dl):
        u12 = [np.zeros(4 * 2 * (7 if m == 14 else 6) * Vh * 2) if m else np.zeros(2) for m in (modes + [0])[:2]]
        sgn = np.zeros(2 * Vh, dtype=np.uint16)
        stats = np.zeros(6)
        assert emu.emu_compress(nl, (C.c_int * 4)(*X), dp(links), dp(u12[0]), dp(u12[1]),
                                sgn.ctypes.data_as(C.POINTER(C.c_ushort)), modes[0], (modes + [0])[1], dp(stats)) == 0
        rc = 0
        for j in range(nl):
            if not modes[j]:
                continue
            dev, smin, smax = stats[j], stats[2 + j], stats[4 + j]
            assert dev <= 2e-14 and smin > 0, (r, j, dev)
            if modes[j] == 12:
                assert smax - smin <= 1e-13 * smax
                rec[r].inv[j] = 1.0 / np.sqrt(0.5 * (smin + smax))
                rc |= 1 << j
            else:
                rec[r].inv[j] = 1.0
                rc |= 4 << j
            rec[r].u12[j] = dp(u12[j])
        rec[r].sgn = sgn.ctypes.data_as(C.POINTER(C.c_ushort))
        keep += [u12, sgn]
        rcs.add(rc)
    assert len(rcs) == 1
    return rec, keep, rcs.pop()


def run_sweeps(emu, L, G, nlinks, lat, g, x, r0, modes=None, transport="p2p"):
    import qex_b200 as qb
    world = int(np.prod(G))
    X = [L[d] // G[d] for d in range(4)]
    Vh = int(np.prod(X)) // 2
    blocks = [block_index(L, G, r) for r in range(world)]
    plans = [qb.plan_halo(L, G, r, nlinks) for r in range(world)]
    ndir = len(plans[0][0])
    cap = max(p[1] for p in plans)
    halo = (EmuHalo * (4 * world))()
    for r, (plan, _) in enumerate(plans):
        assert len(plan) == ndir
        for i, h in enumerate(plan):
            halo[4 * r + i] = EmuHalo(h["mu"], h["depth"], h["faceh"], h["rank_plus"], h["rank_minus"], h["off_f"], h["off_b"])
    comm_mask = sum(1 << d for d in range(4) if G[d] > 1)
    dl = [links_to_dev(g[:, b]) for b in blocks]
    dx = [vec_to_dev(x[b]) for b in blocks]
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    arr = lambda xs: (C.POINTER(C.c_double) * world)(*[dp(a) for a in xs])
    Xc = (C.c_int * 4)(*X)
    ntiles = (Vh + 127) // 128
    rec, keep, rc_bits = (None, None, 0) if modes is None else compress(emu, dl, X, nlinks, modes)
    for parity in (0, 1):
        tiles = []
        for r in range(world):
            t = qb.plan_tiles(L, G, r, nlinks, parity)
            tiles.append(np.array([ti | (need << 24) for ti, need in t], dtype=np.uint32))
            assert len(t) == ntiles
        tarr = (C.POINTER(C.c_uint) * world)(*[t.ctypes.data_as(C.POINTER(C.c_uint)) for t in tiles])
        for a, b, sgn, post in ((0.0, 0.0, 1.0, 1.0), (0.7, -0.3, 1.0, 1.0), (0.0, 0.04, -1.0, 1.0)):
            dr = [vec_to_dev(r0[bk]) for bk in blocks]
            recv = [np.full(2 * cap, np.nan) for _ in range(world)]      # anything read but never packed shows as NaN
            if transport == "p2p":
                rc = emu.emu_sweep_mr(world, nlinks // 4, rc_bits, Xc, comm_mask, parity, arr(dl), rec, arr(dx), arr(dr), a, b, sgn,
                                      post, halo, ndir, arr(recv), tarr, ntiles)
            else:
                send = [np.full(2 * cap, np.nan) for _ in range(world)]
                rc = emu.emu_sweep_nccl(world, nlinks // 4, rc_bits, Xc, comm_mask, parity, arr(dl), rec, arr(dx), arr(dr), a, b,
                                        sgn, post, halo, ndir, arr(send), arr(recv))
            assert rc == 0
            want = r0.copy()
            if sgn > 0:
                lat.stagD2(g, want, x, a, b, parity)
            else:       # the second sweep of stagD2ee: r = b x - hops(x) (stagD.nim:433-469 with 4 m^2 -> b)
                hop = np.zeros_like(r0)
                lat.stagD2(g, hop, x, 0.0, 0.0, parity)
                sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
                want[sl] = b * x[sl] - hop[sl]
            for r in range(world):
                got = vec_from_dev(dr[r])
                w = want[blocks[r]]
                sl = slice(0, Vh) if parity == 0 else slice(Vh, 2 * Vh)
                osl = slice(Vh, 2 * Vh) if parity == 0 else slice(0, Vh)
                assert np.isfinite(got).all()
                assert rel_l2(got[sl], w[sl]) <= 1e-13, (r, parity, a, b)
                assert np.array_equal(got[osl], r0[blocks[r]][osl])       # the other parity is untouched


@pytest.mark.parametrize("L,G,nlinks,modes", [
    ((8, 8, 8, 16), (1, 1, 1, 2), 4, [12]),
    ((8, 8, 12, 8), (1, 1, 2, 1), 8, [12, 12]),
    ((8, 8, 16, 16), (1, 1, 2, 2), 8, [0, 12]),      # fat 18, Naik 12: what a real HISQ set allows
    ((16, 8, 8, 8), (4, 1, 1, 1), 4, [12]),
    ((8, 8, 8, 16), (1, 1, 1, 2), 4, [14]),
    ((8, 8, 16, 16), (1, 1, 2, 2), 8, [14, 14]),
    ((8, 8, 16, 16), (1, 1, 2, 2), 8, [0, 14]),
    ((8, 16, 16, 8), (1, 2, 2, 1), 8, [12, 14]),
    ((8, 8, 8, 8), (1, 1, 1, 1), 8, [12, 12]),        # one rank: k_sweep's path (stencil_local only, local wrap of the sign bits)
])
def test_multi_rank_sweep_with_12_and_14_real_links(emu, L, G, nlinks, modes):
    """The same with the compressed link forms: k_compress / k_sign_bwd's site functions build rows 0,1 (+ sign bits or
    the factor phi) per rank, the sweep rebuilds the third row in `load_link_k` -- also in the pack items and for the
    sign bits of backward hops next to a rank boundary.  12 reals: links SU(3) to machine precision with phases and
    BC; 14 reals: U(3) links with a determinant phase that changes from site to site.  Bar: the 18-real oracle, 1e-13."""
    from util import make_exact_inputs
    lat, g, (x, r0) = make_exact_inputs(L, nlinks)
    if 14 in modes:
        th = np.random.default_rng(99).uniform(0, 2 * np.pi, g.shape[:2])
        gc = (g[..., 0] + 1j * g[..., 1]) * np.exp(1j * th)[:, :, None, None]
        if 12 in modes:       # keep the 12-real class SU(3): phases on the other class only
            keep = [k for k in range(nlinks) if modes[k % 2 if nlinks == 8 else 0] == 12]
            gc[keep] = (g[..., 0] + 1j * g[..., 1])[keep]
        g = np.ascontiguousarray(np.stack([gc.real, gc.imag], axis=-1))
    run_sweeps(emu, L, G, nlinks, lat, g, x, r0, modes)


@pytest.mark.parametrize("L,G,nlinks,modes", [
    ((8, 8, 8, 16), (1, 1, 1, 2), 4, None),
    ((8, 8, 12, 8), (1, 1, 2, 1), 8, None),
    ((8, 8, 8, 8), (1, 1, 1, 2), 8, None),           # local extent 4: region width = the whole extent
    ((8, 8, 16, 16), (1, 1, 2, 2), 8, None),         # corners belong to exactly one region
    ((16, 8, 8, 8), (4, 1, 1, 1), 4, None),
    ((8, 16, 16, 8), (1, 2, 2, 1), 8, None),
    ((8, 16, 16, 16), (1, 2, 2, 2), 8, None),        # three split directions: three regions
    ((8, 8, 16, 32), (1, 1, 2, 4), 4, None),
    ((8, 8, 16, 16), (1, 1, 2, 2), 8, [12, 12]),
    ((8, 8, 8, 16), (1, 1, 1, 2), 4, [14]),
])
def test_nccl_transport_phases_on_the_host(emu, L, G, nlinks, modes):
    """The second transport (pack -> ncclSend/ncclRecv -> interior k_sweep -> k_exterior, the 3-phase shape of
    stagD.nim:355-393) with its site functions on the host: pack_item into send buffers, the message pairing of
    comm.cu halo_start as copies, stencil_local with off-rank hops skipped, exterior_site over the disjoint boundary
    regions (plan_regions) -- against the oracle."""
    from util import make_exact_inputs
    lat, g, (x, r0) = (make_inputs if modes is None else make_exact_inputs)(L, nlinks)
    if modes and 14 in modes:
        th = np.random.default_rng(7).uniform(0, 2 * np.pi, g.shape[:2])
        gc = (g[..., 0] + 1j * g[..., 1]) * np.exp(1j * th)[:, :, None, None]
        g = np.ascontiguousarray(np.stack([gc.real, gc.imag], axis=-1))
    run_sweeps(emu, L, G, nlinks, lat, g, x, r0, modes, transport="nccl")


def _random_geometries(n, seed=20261105):
    import random
    rnd = random.Random(seed)
    grids = [(1, 1, 1, 2), (1, 1, 2, 1), (1, 2, 1, 1), (2, 1, 1, 1), (1, 1, 2, 2), (2, 2, 1, 1), (1, 2, 1, 2), (2, 1, 2, 1),
             (1, 1, 1, 4), (4, 1, 1, 1), (1, 4, 1, 1), (1, 1, 2, 4), (1, 2, 2, 2), (2, 2, 2, 1), (2, 1, 1, 4), (1, 1, 4, 2)]
    out = []
    while len(out) < n:
        G, nlinks = rnd.choice(grids), rnd.choice([4, 8])
        X = [rnd.choice([e for e in (2, 4, 6, 8) if e >= (4 if nlinks == 8 else 2)]) for _ in range(4)]
        vol = int(np.prod(X)) * int(np.prod(G))
        if 64 <= vol <= 20000:
            out.append((tuple(X[d] * G[d] for d in range(4)), G, nlinks, rnd.choice(["p2p", "nccl"])))
    return out


@pytest.mark.parametrize("L,G,nlinks,transport", _random_geometries(14))
def test_random_geometries(emu, L, G, nlinks, transport):
    """Seeded random lattices, rank grids (up to 4 ranks in one direction, 8 in all), link sets and transports: local
    extents down to 2 (naive) / 4 (fat+Naik, the library's own limit), every direction split in some case."""
    lat, g, (x, r0) = make_inputs(L, nlinks)
    run_sweeps(emu, L, G, nlinks, lat, g, x, r0, None, transport)
