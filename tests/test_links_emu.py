"""Host emulation of the device link pipeline's per-site functions (qex_b200/csrc/linkops.cuh compiled
with g++) against the oracle: index arithmetic of the staple gather, 3x3 algebra, projectU.  No GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "emu", "_emu_links.so")


@pytest.fixture(scope="module")
def emu():
    src = os.path.join(HERE, "emu", "emu_links.cpp")
    hdr = os.path.join(HERE, "..", "qex_b200", "csrc", "linkops.cuh")
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++",
                               "-I/usr/local/cuda/include", src, "-o", SO])
    lib = C.CDLL(SO)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    lib.emu_staple_field.argtypes = [dp, dp, C.c_int, C.c_int, ip, dp]
    lib.emu_project_field.argtypes = [dp, ip, dp]
    lib.emu_hop.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, ip, ip]
    lib.emu_project_u_deriv_field.argtypes = [dp, dp, ip, dp]
    lib.emu_staple_deriv_field.argtypes = [dp, dp, dp, C.c_int, dp, C.c_int, dp, ip]
    lib.emu_project_tah_field.argtypes = [dp, ip, dp]
    lib.emu_staple_field_mr.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.c_int, C.POINTER(C.c_ubyte), C.c_longlong,
                                        C.c_longlong, C.c_int, C.c_int, ip, dp]
    lib.emu_naik_deriv_field.argtypes = [dp, dp, dp, C.c_int, C.c_double, ip]
    lib.emu_staple_deriv_pipe_field.argtypes = [dp, dp, dp, C.c_int, dp, C.c_int, dp, ip, C.c_int, C.c_int, C.c_int]
    return lib


def to_dev(f):
    """oracle field [vol,3,3,2] -> device layout [2][9][Vh][2]"""
    vol = f.shape[0]
    return np.ascontiguousarray(f.reshape(2, vol // 2, 9, 2).transpose(0, 2, 1, 3))


def from_dev(d):
    vh = d.shape[2]
    return np.ascontiguousarray(d.transpose(0, 2, 1, 3)).reshape(2 * vh, 3, 3, 2)


def dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


L = (4, 6, 4, 8)


def test_hop_matches_oracle_coords(emu):
    lat = oracle.Lattice(L)
    X = (C.c_int * 4)(*L)
    co = lat.coords
    vh = lat.half
    out = (C.c_int * 6)()
    rng = np.random.default_rng(1)
    for i in rng.integers(0, lat.vol, 300):
        par, idx = divmod(int(i), vh)
        for mu in range(4):
            for d in (+1, -1, +3, -3):
                emu.emu_hop(idx, par, mu, d, X, out)
                assert tuple(out[:4]) == tuple(int(co[k][i]) for k in range(4))
                y = [int(co[k][i]) for k in range(4)]
                y[mu] = (y[mu] + d) % L[mu]
                j = out[4] * vh + out[5]
                assert tuple(int(co[k][j]) for k in range(4)) == tuple(y)


def test_staple_pair_matches_makeImpLinks(emu):
    """threeStaple-only makeImpLinks = sum over nu of the staple pair (fat7l.nim:24-76,139)."""
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(987654321))
    want = lat.makeImpLinks(g, (0.0, 1.0, 0.0, 0.0, 0.0))
    X = (C.c_int * 4)(*L)
    gd = [to_dev(g[mu]) for mu in range(4)]
    for mu in range(4):
        acc = np.zeros_like(gd[0])
        for nu in range(4):
            if nu == mu:
                continue
            out = np.zeros_like(gd[0])
            emu.emu_staple_field(dp(gd[nu]), dp(gd[mu]), nu, mu, X, dp(out))
            acc += out
        got = from_dev(acc)
        err = np.linalg.norm(got - want[mu]) / np.linalg.norm(want[mu])
        assert err < 1e-14, (mu, err)


def test_project_u_matches_oracle(emu):
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(5))
    fat = lat.makeImpLinks(g, (0.625, -1.0 / 16, 1.0 / 64, -1.0 / 384, 0.0))
    want = lat.projectU(fat)
    X = (C.c_int * 4)(*L)
    for mu in range(4):
        out = np.zeros_like(to_dev(fat[mu]))
        emu.emu_project_field(dp(to_dev(fat[mu])), X, dp(out))
        got = from_dev(out)
        err = np.linalg.norm(got - want[mu]) / np.linalg.norm(want[mu])
        assert err < 1e-13, (mu, err)
        # and the result is unitary
        m = got[..., 0] + 1j * got[..., 1]
        dev = np.abs(np.einsum("xab,xcb->xac", m, m.conj()) - np.eye(3)).max()
        assert dev < 1e-12


def test_project_u_deriv_matches_oracle(emu):
    """projectUderiv (matrixFunctions.nim:321-357: rsqrt, inverse, sylsolve) on smeared, non-unitary links."""
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(5))
    x = lat.makeImpLinks(g, (0.5, 0.25 / 6, 0.0, 0.0, 0.0))[0]           # like (1-a) U + a/6 staples
    chain = np.random.default_rng(1).standard_normal(x.shape)
    want = lat.projectUderiv(x, chain)
    X = (C.c_int * 4)(*L)
    out = np.zeros_like(to_dev(x))
    emu.emu_project_u_deriv_field(dp(to_dev(x)), dp(to_dev(chain)), X, dp(out))
    err = np.linalg.norm(from_dev(out) - want) / np.linalg.norm(want)
    assert err < 1e-12, err


def test_staple_deriv_matches_oracle(emu):
    """symStapleDeriv (smearutil.nim:24-54) gathered in place, all direction pairs, accumulating."""
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(9))
    rng = np.random.default_rng(2)
    c = rng.standard_normal(g[0].shape)
    X = (C.c_int * 4)(*L)
    for d1 in range(4):
        for d2 in range(4):
            if d1 == d2:
                continue
            f1, f2 = rng.standard_normal(c.shape), rng.standard_normal(c.shape)
            w1, w2 = f1.copy(), f2.copy()
            lat.sym_staple_deriv(w1, w2, g[d1], d1, g[d2], d2, c)
            e1, e2 = to_dev(f1), to_dev(f2)
            emu.emu_staple_deriv_field(dp(e1), dp(e2), dp(to_dev(g[d1])), d1, dp(to_dev(g[d2])), d2, dp(to_dev(c)), X)
            assert np.linalg.norm(from_dev(e1) - w1) / np.linalg.norm(w1) < 1e-14, (d1, d2)
            assert np.linalg.norm(from_dev(e2) - w2) / np.linalg.norm(w2) < 1e-14, (d1, d2)


@pytest.mark.parametrize("late", [0, 1])
@pytest.mark.parametrize("nblocks,pb", [(2, 64), (6, 32), (64, 8)])
def test_staple_deriv_ring_pipeline_is_bit_identical(emu, nblocks, pb, late):
    """staple_deriv_pipe_thread (the body of k_staple_deriv_pipe) on a host ring, with requests delivered at once
    (late=0) and as late as its waits allow (late=1): slot reuse order, wait counts, the hand-over from a site to
    the thread's next site and the ragged last tile -- bit for bit the plain gather's result."""
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(9))
    rng = np.random.default_rng(2)
    c = rng.standard_normal(g[0].shape)
    X = (C.c_int * 4)(*L)
    for d1, d2 in ((0, 1), (1, 0), (2, 3), (3, 0), (1, 2)):
        f1, f2 = rng.standard_normal(c.shape), rng.standard_normal(c.shape)
        e1, e2, p1, p2 = to_dev(f1), to_dev(f2), to_dev(f1), to_dev(f2)
        G1, G2, Cc = to_dev(g[d1]), to_dev(g[d2]), to_dev(c)
        emu.emu_staple_deriv_field(dp(e1), dp(e2), dp(G1), d1, dp(G2), d2, dp(Cc), X)
        emu.emu_staple_deriv_pipe_field(dp(p1), dp(p2), dp(G1), d1, dp(G2), d2, dp(Cc), X, nblocks, pb, late)
        assert np.array_equal(p1, e1) and np.array_equal(p2, e2), (d1, d2)
        assert not np.array_equal(e1, to_dev(f1))


def test_project_tah(emu):
    rng = np.random.default_rng(3)
    lat = oracle.Lattice(L)
    x = rng.standard_normal((lat.vol, 3, 3, 2))
    out = np.zeros_like(to_dev(x))
    emu.emu_project_tah_field(dp(to_dev(x)), (C.c_int * 4)(*L), dp(out))
    m = x[..., 0] + 1j * x[..., 1]
    w = 0.5 * (m - m.conj().transpose(0, 2, 1))
    w = w - np.einsum("xaa->x", w)[:, None, None] / 3.0 * np.eye(3)
    got = from_dev(out)
    assert np.abs(got[..., 0] + 1j * got[..., 1] - w).max() < 1e-15


def test_naik_deriv_matches_oracle(emu):
    """Naik term of fat7lDeriv (fat7lderiv.nim:300-326): oracle with every fat coefficient zero."""
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(9))
    mid = np.random.default_rng(5).standard_normal(g.shape)
    d, ld = lat.gauge(4), lat.gauge(4)
    lat.fat7lDeriv(d, g, mid, (0.0, 0.0, 0.0, 0.0, 0.0), ld, g, mid, 0.3)
    X = (C.c_int * 4)(*L)
    for mu in range(4):
        out = np.zeros_like(to_dev(g[mu]))
        emu.emu_naik_deriv_field(dp(out), dp(to_dev(g[mu])), dp(to_dev(mid[mu])), mu, 0.3, X)
        assert np.linalg.norm(from_dev(out) - ld[mu]) / np.linalg.norm(ld[mu]) < 1e-14, mu


@pytest.mark.parametrize("G", [(1, 1, 1, 2), (1, 1, 2, 2), (2, 1, 1, 2)])
def test_split_lattice_staple_reads_the_right_peer(emu, qb, G):
    """The multi-rank path of the staple gather (m3_ldx<MR>, LSite.nb, qexb200_plan_peers) on the host: every rank's
    arena is a host buffer, off-rank neighbours (incl. the corner x - nu + mu two ranks away) must come from the
    owner's buffer.  Compared with the oracle's global 3-staple."""
    Lg = (8, 4, 8, 8)
    lat = oracle.Lattice(Lg)
    g = lat.random_gauge(lat.new_rng(21))
    want = lat.makeImpLinks(g, (0.0, 1.0, 0.0, 0.0, 0.0))
    lo = oracle.QexLayout(Lg, G, (1, 1, 1, 1))
    nr = int(np.prod(G))
    X = [Lg[d] // G[d] for d in range(4)]
    vloc = int(np.prod(X))
    # per-rank arenas: 4 fields in the device layout, back to back
    arenas = []
    for r in range(nr):
        fields = [to_dev(lo.scatter(g[mu])[r].reshape(vloc, 3, 3, 2)) for mu in range(4)]
        arenas.append(np.ascontiguousarray(np.stack(fields)))
    fbytes = arenas[0][0].nbytes
    ptrs = (C.c_char_p * nr)(*[C.cast(a.ctypes.data, C.c_char_p) for a in arenas])
    Xc = (C.c_int * 4)(*X)
    for r in range(nr):
        peer = (C.c_ubyte * 81)(*qb.plan_peers(Lg, G, r))
        for mu in range(4):
            acc = np.zeros_like(arenas[0][0])
            for nu in range(4):
                if nu == mu:
                    continue
                out = np.zeros_like(acc)
                emu.emu_staple_field_mr(ptrs, nr, r, peer, nu * fbytes, mu * fbytes, nu, mu, Xc, dp(out))
                acc += out
            mine = lo.scatter(want[mu])[r].reshape(vloc, 3, 3, 2)
            err = np.linalg.norm(from_dev(acc) - mine) / np.linalg.norm(mine)
            assert err < 1e-14, (G, r, mu, err)
