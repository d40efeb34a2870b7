"""GPU parity tests proper: the CUDA path, called through the C ABI with native QEX host memory,
against the CPU oracle on the same seeded inputs.  Bars: Dslash rel-L2 <= 1e-13,
CG iterations within +-1 and true residual <= requested."""
import numpy as np
import pytest

from util import make_inputs, make_exact_inputs, rel_l2

pytestmark = pytest.mark.gpu

DSLASH_TOL = 1e-13


@pytest.fixture()
def ctx_factory(qb):
    made = []

    def make(L, inner=(1, 1, 1, 1)):
        if qb.is_initialized():
            qb.finalize()
        qb.init(L)
        qb.set_host_layout(inner)
        made.append(1)
        return qb
    yield make
    if qb.is_initialized():
        qb.finalize()


CASES = [
    ((8, 8, 8, 8), 4, (1, 1, 1, 1)),
    ((8, 8, 8, 8), 4, (1, 2, 2, 2)),      # QEX's default V=8 lane geometry on 8^4
    ((8, 8, 8, 8), 8, (1, 2, 2, 2)),
    ((8, 8, 8, 16), 4, (1, 1, 1, 1)),     # BASELINE config shape 8^3 x 16
    ((4, 6, 8, 10), 8, (1, 1, 1, 1)),     # ragged extents; Naik wraps more than once in x0 (L=4)
    ((12, 4, 4, 4), 4, (2, 1, 1, 1)),     # odd outer extent: lane checkerboard twist in the host layout
    ((16, 16, 16, 16), 8, (1, 2, 2, 2)),
]


@pytest.mark.parametrize("L,nlinks,inner", CASES)
def test_stagD2_parity(orc, ctx_factory, L, nlinks, inner):
    """stagD2 (stagD.nim:348-395) on both parities, with a, b != 0."""
    qb = ctx_factory(L, inner)
    lat, g, (x, r0) = make_inputs(L, nlinks)
    lo = orc.QexLayout(L, (1, 1, 1, 1), inner)
    s = qb.Staggered([lo.scatter(g[k])[0] for k in range(nlinks)])
    for parity in (0, 1):
        for a, b in ((0.0, 0.0), (0.7, -0.3)):
            want = r0.copy()
            lat.stagD2(g, want, x, a, b, parity)
            got = lo.scatter(r0)[0]
            qb.stagD2(s, parity, got, lo.scatter(x)[0], a, b)
            got = lo.gather([got])
            sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
            osl = slice(lat.half, lat.vol) if parity == 0 else slice(0, lat.half)
            assert rel_l2(got[sl], want[sl]) <= DSLASH_TOL
            assert np.array_equal(got[osl], r0[osl])      # the other parity is untouched
    s.free()


@pytest.mark.parametrize("L,nlinks,inner", CASES[:5])
def test_D_Ddag_D2ee_parity(orc, ctx_factory, L, nlinks, inner):
    qb = ctx_factory(L, inner)
    lat, g, (x, _) = make_inputs(L, nlinks)
    lo = orc.QexLayout(L, (1, 1, 1, 1), inner)
    s = qb.Staggered([lo.scatter(g[k])[0] for k in range(nlinks)])
    xq = lo.scatter(x)[0]
    m = 0.1
    for name in ("D", "Ddag"):
        want = lat.vector()
        getattr(lat, name)(g, want, x, m)
        got = np.zeros_like(xq)
        getattr(s, name)(got, xq, m)
        assert rel_l2(lo.gather([got]), want) <= DSLASH_TOL
    for parity, fn in ((0, qb.stagD2ee), (1, qb.stagD2oo)):
        want = lat.vector()
        lat.stagD2xx(g, want, x, m * m, parity)
        got = np.zeros_like(xq)
        fn(s, got, xq, m * m)
        sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
        assert rel_l2(lo.gather([got])[sl], want[sl]) <= DSLASH_TOL
    # eoReconstruct (stagD.nim:582-586)
    want = x.copy()
    b = make_inputs(L, 4, seed=11, nvec=1)[2][0]
    lat.eoReconstruct(g, want, b, m)
    got = xq.copy()
    s.eoReconstruct(got, lo.scatter(b)[0], m)
    assert rel_l2(lo.gather([got]), want) <= DSLASH_TOL
    s.free()


@pytest.mark.parametrize("L,nlinks,inner", [CASES[0], CASES[2], CASES[4], CASES[6]])
def test_recon12_parity(orc, ctx_factory, L, nlinks, inner):
    """12-real link storage (rows 0,1 + rebuilt third row, include/qex_b200.h qexb200_gauge_recon):
    phased +-SU(3) one-hop links and 0.1*SU(3) three-hop links qualify, the Dslash stays within the
    1e-13 bar of the 18-real oracle and the CG iteration count within +-1."""
    qb = ctx_factory(L, inner)
    lat, g, (x, r0) = make_exact_inputs(L, nlinks)
    lo = orc.QexLayout(L, (1, 1, 1, 1), inner)
    s = qb.Staggered([lo.scatter(g[k])[0] for k in range(nlinks)])
    got_rc, dev = s.recon(12)
    assert got_rc == ([12, 12] if nlinks == 8 else [12, 18]), (got_rc, dev)
    assert max(dev[:nlinks // 4]) <= 2e-14
    for parity in (0, 1):
        want = r0.copy()
        lat.stagD2(g, want, x, 0.7, -0.3, parity)
        got = lo.scatter(r0)[0]
        qb.stagD2(s, parity, got, lo.scatter(x)[0], 0.7, -0.3)
        got = lo.gather([got])
        sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
        assert rel_l2(got[sl], want[sl]) <= DSLASH_TOL
    m, r2req = 0.1, 1e-10
    want = lat.vector()
    it_o, _ = lat.solveXX(g, want, x, m, 0, r2req, 10000)
    got = lo.scatter(lat.vector())[0]
    sp = qb.SolverParams(r2req=r2req, maxits=10000)
    s.solveEE(got, lo.scatter(x)[0], m, sp)
    assert abs(sp.iterations - it_o) <= 1, (sp.iterations, it_o)
    assert rel_l2(lo.gather([got])[:lat.half], want[:lat.half]) <= 10 * np.sqrt(r2req)
    # back to 18 reals: bit-identical to a set that never was compressed
    assert s.recon(18)[0] == [18, 18]
    a = lo.scatter(r0)[0]
    qb.stagD2(s, 0, a, lo.scatter(x)[0], 0.0, 0.0)
    s2 = qb.Staggered([lo.scatter(g[k])[0] for k in range(nlinks)])
    b = lo.scatter(r0)[0]
    qb.stagD2(s2, 0, b, lo.scatter(x)[0], 0.0, 0.0)
    assert np.array_equal(a, b)
    s.free()
    s2.free()


def test_recon12_refuses_non_unitary_links(orc, ctx_factory):
    """Smeared (non-unitary) fat links must stay at 18 reals while the Naik class compresses; device
    rephase keeps the request (sticky) and re-derives the sign bits."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    # QEX's own projectSU output is unitary to ~1e-11 only: above the 2e-14 bar, so it is refused
    # (reading it as 12 reals would cost the 1e-13 parity with the 18-real reference)
    lat, g, (x, _) = make_inputs(L, 4)
    s = qb.Staggered([g[k] for k in range(4)])
    got_rc, dev = s.recon(12)
    assert got_rc == [18, 18] and 2e-14 < dev[0] < 1e-8, (got_rc, dev)
    s.free()
    lat, g, (x, _) = make_exact_inputs(L, 8)
    rs = np.random.default_rng(5)
    g = g.copy()
    g[0::2] += 0.05 * rs.standard_normal(g[0::2].shape)     # "fat" links: no longer s*SU(3)
    s = qb.Staggered([g[k] for k in range(8)])
    got_rc, dev = s.recon(12)
    assert got_rc == [18, 12] and dev[0] > 1e-3
    assert s.recon(1218)[0] == [18, 18] and s.recon(1812)[0] == [18, 12]
    want, got = lat.vector(), lat.vector()
    lat.D(g, want, x, 0.1)
    s.D(got, x, 0.1)
    assert rel_l2(got, want) <= DSLASH_TOL
    s.free()
    # sticky across rephase: plain SU(3) in, phases applied on the device
    from qex_b200.synth import random_su3_links
    g0 = np.ascontiguousarray(np.stack(random_su3_links(lat.vol, 4, seed=9)))
    g1 = g0.copy()
    lat.setBC(g1)
    lat.stagPhase(g1)
    s = qb.newStag([g0[k] for k in range(4)])
    assert s.recon(12)[0][0] == 12
    s.rephase()
    lat.D(g1, want, x, 0.0)
    s.D(got, x, 0.0)
    assert rel_l2(got, want) <= DSLASH_TOL
    assert s.recon(12)[0][0] == 12
    s.free()


def test_point_source_and_ramp(orc, ctx_factory):
    """The reference's own smoke inputs: point source v{0}[0]=1 (testStagProp.nim:37) and the
    coordinate-ramp source on unit gauge with BC + phases (stagD.nim:703-714)."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat = orc.Lattice(L)
    g = lat.unit_gauge()
    lat.setBC(g)
    lat.stagPhase(g)
    s = qb.newStag([g[k] for k in range(4)])
    v = lat.vector()
    v[0, 0, 0] = 1.0
    ramp = lat.vector()
    c = lat.coords.astype(float)
    ramp[:, 0, 0] = c[0] * 1000 + c[1] * 100 + c[2] * 10 + c[3]
    for src in (v, ramp):
        want, got = lat.vector(), lat.vector()
        lat.D(g, want, src, 0.5)
        s.D(got, src, 0.5)
        assert np.isfinite(got).all()
        assert rel_l2(got, want) <= DSLASH_TOL
    s.free()


def test_device_rephase_matches_oracle(orc, ctx_factory):
    """Staggered.rephase on the device copy == setBC + stagPhase of the oracle (stagD.nim:72-80)."""
    L = (4, 6, 8, 10)
    qb = ctx_factory(L)
    lat = orc.Lattice(L)
    rng = lat.new_rng(3)
    g0 = lat.random_gauge(rng)
    x = lat.gaussian_vector(rng)
    g1 = g0.copy()
    lat.setBC(g1)
    lat.stagPhase(g1)
    s = qb.newStag([g0[k] for k in range(4)])
    s.rephase()
    want, got = lat.vector(), lat.vector()
    lat.D(g1, want, x, 0.0)
    s.D(got, x, 0.0)
    assert rel_l2(got, want) <= DSLASH_TOL
    s.free()


@pytest.mark.parametrize("L,nlinks,m,r2req", [
    ((8, 8, 8, 8), 4, 0.1, 1e-8),        # BASELINE config 1
    ((8, 8, 8, 8), 8, 0.05, 1e-12),
    ((8, 8, 8, 16), 4, 0.05, 1e-12),
    ((16, 16, 16, 16), 4, 0.1, 1e-8),
])
def test_solveEE_parity(orc, ctx_factory, L, nlinks, m, r2req):
    """solveEE / solveOO (stagSolve.nim:57-138): iteration count within +-1 of the oracle's CG,
    same recursive residual, solution close, true residual below the request."""
    qb = ctx_factory(L)
    lat, g, (b, _) = make_inputs(L, nlinks)
    s = qb.Staggered([g[k] for k in range(nlinks)])
    for par_even in (True, False):
        parity = 0 if par_even else 1
        want = lat.vector()
        it_o, r2_o = lat.solveXX(g, want, b, m, parity, r2req, 10000)
        got = np.ones_like(b)      # solveXX must zero r itself (stagSolve.nim:63-64)
        sp = qb.SolverParams(r2req=r2req, maxits=10000)
        s.solveXX(got, b, m, sp, parEven=par_even)
        assert abs(sp.iterations - it_o) <= 1, (sp.iterations, it_o)
        assert sp.r2 <= r2req
        sl = slice(0, lat.half) if par_even else slice(lat.half, lat.vol)
        osl = slice(lat.half, lat.vol) if par_even else slice(0, lat.half)
        assert np.all(got[osl] == 0)
        assert rel_l2(got[sl], want[sl]) <= 10 * np.sqrt(r2req)
        # true residual of the normal equations
        Ax = lat.vector()
        lat.stagD2xx(g, Ax, got, m * m, parity)
        tr = float(((Ax - b)[sl] ** 2).sum() / (b[sl] ** 2).sum())
        assert tr <= 1.01 * r2req
        nev = lat.half
        assert sp.flops == (nlinks * 4 * 72 + 60) * nev * sp.iterations   # stagSolve.nim:92
    s.free()


@pytest.mark.parametrize("L,nlinks,m,r2req", [
    ((8, 8, 8, 8), 4, 0.1, 1e-8),
    ((8, 8, 8, 8), 8, 0.01, 1e-16),
    ((8, 8, 8, 8), 8, 0.05, 1e-12),
    ((4, 6, 8, 10), 4, 0.2, 1e-20),
])
def test_solve_parity(orc, ctx_factory, L, nlinks, m, r2req):
    """Staggered.solve (stagSolve.nim:224-294): D x = b on the full lattice."""
    qb = ctx_factory(L)
    lat, g, (b, _) = make_inputs(L, nlinks)
    s = qb.Staggered([g[k] for k in range(nlinks)])
    want = lat.vector()
    st = lat.solve(g, want, b, m, r2req, 20000)
    got = lat.vector()
    sp = qb.SolverParams(r2req=r2req, maxits=20000)
    s.solve(got, b, m, sp)
    # +-1 per inner solve for solves of a few hundred iterations; a 2500-iteration solve at m=0.01
    # amplifies summation-order rounding (QEX itself differs run to run across thread counts,
    # SURVEY.md 7 "parity of CG iteration counts"), and landing a hair above r2req triggers one more
    # defect-correction pass (stagSolve.nim:259) -- allow 3 % there.
    slack = max(1, st.calls) if st.iterations < 500 else int(0.03 * st.iterations) + 2
    assert abs(sp.iterations - st.iterations) <= slack, (sp.iterations, st.iterations)
    Dx = lat.vector()
    lat.D(g, Dx, got, m)
    tr = lat.norm2(Dx - b) / lat.norm2(b)
    assert tr <= r2req
    assert abs(sp.r2 - tr) <= 1e-3 * tr + 1e-30
    assert rel_l2(got, want) <= 10 * np.sqrt(r2req)
    s.free()


@pytest.mark.parametrize("L,nlinks,ms,r2req", [
    ((8, 8, 8, 8), 4, (0.1, 0.2, 0.5), 1e-12),
    ((8, 8, 8, 16), 8, (0.05, 0.1, 0.3, 1.0), 1e-10),
    ((4, 6, 8, 10), 4, (0.2,), 1e-16),                 # a single mass: plain CG through the multi-shift code
])
def test_multishift_solveXX_parity(orc, ctx_factory, L, nlinks, ms, r2req):
    """solveXX(s, xs, b, ms, sp, subset) (stagSolve.nim:296-345, cgm.nim:84-316): same iteration count
    (+-1) and solutions as the oracle's multi-shift CG, on both subsets; every xs[k] satisfies its own
    shifted system; the untouched parity of xs is preserved."""
    qb = ctx_factory(L)
    lat, g, (b, junk) = make_inputs(L, nlinks)
    s = qb.Staggered([g[k] for k in range(nlinks)])
    shifts = [ms[0]] + [4.0 * (m * m - ms[0] ** 2) for m in ms[1:]]
    for par_even in (True, False):
        parity = 0 if par_even else 1
        sl = slice(0, lat.half) if par_even else slice(lat.half, lat.vol)
        osl = slice(lat.half, lat.vol) if par_even else slice(0, lat.half)
        want = np.zeros((len(ms),) + b.shape)
        it_o, r2_o, b2_o = lat.cgm_xx(g, want, b, ms[0] ** 2, shifts, parity, r2req, 10000)
        got = [junk.copy() for _ in ms]
        sp = qb.SolverParams(r2req=r2req, maxits=10000)
        s.solveXX_multi(got, b, shifts, sp, parEven=par_even)
        assert abs(sp.iterations - it_o) <= 1, (sp.iterations, it_o)
        assert sp.flops == (nlinks * 4 * 72 + 60) * lat.half * sp.iterations
        for k, m in enumerate(ms):
            assert np.array_equal(got[k][osl], junk[osl])
            assert rel_l2(got[k][sl], want[k][sl]) <= 10 * np.sqrt(r2req)
            Ax = lat.vector()
            lat.stagD2xx(g, Ax, got[k], m * m, parity)
            tr = float(((Ax - b)[sl] ** 2).sum() / (b[sl] ** 2).sum())
            assert tr <= 1.01 * r2req, (k, tr)
    # maxits is honoured, and every shifted solution still gets that iteration's update
    got = [lat.vector() for _ in ms]
    want = np.zeros((len(ms),) + b.shape)
    lat.cgm_xx(g, want, b, ms[0] ** 2, shifts, 0, 1e-30, 5)
    sp = qb.SolverParams(r2req=1e-30, maxits=5)
    s.solveXX_multi(got, b, shifts, sp)
    assert sp.iterations == 5
    for k in range(len(ms)):
        assert rel_l2(got[k][:lat.half], want[k][:lat.half]) <= 1e-11
    s.free()


@pytest.mark.parametrize("L,nlinks,ms,r2req", [
    ((8, 8, 8, 8), 4, (0.1, 0.2, 0.5), 1e-12),
    ((8, 8, 8, 8), 8, (0.05, 0.25), 1e-10),
])
def test_multishift_solve_parity(orc, ctx_factory, L, nlinks, ms, r2req):
    """solve(s, xs, b, ms, sp) (stagSolve.nim:347-446) and the reference's own check of it
    (stagSolve.nim:651-661): multi-shift result == repeated single-mass solves."""
    qb = ctx_factory(L)
    lat, g, (b, _) = make_inputs(L, nlinks)
    s = qb.Staggered([g[k] for k in range(nlinks)])
    want = np.zeros((len(ms),) + b.shape)
    st = lat.solve_multi(g, want, b, ms, r2req, 20000)
    got = [lat.vector() for _ in ms]
    sp = qb.SolverParams(r2req=r2req, maxits=20000)
    s.solve_multi(got, b, ms, sp)
    assert abs(sp.iterations - st.iterations) <= max(1, st.calls), (sp.iterations, st.iterations)
    assert sp.r2 <= r2req
    for k, m in enumerate(ms):
        assert rel_l2(got[k], want[k]) <= 10 * np.sqrt(r2req)
        Dx = lat.vector()
        lat.D(g, Dx, got[k], m)
        assert lat.norm2(Dx - b) / lat.norm2(b) <= 2 * r2req
        single = lat.vector()
        s.solve(single, b, m, qb.SolverParams(r2req=r2req, maxits=20000))
        assert rel_l2(got[k], single) <= 20 * np.sqrt(r2req)
    # device-resident entry point gives the same numbers
    fb = qb.DeviceField(b)
    fxs = [qb.DeviceField() for _ in ms]
    sp2 = qb.SolverParams(r2req=r2req, maxits=20000)
    s.dev_solve_multi(fxs, fb, ms, sp2)
    assert sp2.iterations == sp.iterations
    for k in range(len(ms)):
        assert np.array_equal(fxs[k].download(lat.vector()), got[k])
        fxs[k].free()
    fb.free()
    s.free()


def test_solve_structural_identities(orc, ctx_factory):
    """solve(D x0) == x0 (stagSolve.nim:556-574); usePrevSoln restart converges in 0 iterations;
    zero source gives zero (cg.nim:139-144); maxits is honoured."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g, (x0, _) = make_inputs(L, 4)
    s = qb.Staggered([g[k] for k in range(4)])
    m = 0.1
    b = lat.vector()
    lat.D(g, b, x0, m)
    x = lat.vector()
    sp = qb.SolverParams(r2req=1e-22, maxits=5000)
    s.solve(x, b, m, sp)
    assert rel_l2(x, x0) < 1e-9
    sp2 = qb.SolverParams(r2req=1e-20, maxits=5000, usePrevSoln=True)
    s.solve(x, b, m, sp2)
    assert sp2.iterations == 0
    z = lat.vector()
    out = np.ones_like(z)
    sp3 = qb.SolverParams(r2req=1e-10)
    s.solveEE(out, z, m, sp3)
    assert sp3.iterations == 0 and np.all(out == 0)
    sp4 = qb.SolverParams(r2req=1e-30, maxits=7)
    s.solveEE(out, b, m, sp4)
    assert sp4.iterations == 7
    s.free()


def test_device_resident_api(orc, ctx_factory):
    """Device-field entry points give the same numbers as the host-buffer ones; reductions match."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g, (x, y) = make_inputs(L, 8)
    s = qb.Staggered([g[k] for k in range(8)])
    fx, fy, fr = qb.DeviceField(x), qb.DeviceField(y), qb.DeviceField()
    assert abs(fx.norm2() - lat.norm2(x)) <= 1e-12 * lat.norm2(x)
    assert abs(fx.norm2(qb.ODD) - lat.norm2(x, 1)) <= 1e-12 * lat.norm2(x)
    assert abs(fx.redot(fy, qb.EVEN) - lat.redot(x, y, 0)) <= 1e-11 * np.sqrt(lat.norm2(x) * lat.norm2(y))
    s.dev_D(fr, fx, 0.2)
    want = lat.vector()
    lat.D(g, want, x, 0.2)
    assert rel_l2(fr.download(lat.vector()), want) <= DSLASH_TOL
    s.dev_stagD2xx(0, fr, fx, 0.01)
    lat.stagD2xx(g, want, x, 0.01, 0)
    assert rel_l2(fr.download(lat.vector())[:lat.half], want[:lat.half]) <= DSLASH_TOL
    # CG determinism: two identical solves give bit-identical results
    sp = qb.SolverParams(r2req=1e-10)
    s.dev_solveXX(fr, fx, 0.1, sp)
    a = fr.download(lat.vector()).copy()
    sp = qb.SolverParams(r2req=1e-10)
    s.dev_solveXX(fr, fx, 0.1, sp)
    assert np.array_equal(a, fr.download(lat.vector()))
    for f in (fx, fy, fr):
        f.free()
    s.free()


def test_large_lattice_properties(qb, ctx_factory):
    """BASELINE config 2 size (32^4) without the oracle: anti-Hermiticity, D^+D block identity and a
    CG round trip, on synthetic SU(3) links generated with numpy."""
    from qex_b200.synth import random_su3_links, gaussian_vector
    L = (32, 32, 32, 32)
    qb = ctx_factory(L)
    vol = int(np.prod(L))
    g = random_su3_links(vol, 4, seed=1)
    s = qb.Staggered(g)
    s.rephase()
    x, y = gaussian_vector(vol, 2), gaussian_vector(vol, 3)
    fx, fy, fa, fb = qb.DeviceField(x), qb.DeviceField(y), qb.DeviceField(), qb.DeviceField()
    s.dev_D(fa, fx, 0.0)
    s.dev_D(fb, fy, 0.0)
    # Re<y, D x> = -Re<D y, x>
    lhs, rhs = fy.redot(fa), fb.redot(fx)
    assert abs(lhs + rhs) <= 1e-12 * np.sqrt(fx.norm2() * fy.norm2()) * 8
    m = 0.1
    sp = qb.SolverParams(r2req=1e-16, maxits=5000)
    s.dev_D(fb, fx, m)               # b = D x
    s.dev_solve(fa, fb, m, sp)       # solve D a = b
    assert sp.r2 <= 1e-16 and sp.iterations > 10
    out = fa.download(np.zeros_like(x))
    assert rel_l2(out, x) < 1e-6
    for f in (fx, fy, fa, fb):
        f.free()
    s.free()


def test_remaining_stagD_flavours(orc, ctx_factory):
    """peqDdag, eoReduce, stagDb, stagD1, stagD1x (stagD.nim:414-427,572-580) -- parameter choices of the sweep."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g, (x, r0) = make_inputs(L, 8)
    s = qb.Staggered([g[k] for k in range(8)])
    m = 0.13
    dd = lat.vector()
    lat.Ddag(g, dd, x, m)
    got = r0.copy()
    s.peqDdag(got, x, m)
    assert rel_l2(got, r0 + dd) <= DSLASH_TOL
    got = r0.copy()
    s.eoReduce(got, x, m)
    assert rel_l2(got[:lat.half], dd[:lat.half]) <= DSLASH_TOL and np.array_equal(got[lat.half:], r0[lat.half:])
    fr, fx = qb.DeviceField(r0), qb.DeviceField(x)
    s.peqDdag(fr, fx, m)
    assert rel_l2(fr.download(lat.vector()), r0 + dd) <= DSLASH_TOL
    fr.free(); fx.free()
    for parity in (0, 1):
        sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
        hop = lat.vector()
        lat.stagD2(g, hop, x, 0.0, 0.0, parity)
        got = r0.copy()
        qb.stagD1(s, parity, got, x)
        assert rel_l2(got[sl], hop[sl]) <= DSLASH_TOL
        got = r0.copy()
        qb.stagD1x(s, parity, got, x)
        assert rel_l2(got[sl], -hop[sl]) <= DSLASH_TOL
        want = lat.vector()
        lat.stagD2(g, want, x, 0.0, m / (0.5 * 0.7), parity)
        got = r0.copy()
        qb.stagDb(s, parity, got, x, m, 0.7)
        assert rel_l2(got[sl], want[sl]) <= DSLASH_TOL
    s.free()


def test_solveEO_starts_from_the_callers_guess(orc, ctx_factory):
    """solveEO (stagSolve.nim:17-38): cgSolve(r, x, stagD2ee) with r's content as the initial guess -- against the
    oracle's CG from the same guess; a good guess must cost fewer iterations than the zero start."""
    L, m, r2req = (8, 8, 8, 16), 0.1, 1e-12
    qb = ctx_factory(L)
    lat, g, (b, g0) = make_inputs(L, 4)
    s = qb.Staggered([g[k] for k in range(4)])
    for guess in (0.01 * g0, None):
        if guess is None:       # the solution of a looser solve as the guess
            guess = lat.vector()
            lat.cg_xx(g, guess, b, m * m, 0, 1e-6, 10000)
        want = guess.copy()
        it_o, r2_o, b2_o = lat.cg_xx(g, want, b, m * m, 0, r2req, 10000)
        got = guess.copy()
        sp = qb.SolverParams(r2req=r2req, maxits=10000)
        s.solveEO(got, b, m, sp)
        assert abs(sp.iterations - it_o) <= 1, (sp.iterations, it_o)
        assert rel_l2(got[:lat.half], want[:lat.half]) <= 10 * np.sqrt(r2req)
        assert np.array_equal(got[lat.half:], guess[lat.half:])       # the odd part is not touched
        assert sp.r2 <= r2req
        assert sp.flops == (4 * 4 * 72 + 60) * lat.half * sp.iterations
    zero = lat.vector()
    it_zero, _, _ = lat.cg_xx(g, zero, b, m * m, 0, r2req, 10000)
    assert sp.iterations < it_zero
    # device-resident flavour, bit-identical with the host-buffer call
    dr, dx = qb.DeviceField(guess), qb.DeviceField(b)
    sp2 = qb.SolverParams(r2req=r2req, maxits=10000)
    s.solveEO(dr, dx, m, sp2)
    assert sp2.iterations == sp.iterations
    assert np.array_equal(dr.download(lat.vector())[:lat.half], got[:lat.half])
    dr.free(); dx.free(); s.free()


def test_cgstate_restart_continues_the_same_krylov_sequence(orc, ctx_factory):
    """CgState (cg.nim:13-53,256-283): solve() called again without reset continues where it stopped -- capped at
    20 iterations and then released it must give the iterates of one uninterrupted solve bit for bit; after
    reset() it starts over from the current x (fewer iterations than from zero)."""
    L, m, r2req = (8, 8, 8, 8), 0.1, 1e-12
    qb = ctx_factory(L)
    lat, g, (b, _) = make_inputs(L, 8)
    s = qb.Staggered([g[k] for k in range(8)])
    db = qb.DeviceField(b)
    x1, x2 = qb.DeviceField(), qb.DeviceField()
    one = qb.SolverParams(r2req=r2req, maxits=10000)
    s.dev_solveXX(x1, db, m, one, parEven=True)
    x2.zero()
    cg = qb.CgState(s, x2, db, parEven=True)
    sp = qb.SolverParams(r2req=r2req, maxits=20)
    cg.solve(m, sp)
    assert sp.iterations == 20 and sp.r2 > r2req
    mid = x2.download(lat.vector()).copy()
    sp = qb.SolverParams(r2req=r2req, maxits=10000)
    cg.solve(m, sp)
    assert sp.iterations == one.iterations                   # cumulative, like sp.finalIterations = state.iterations
    assert sp.r2 == one.r2
    a, c = x1.download(lat.vector()), x2.download(lat.vector())
    assert np.array_equal(a[:lat.half], c[:lat.half])
    assert not np.array_equal(mid[:lat.half], c[:lat.half])
    # a converged state does nothing more
    sp3 = qb.SolverParams(r2req=r2req, maxits=10000)
    cg.solve(m, sp3)
    assert sp3.iterations == one.iterations
    # reset: b2 and r are recomputed from the current x; tightening the request now takes a few iterations only
    cg.reset()
    sp = qb.SolverParams(r2req=1e-16, maxits=10000)
    cg.solve(m, sp)
    assert 0 < sp.iterations < one.iterations // 2
    want = lat.vector()
    lat.cg_xx(g, want, b, m * m, 0, 1e-16, 10000)
    assert rel_l2(x2.download(lat.vector())[:lat.half], want[:lat.half]) <= 1e-7
    cg.free()
    for d in (db, x1, x2, s):
        d.free()


def test_reunit_makes_qex_links_exact_and_12_real_storage_applies(orc, ctx_factory):
    """reunit (staghmc_sh.nim:247-257 = checkSU, projectSU, checkSU): g.random leaves links unitary to ~1e-11 (the
    12-real form refuses them); after reunit they are unitary to ~1e-15, the 12-real form is accepted, and the
    Dslash on them matches the oracle's 18-real Dslash on the SAME (downloaded) links to 1e-13."""
    L = (8, 8, 8, 16)
    qb = ctx_factory(L)
    lat = orc.Lattice(L)
    r = qb.newRNGField(987654321)
    g = qb.DeviceGauge(nlinks=4)
    g.random(r)
    got_rc, dev = g.recon(12)
    assert got_rc[0] == 18 and dev[0] > 2e-14               # as generated: refused
    before, after = g.reunit()
    assert 1e-13 < before[1] < 1e-9 and before[0] <= before[1]
    assert after[1] < 5e-15 and after[0] <= after[1]
    links = [lat.gauge(1)[0] for _ in range(4)]
    g.download(links)
    go = lat.random_gauge(lat.new_rng(987654321), 4)
    assert 0 < np.abs(np.stack(links) - go).max() < 1e-9     # moved, by the size of the old deviation only
    g.rephase()
    got_rc, dev = g.recon(12)
    assert got_rc[0] == 12 and dev[0] <= 2e-14, (got_rc, dev)
    gl = np.ascontiguousarray(np.stack(links))
    lat.setBC(gl); lat.stagPhase(gl)
    x = lat.gaussian_vector(lat.new_rng(3))
    s = qb.newStag(g)
    for parity in (0, 1):
        want = lat.vector()
        lat.stagD2(gl, want, x, 0.0, 0.0, parity)
        got = lat.vector()
        qb.stagD2(s, parity, got, x, 0.0, 0.0)
        sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
        assert rel_l2(got[sl], want[sl]) <= DSLASH_TOL
    r.free(); g.free()


def test_gauge_handle_cache_of_the_glue(orc, ctx_factory):
    """qexb200_gauge_cached: the device copy is keyed on the host link pointers and the
    caller's update counter -- second call: no upload, same handle; bumped counter: re-upload into the same handle
    (new results); other pointers: another handle; the ninth set evicts the least recently used one."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g, (x, _) = make_inputs(L, 4)
    links = [g[k].copy() for k in range(4)]
    s1, hit = qb.cachedStag(links, 1)
    assert not hit
    want, got = lat.vector(), lat.vector()
    lat.D(g, want, x, 0.1)
    s1.D(got, x, 0.1)
    assert rel_l2(got, want) <= DSLASH_TOL
    n0 = qb.launch_count()
    s2, hit = qb.cachedStag(links, 1)
    assert hit and s2.h == s1.h and qb.launch_count() == n0          # nothing ran: no re-layout kernel
    for l in links:                                                   # the caller rewrites its links in place ...
        l *= -1.0
    s3, hit = qb.cachedStag(links, 1)
    s3.D(got, x, 0.1)
    lat.D(-g, want, x, 0.1)
    # ... same counter: the sampled fingerprint of the host links sees a whole-field rewrite and uploads again
    # (qudaSolveXX uploads every call, qudaWrapperImpl.nim:198-261: a forgotten bump must not solve on stale links)
    assert not hit and s3.h == s1.h and rel_l2(got, want) <= DSLASH_TOL
    s4, hit = qb.cachedStag(links, 2)                                 # ... bumped counter: uploaded again, same handle
    assert not hit and s4.h == s1.h
    s4.D(got, x, 0.1)
    assert rel_l2(got, want) <= DSLASH_TOL
    links[0][5, 1, 1, 0] += 1.0                                       # a change between the samples is NOT seen: the
    _, hit = qb.cachedStag(links, 2)                                  # counter stays the contract (documented limit)
    assert hit
    links[0][5, 1, 1, 0] -= 1.0
    others = [[g[k].copy() for k in range(4)] for _ in range(8)]
    hs = [qb.cachedStag(o, 1)[0].h for o in others]
    assert len(set(hs)) == 8 and s1.h not in hs
    _, hit = qb.cachedStag(links, 2)                                  # evicted by the eight newer sets
    assert not hit
    _, hit = qb.cachedStag(others[-1], 1)
    assert hit
    qb.invalidateStag(others[-1])                                     # the glue's finaliser: entry and handle are gone
    _, hit = qb.cachedStag(others[-1], 1)
    assert not hit
    qb.invalidateStag([g[k].copy() for k in range(4)])                # unknown pointers: not an error


def test_sloppySolve_is_accepted_and_reported_as_fp64(orc, ctx_factory):
    """SolverParams.sloppySolve (solverBase.nim:8-9,15; qudaWrapperImpl.nim:194-197) crosses the boundary.  There are no
    mixed-precision kernels: SloppySingle / SloppyHalf run the same fp64 solve (bit-identical solution, same iteration
    count) and the library says so in sloppy_used; an out-of-range value is an error, not a silent default."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g, (x, _) = make_inputs(L, 4)
    s = qb.Staggered([g[k] for k in range(4)])
    sols, its = [], []
    for sloppy in (0, 1, 2):
        sp = qb.SolverParams(r2req=1e-10, sloppySolve=sloppy)
        sol = lat.vector()
        s.solve(sol, x, 0.1, sp)
        assert sp.sloppyUsed == 0
        sols.append(sol); its.append(sp.iterations)
    assert its[0] == its[1] == its[2]
    assert np.array_equal(sols[0], sols[1]) and np.array_equal(sols[0], sols[2])
    with pytest.raises(qb.QexB200Error, match="sloppy_solve"):
        s.solve(lat.vector(), x, 0.1, qb.SolverParams(r2req=1e-10, sloppySolve=7))
    s.free()
