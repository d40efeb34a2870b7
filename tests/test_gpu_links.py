"""GPU parity of the device link pipeline through the C ABI: makeImpLinks
(src/gauge/fat7l.nim:78-162), HISQ (src/physics/hisqLinks.nim:32-39), nHYP (src/gauge/hypsmear.nim:49-162),
projectU, plaq -- against the oracle on the same seeded links.  Bar: rel-L2 <= 1e-13 per link field."""
import numpy as np
import pytest

import oracle
from util import SEED, rel_l2, hisq_reference, hisq_bar

pytestmark = pytest.mark.gpu

TOL = 1e-13


@pytest.fixture()
def ctx_factory(qb):
    def make(L, inner=(1, 1, 1, 1)):
        if qb.is_initialized():
            qb.finalize()
        qb.init(L)
        qb.set_host_layout(inner)
        return qb
    yield make
    if qb.is_initialized():
        qb.finalize()


def thin_links(L, phases=True, seed=SEED):
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(seed), 4)
    if phases:
        lat.setBC(g)
        lat.stagPhase(g)
    return lat, g


def up(qb, lo, g):
    return qb.DeviceGauge([lo.scatter(g[k])[0] for k in range(g.shape[0])])


def down(lat, lo, dg):
    parts = [lo.scatter(lat.gauge(1)[0])[0] for _ in range(dg.nlinks)]
    dg.download(parts)
    return np.stack([lo.gather([p]) for p in parts])


GEOMS = [((8, 8, 8, 8), (1, 1, 1, 1)), ((4, 6, 8, 10), (1, 1, 1, 1)), ((8, 8, 8, 8), (1, 2, 2, 2))]


@pytest.mark.parametrize("L,inner", GEOMS)
def test_plaq_and_roundtrip(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat, g = thin_links(L, phases=False)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    dg = up(qb, lo, g)
    assert np.allclose(dg.plaq(), lat.plaq(g), rtol=0, atol=1e-14)
    assert np.array_equal(down(lat, lo, dg), g)       # upload -> download is a pure permutation
    dg.free()


def test_fat7_unit_gauge_known_answer(ctx_factory):
    """The reference's own check (fat7l.nim:185-214): on the unit gauge every plaquette of the fat links
    is c^4/6 with c = oneLink + 6 threeStaple + 24 fiveStaple + 48 sevenStaple + 6 lepage."""
    L = (4, 4, 4, 4)
    qb = ctx_factory(L)
    lat = oracle.Lattice(L)
    lo = oracle.QexLayout(L)
    dg = up(qb, lo, lat.unit_gauge())
    fl = qb.DeviceGauge(nlinks=4)
    for c in [(1, 0, 0, 0, 0), (0, 1, 0, 0, 0), (0, 0, 1, 0, 0), (0, 0, 0, 1, 0), (0, 0, 0, 0, 1), (1, 1, 1, 1, 1)]:
        qb.makeImpLinks(fl, dg, qb.Fat7lCoefs(*[float(t) for t in c]))
        cc = c[0] + 6 * c[1] + 24 * c[2] + 48 * c[3] + 6 * c[4]
        want = cc ** 4 / 6.0
        assert np.allclose(fl.plaq(), want, rtol=1e-13), (c, fl.plaq(), want)
    fl.free(); dg.free()


COEFS = [
    (0.3, -0.2, 0.0, 0.0, 0.0),                          # 3-staple only
    (0.3, -0.2, 0.0, 0.0, 0.15),                         # + Lepage
    (0.3, -0.2, 0.11, 0.0, 0.0),                         # + 5-staple, no 7
    (0.625, -1 / 16., 1 / 64., -1 / 384., 0.0),          # HISQ first level
    (1.0, -1 / 16., 1 / 64., -1 / 384., -1 / 8.),        # HISQ second level (all five terms)
    (0.5, 0.0, 0.0, 0.0, 0.0),                           # one-link only
]


@pytest.mark.parametrize("L,inner", GEOMS[:2])
def test_makeImpLinks_parity(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat, g = thin_links(L)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    dg = up(qb, lo, g)
    fl, ll = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    for c in COEFS:
        want = lat.makeImpLinks(g, c)
        qb.makeImpLinks(fl, dg, qb.Fat7lCoefs(*c))
        got = down(lat, lo, fl)
        for mu in range(4):
            assert rel_l2(got[mu], want[mu]) <= TOL, (c, mu)
    # with Naik links (fat7l.nim:156-167)
    c = COEFS[4]
    wfl, wll = lat.makeImpLinks(g, c, gfLong=g, naik=-1 / 24.)
    qb.makeImpLinks(fl, dg, qb.Fat7lCoefs(*c), ll, dg, -1 / 24.)
    gfl, gll = down(lat, lo, fl), down(lat, lo, ll)
    for mu in range(4):
        assert rel_l2(gfl[mu], wfl[mu]) <= TOL
        assert rel_l2(gll[mu], wll[mu]) <= TOL
    for d in (dg, fl, ll):
        d.free()


@pytest.mark.parametrize("L,inner", GEOMS)
def test_projectU_parity(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat, g = thin_links(L)
    fat = lat.makeImpLinks(g, COEFS[3])
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    src, dst = up(qb, lo, fat), qb.DeviceGauge(nlinks=4)
    dst.projectU(src)
    got, want = down(lat, lo, dst), lat.projectU(fat)
    for mu in range(4):
        assert rel_l2(got[mu], want[mu]) <= TOL
    m = got[..., 0] + 1j * got[..., 1]
    assert np.abs(np.einsum("kxab,kxcb->kxac", m, m.conj()) - np.eye(3)).max() < 1e-12
    src.free(); dst.free()


@pytest.mark.parametrize("L,inner", GEOMS)
def test_hisq_smear_parity_and_solve(ctx_factory, L, inner):
    """HisqCoefs.smear on the device = the oracle's, and the operator built from the device-resident
    result (newStag3(fl, ll), src/physics/hisqLinks.nim:47-72 / tests/examples/testStagProp.nim:28-36) gives the
    oracle's Dslash and CG on the oracle's HISQ links."""
    qb = ctx_factory(L, inner)
    lat, g = thin_links(L)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    # Bar: QEX's closed-form projectU is itself ~1e-13 (L2; single sites 1e-11) from the exact polar factor on
    # these links, so the device is held to 10x the oracle's own distance from the exact-projection chain, both
    # against the oracle and against the exact chain (tests/util.py hisq_reference).
    wfl, wll, efl, ell = hisq_reference(lat, g, oracle.hisq_coefs())
    tol_fl = [hisq_bar(wfl, efl)] * 4
    tol_ll = [hisq_bar(wll, ell)] * 4
    assert max(tol_fl + tol_ll) < 1e-11
    dg = up(qb, lo, g)
    fl, ll = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    qb.HisqCoefs().init().smear(dg, fl, ll)
    gfl, gll = down(lat, lo, fl), down(lat, lo, ll)
    for mu in range(4):
        assert rel_l2(gfl[mu], wfl[mu]) <= tol_fl[mu], mu
        assert rel_l2(gll[mu], wll[mu]) <= tol_ll[mu], mu
        assert rel_l2(gfl[mu], efl[mu]) <= tol_fl[mu] and rel_l2(gll[mu], ell[mu]) <= tol_ll[mu], mu
        # and site by site the typical deviation is at rounding level (a wrong term would show here)
        site = np.linalg.norm((gfl[mu] - wfl[mu]).reshape(lat.vol, -1), axis=1) / np.linalg.norm(wfl[mu].reshape(lat.vol, -1), axis=1)
        assert np.median(site) < 1e-14, mu
    assert np.allclose(fl.plaq(), lat.plaq(wfl), rtol=1e-11, atol=1e-15)
    s = qb.newStag3(fl, ll)
    gg = np.zeros((8,) + wfl.shape[1:])
    gg[0::2], gg[1::2] = wfl, wll
    x = lat.gaussian_vector(lat.new_rng(11))
    want, got = lat.vector(), lo.scatter(lat.vector())[0]
    lat.D(gg, want, x, 0.05)
    s.D(got, lo.scatter(x)[0], 0.05)
    assert rel_l2(lo.gather([got]), want) <= max(tol_fl)      # oracle links vs device links: same bar
    sol, ref = lo.scatter(lat.vector())[0], lat.vector()
    sp = qb.SolverParams(r2req=1e-10)
    s.solve(sol, lo.scatter(x)[0], 0.05, sp)
    st = lat.solve(gg, ref, x, 0.05, 1e-10)
    assert abs(sp.iterations - st.iterations) <= 1
    assert rel_l2(lo.gather([sol]), ref) < 1e-4
    for d in (s, dg, fl, ll):
        d.free()


@pytest.mark.parametrize("L,inner", GEOMS)
def test_hyp_smear_parity_and_solve(ctx_factory, L, inner):
    """nHYP as staghmc_spv uses it (src/stagg_pv_hmc/staghmc_spv.nim:367-386): smear the thin links, then
    setBC + stagPhase on the smeared set, then solve -- all on the device."""
    qb = ctx_factory(L, inner)
    lat, g = thin_links(L, phases=False)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    alpha = (0.4, 0.5, 0.5)                      # tests/extra/staghmc_sh/ref.0:62-66
    want = lat.hyp_smear(g, alpha)
    dg, fl = up(qb, lo, g), qb.DeviceGauge(nlinks=4)
    qb.HypCoefs(*alpha).smear(dg, fl)
    got = down(lat, lo, fl)
    for mu in range(4):
        assert rel_l2(got[mu], want[mu]) <= TOL, mu
    assert np.allclose(fl.plaq(), lat.plaq(want), rtol=1e-12, atol=1e-15)
    fl.rephase()
    lat.setBC(want)
    lat.stagPhase(want)
    s = qb.newStag(fl)
    x = lat.gaussian_vector(lat.new_rng(3))
    sol, ref = lo.scatter(lat.vector())[0], lat.vector()
    sp = qb.SolverParams(r2req=1e-12)
    s.solve(sol, lo.scatter(x)[0], 0.1, sp)
    st = lat.solve(want, ref, x, 0.1, 1e-12)
    assert abs(sp.iterations - st.iterations) <= 1
    assert rel_l2(lo.gather([sol]), ref) < 1e-5
    for d in (s, dg, fl):
        d.free()


@pytest.mark.parametrize("L,grid", [((8, 8, 8, 8), "6"), ((4, 6, 8, 10), "2"), ((16, 16, 16, 16), None)])
def test_staged_staple_kernel_is_bit_identical(ctx_factory, monkeypatch, L, grid):
    """The shared-memory-staged staple kernel (k_staple_pipe, opt-in with QEXB200_STAPLE_PIPE=1) does the arithmetic of the
    register-gather kernel in the same order: HISQ and nHYP links must agree bit for bit, also when a block
    walks many tiles (QEXB200_PIPE_GRID) and when the last tile is ragged."""
    qb = ctx_factory(L)
    vol = int(np.prod(L))
    from qex_b200.synth import random_su3_links
    g = qb.DeviceGauge(random_su3_links(vol, 4, seed=5))
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("QEXB200_STAPLE_PIPE", mode)
        if grid and mode == "1":
            monkeypatch.setenv("QEXB200_PIPE_GRID", grid)
        fl, ll, hl = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
        qb.HisqCoefs().init().smear(g, fl, ll)
        qb.HypCoefs(0.4, 0.5, 0.5).smear(g, hl)
        res = []
        for dg in (fl, ll, hl):
            parts = [np.zeros((vol, 3, 3, 2)) for _ in range(4)]
            dg.download(parts)
            res.append(np.stack(parts))
            dg.free()
        out[mode] = res
    monkeypatch.delenv("QEXB200_PIPE_GRID", raising=False)
    for a, b in zip(out["1"], out["0"]):
        assert np.isfinite(a).all() and np.array_equal(a, b)
    g.free()


def test_recon14_on_projected_links(ctx_factory):
    """14-real link storage (rows 0,1 + one complex factor per link, include/qex_b200.h qexb200_gauge_recon) for
    links that are U(3) rather than SU(3): nHYP links smeared on the device.  The 12-real form must refuse them,
    the 14-real form must take them, the Dslash stays within the 1e-13 bar of the 18-real oracle, the solve
    within +-1 iteration, and the request survives a fresh smear + rephase."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g = thin_links(L, phases=False)
    lo = oracle.QexLayout(L)
    alpha = (0.4, 0.5, 0.5)
    want = lat.hyp_smear(g, alpha)
    lat.setBC(want)
    lat.stagPhase(want)
    dg, fl = up(qb, lo, g), qb.DeviceGauge(nlinks=4)
    hyp = qb.HypCoefs(*alpha)
    hyp.smear(dg, fl)
    fl.rephase()
    s = qb.newStag(fl)
    x = lat.gaussian_vector(lat.new_rng(3))
    ref, r18, r14 = lat.vector(), lat.vector(), lat.vector()
    lat.D(want, ref, x, 0.1)
    s.D(r18, x, 0.1)
    got_rc, dev = s.recon(12)
    assert got_rc[0] == 18 and dev[0] > 1e-3          # det phase != +-1: not s*SU(3)
    got_rc, dev = s.recon(14)
    assert got_rc[0] == 14 and dev[0] <= 2e-14, (got_rc, dev)
    s.D(r14, x, 0.1)
    assert rel_l2(r14, ref) <= TOL and rel_l2(r14, r18) <= 1e-14
    for parity in (0, 1):                               # a, b terms and both parities through stagD2
        a, b = lat.vector(), lat.vector()
        a[:] = x; b[:] = x
        lat.stagD2(want, a, x, 0.7, -0.3, parity)
        qb.stagD2(s, parity, b, x, 0.7, -0.3)
        sl = slice(0, lat.half) if parity == 0 else slice(lat.half, lat.vol)
        assert rel_l2(b[sl], a[sl]) <= TOL
    sol, solref = lat.vector(), lat.vector()
    sp = qb.SolverParams(r2req=1e-12)
    s.solve(sol, x, 0.1, sp)
    st = lat.solve(want, solref, x, 0.1, 1e-12)
    assert abs(sp.iterations - st.iterations) <= 1
    assert rel_l2(sol, solref) < 1e-5
    # sticky: smear again into the same set, phases on the device, still 14 reals and still right
    hyp.smear(dg, fl)
    fl.rephase()
    assert s.recon(14)[0][0] == 14
    s.D(r14, x, 0.1)
    assert rel_l2(r14, ref) <= TOL
    assert s.recon(18)[0] == [18, 18]
    s.D(r14, x, 0.1)
    assert np.array_equal(r14, r18)
    for d in (s, dg, fl):
        d.free()


def test_recon14_on_hisq_long_links(ctx_factory):
    """HISQ: the fat links are sums of staples (stay at 18 reals), the Naik links are naik * W W W with W in U(3)
    (src/gauge/fat7l.nim:156-167 on the projected links of src/physics/hisqLinks.nim:32-39): 14 reals."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    # Links near the unit gauge, like a thermalised configuration.  On a hot start the closed-form projectU of the
    # reference leaves a few per cent of the W links unitary to 1e-11 only (nearly singular fat links); the Naik
    # products then miss the 2e-14 bar and the library keeps them at 18 reals, as it must.
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    g = lat.projectU(lat.unit_gauge() + 0.15 * np.random.default_rng(8).standard_normal(lat.unit_gauge().shape))
    lat.setBC(g)
    lat.stagPhase(g)
    dg, fl, ll = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    qb.HisqCoefs().init().smear(dg, fl, ll)
    s = qb.newStag3(fl, ll)
    x = lat.gaussian_vector(lat.new_rng(4))
    r18, r14 = lat.vector(), lat.vector()
    s.D(r18, x, 0.05)
    got_rc, dev = s.recon(12)
    assert got_rc == [18, 18]
    got_rc, dev = s.recon(14)
    assert got_rc == [18, 14] and dev[1] <= 2e-14 and dev[0] > 1e-3, (got_rc, dev)
    s.D(r14, x, 0.05)
    assert rel_l2(r14, r18) <= 1e-14
    assert s.recon(1814)[0] == [18, 14] and s.recon(1418)[0] == [18, 18]
    fat, lng, efat, elng = hisq_reference(lat, g, oracle.hisq_coefs())
    links = np.empty((8,) + fat.shape[1:])
    links[0::2], links[1::2] = fat, lng
    ref = lat.vector()
    lat.D(links, ref, x, 0.05)
    assert rel_l2(r14, ref) <= max(hisq_bar(fat, efat), hisq_bar(lng, elng), TOL)   # device links vs oracle links
    sp, sp18 = qb.SolverParams(r2req=1e-10), qb.SolverParams(r2req=1e-10)
    a, b = lat.vector(), lat.vector()
    s.recon(1814)
    s.solve(a, x, 0.05, sp)
    s.recon(18)
    s.solve(b, x, 0.05, sp18)
    assert abs(sp.iterations - sp18.iterations) <= 1 and rel_l2(a, b) < 1e-6
    for d in (s, dg, fl, ll):
        d.free()


def test_hyp_unit_gauge_is_fixed_point(ctx_factory):
    L = (4, 4, 4, 4)
    qb = ctx_factory(L)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    dg, fl = up(qb, lo, lat.unit_gauge()), qb.DeviceGauge(nlinks=4)
    qb.HypCoefs(0.4, 0.5, 0.5).smear(dg, fl)
    assert np.abs(down(lat, lo, fl) - lat.unit_gauge()).max() < 1e-14
    dg.free(); fl.free()


def test_link_pipeline_errors(ctx_factory):
    qb = ctx_factory((4, 4, 4, 4))
    lat, lo = oracle.Lattice((4, 4, 4, 4)), oracle.QexLayout((4, 4, 4, 4))
    dg, empty = up(qb, lo, lat.unit_gauge()), qb.DeviceGauge(nlinks=4)
    with pytest.raises(qb.QexB200Error):
        qb.makeImpLinks(dg, dg, qb.Fat7lCoefs(1, 0, 0, 0, 0))      # in place is refused
    with pytest.raises(qb.QexB200Error):
        qb.HypCoefs(0.4, 0.5, 0.5).smear(empty, dg)                # source holds no links
    with pytest.raises(qb.QexB200Error):
        empty.plaq()
    dg.free(); empty.free()
