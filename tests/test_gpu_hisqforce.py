"""GPU parity of fat7lDeriv (src/gauge/fat7lderiv.nim:122-335) and the HISQ smearing force (smearGetForce / smearedForce,
src/gauge/hisqsmear.nim:51-84) against the oracle, which passes the reference's own finite-difference checks
(fat7lderiv.nim:366-398, hisqsmear.nim:143-184; tests/test_oracle_golden.py::test_fat7_and_hisq_force_are_gradients)."""
import numpy as np
import pytest

import oracle
from util import SEED, rel_l2, hisq_reference, hisq_bar

pytestmark = pytest.mark.gpu


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


def up(qb, lo, g):
    return qb.DeviceGauge([lo.scatter(g[k])[0] for k in range(g.shape[0])])


def down(lat, lo, dg):
    parts = [lo.scatter(lat.gauge(1)[0])[0] for _ in range(dg.nlinks)]
    dg.download(parts)
    return np.stack([lo.gather([p]) for p in parts])


COEFS = [(1, 0, 0, 0, 0), (0, 1, 0, 0, 0), (0, 0, 1, 0, 0), (0, 0, 0, 1, 0), (0, 0, 0, 0, 1),      # fat7lderiv.nim:366-386
         (1, 1, 1, 1, 0), (0.625, -1 / 16., 1 / 64., -1 / 384., 0.0), (1.0, -1 / 16., 1 / 64., -1 / 384., -1 / 8.)]


@pytest.mark.parametrize("L,inner", [((8, 8, 8, 8), (1, 1, 1, 1)), ((4, 6, 8, 10), (1, 1, 1, 1))])
def test_fat7lDeriv_parity(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L, (1, 1, 1, 1), inner)
    g = lat.random_gauge(lat.new_rng(SEED), 4)
    lat.setBC(g); lat.stagPhase(g)
    rng = np.random.default_rng(2)
    mid, llmid, start = rng.standard_normal(g.shape), rng.standard_normal(g.shape), rng.standard_normal(g.shape)
    dg, dm, dlm = up(qb, lo, g), up(qb, lo, mid), up(qb, lo, llmid)
    for c in COEFS:
        want = start.copy()                               # accumulates
        lat.fat7lDeriv(want, g, mid, [float(t) for t in c])
        dd = up(qb, lo, start)
        qb.fat7lDeriv(dd, dg, dm, qb.Fat7lCoefs(*[float(t) for t in c]))
        got = down(lat, lo, dd)
        for mu in range(4):
            assert rel_l2(got[mu], want[mu]) <= 1e-13, (c, mu)
        dd.free()
    c = COEFS[-1]
    want, wll = start.copy(), 0.5 * start
    lat.fat7lDeriv(want, g, mid, c, wll, g, llmid, -1 / 24.)
    dd, dl = up(qb, lo, start), up(qb, lo, 0.5 * start)
    qb.fat7lDeriv(dd, dg, dm, qb.Fat7lCoefs(*c), dl, dg, dlm, -1 / 24.)
    got, gll = down(lat, lo, dd), down(lat, lo, dl)
    for mu in range(4):
        assert rel_l2(got[mu], want[mu]) <= 1e-13 and rel_l2(gll[mu], wll[mu]) <= 1e-13, mu
    for d in (dg, dm, dlm, dd, dl):
        d.free()


def conditioning(lat, g, hc, ch, chl, want):
    """The oracle's own response to 1-ulp changes of the thin links (three samples): on random links a few sites
    have nearly singular fat links (smallest singular value 3e-4), the derivative of the polar factor there goes
    like 1/(s_i + s_j), and the force -- spread over their neighbours -- moves by ~1e-10 in L2 under ANY change
    of rounding.  Returns (L2 response, median per-site response)."""
    l2, med = 0.0, 0.0
    for seed in range(3):
        sgn = np.random.default_rng(seed).choice([-1.0, 1.0], size=g.shape)
        fp = lat.hisq_force(g * (1.0 + 1.1e-16 * sgn), hc, ch, chl)
        site = np.linalg.norm((fp - want).reshape(4 * lat.vol, -1), axis=1) / np.linalg.norm(want.reshape(4 * lat.vol, -1), axis=1)
        l2, med = max(l2, rel_l2(fp, want)), max(med, float(np.median(site)))
    return l2, med


@pytest.mark.parametrize("L,inner,smooth", [((8, 8, 8, 8), (1, 1, 1, 1), False), ((4, 6, 8, 10), (1, 1, 1, 1), False),
                                            ((8, 8, 8, 8), (1, 2, 2, 2), False), ((8, 8, 8, 8), (1, 1, 1, 1), True)])
def test_hisq_smeared_force_parity(ctx_factory, L, inner, smooth):
    qb = ctx_factory(L, inner)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L, (1, 1, 1, 1), inner)
    if smooth:      # links near the unit gauge: fat links far from singular, the chain is well conditioned
        g = lat.projectU(lat.unit_gauge() + 0.15 * np.random.default_rng(8).standard_normal(lat.unit_gauge().shape))
    else:
        g = lat.random_gauge(lat.new_rng(SEED), 4)
    lat.setBC(g); lat.stagPhase(g)
    rng = np.random.default_rng(3)
    ch, chl = rng.standard_normal(g.shape), rng.standard_normal(g.shape)
    hc = oracle.hisq_coefs()
    want = lat.hisq_force(g, hc, ch, chl)
    if smooth:
        bar_l2, bar_med = 1e-12, 1e-13
    else:
        r_l2, r_med = conditioning(lat, g, hc, ch, chl, want)
        assert r_l2 < 1e-8
        bar_l2, bar_med = 4 * r_l2, 4 * r_med
    dg, su, sul, du = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    dch, dchl = up(qb, lo, ch), up(qb, lo, chl)
    force = qb.HisqCoefs().init().smearGetForce(dg, su, sul)
    du.zero()
    force(du, dch, dchl)
    got = down(lat, lo, du)
    site = np.linalg.norm((got - want).reshape(4 * lat.vol, -1), axis=1) / np.linalg.norm(want.reshape(4 * lat.vol, -1), axis=1)
    assert rel_l2(got, want) <= bar_l2, (rel_l2(got, want), bar_l2)
    assert np.median(site) <= bar_med, (float(np.median(site)), bar_med)
    force(du, dch, dchl)                                  # accumulates: twice the force now
    assert rel_l2(down(lat, lo, du), 2 * want) <= bar_l2
    force.free()
    for d in (dg, su, sul, du, dch, dchl):
        d.free()


def test_device_hisq_force_is_the_gradient(ctx_factory):
    """The reference's own check (hisqsmear.nim:143-160) run on the device: central differences of
    sum ch.fl + chl.ll under the DEVICE smearing against sum dg.f with the DEVICE force."""
    L = (4, 4, 4, 4)
    qb = ctx_factory(L)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    g = lat.random_gauge(lat.new_rng(7), 4)
    lat.stagPhase(g)
    rng = np.random.default_rng(0)
    ch, chl, dg_ = rng.standard_normal(g.shape), rng.standard_normal(g.shape), rng.standard_normal(g.shape)
    hc = qb.HisqCoefs().init()

    def action(gg):
        a, b, c = up(qb, lo, gg), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
        hc.smear(a, b, c)
        v = float(np.sum(ch * down(lat, lo, b)) + np.sum(chl * down(lat, lo, c)))
        for d in (a, b, c):
            d.free()
        return v
    eps = 1e-6
    fd = (action(g + eps * dg_) - action(g - eps * dg_)) / (2 * eps)
    a, su, sul, du = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    dch, dchl = up(qb, lo, ch), up(qb, lo, chl)
    force = hc.smearGetForce(a, su, sul)
    du.zero()
    force(du, dch, dchl)
    an = float(np.sum(down(lat, lo, du) * dg_))
    assert abs(fd - an) <= 1e-6 * abs(an), (fd, an)
    force.free()
    for d in (a, su, sul, du, dch, dchl):
        d.free()


def test_fermion_force_step_of_hisqhmc(ctx_factory):
    """fermionForce of src/examples/hisqhmc.nim:496-545 on the device: smearRephase (:405-414) -> solve -> one-hop and
    three-hop outer products p (x) p(+mu)^+, p (x) p(+3mu)^+ (one pass), odd sites negated -> HISQ smearedForce on the
    phased thin links -> projectTAH(ff g^+).  Near-unit links (well conditioned chain, see above)."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    u = lat.projectU(lat.unit_gauge() + 0.15 * np.random.default_rng(8).standard_normal(lat.unit_gauge().shape))
    phi = lat.gaussian_vector(lat.new_rng(17))
    mass, r2req = 0.1, 1e-20
    hc = oracle.hisq_coefs()
    # oracle
    up_ = u.copy()
    lat.setBC(up_); lat.stagPhase(up_)                     # g.rephase
    su, sul = lat.hisq_smear(up_, hc)
    gg = np.zeros((8,) + su.shape[1:])
    gg[0::2], gg[1::2] = su, sul
    psi = lat.vector()
    st = lat.solve(gg, psi, phi, mass, r2req)
    f13 = lat.gauge(8)
    lat.outer(f13, psi, psi, 1.0, -1.0, accumulate=False)
    ff = lat.hisq_force(up_, hc, np.ascontiguousarray(f13[0::2]), np.ascontiguousarray(f13[1::2]))
    m = (ff[..., 0] + 1j * ff[..., 1]) @ np.conj(up_[..., 0] + 1j * up_[..., 1]).transpose(0, 1, 3, 2)
    t = 0.5 * (m - np.conj(m).transpose(0, 1, 3, 2))
    t = t - np.einsum("kxaa->kx", t)[..., None, None] / 3.0 * np.eye(3)
    want = np.stack([t.real, t.imag], axis=-1)
    # device
    du = up(qb, lo, u)
    du.rephase()
    dsu, dsul = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    force = qb.HisqCoefs().init().smearGetForce(du, dsu, dsul)
    s = qb.newStag3(dsu, dsul)
    dphi, dpsi = qb.DeviceField(phi), qb.DeviceField()
    sp = qb.SolverParams(r2req=r2req)
    s.dev_solve(dpsi, dphi, mass, sp)
    assert abs(sp.iterations - st.iterations) <= 2
    f8, f1, f3, dff = qb.DeviceGauge(nlinks=8), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    f8.outer(dpsi, dpsi, 1.0, -1.0, accumulate=False)
    f8.split3(f1, f3)
    dff.zero()
    force(dff, f1, f3)
    dff.projTAH(du)
    got = down(lat, lo, dff)
    assert rel_l2(got, want) <= 1e-8, rel_l2(got, want)
    force.free()
    for d in (s, du, dsu, dsul, dphi, dpsi, f8, f1, f3, dff):
        d.free()
