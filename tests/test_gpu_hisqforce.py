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


@pytest.mark.parametrize("L,inner", [((8, 8, 8, 8), (1, 1, 1, 1)), ((4, 6, 8, 10), (1, 1, 1, 1)), ((8, 8, 8, 8), (1, 2, 2, 2))])
def test_hisq_smeared_force_parity(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L, (1, 1, 1, 1), inner)
    g = lat.random_gauge(lat.new_rng(SEED), 4)
    lat.setBC(g); lat.stagPhase(g)
    rng = np.random.default_rng(3)
    ch, chl = rng.standard_normal(g.shape), rng.standard_normal(g.shape)
    hc = oracle.hisq_coefs()
    want = lat.hisq_force(g, hc, ch, chl)
    wfl, wll, efl, ell = hisq_reference(lat, g, hc)
    bar = max(hisq_bar(wfl, efl), hisq_bar(wll, ell))     # the chain goes through the same closed-form projection
    dg, su, sul, du = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    dch, dchl = up(qb, lo, ch), up(qb, lo, chl)
    force = qb.HisqCoefs().init().smearGetForce(dg, su, sul)
    gfl, gll = down(lat, lo, su), down(lat, lo, sul)
    assert rel_l2(gfl, wfl) <= bar and rel_l2(gll, wll) <= bar
    du.zero()
    force(du, dch, dchl)
    got = down(lat, lo, du)
    assert rel_l2(got, want) <= 10 * bar, (rel_l2(got, want), bar)
    site = np.linalg.norm((got - want).reshape(4 * lat.vol, -1), axis=1) / np.linalg.norm(want.reshape(4 * lat.vol, -1), axis=1)
    assert np.median(site) < 1e-13, float(np.median(site))
    force(du, dch, dchl)                                  # accumulates: twice the force now
    assert rel_l2(down(lat, lo, du), 2 * want) <= 10 * bar
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
