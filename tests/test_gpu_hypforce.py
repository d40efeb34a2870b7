"""GPU parity of the nHYP smearing force (smearGetForce / smearedForce, src/gauge/hypsmear.nim:49-262; projectUderiv,
src/maths/matrixFunctions.nim:321-357; symStapleDeriv, src/gauge/smearutil.nim:24-54) and projTAH
(src/stagg_pv_hmc/staghmc_spv_gforce.nim:256-291) against the oracle, whose chain is itself pinned by finite differences
(tests/test_oracle_golden.py::test_hyp_force_is_the_gradient)."""
import numpy as np
import pytest

import oracle
from util import SEED, rel_l2, make_inputs

pytestmark = pytest.mark.gpu
ALPHA = (0.4, 0.5, 0.5)
BAR = 1e-13          # nHYP's projected sums are well conditioned (measured 2e-15 on 2 ranks)
SITE_BAR = 1e-14     # median per-site deviation


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


def check(got, want, vol):
    for mu in range(4):
        assert rel_l2(got[mu], want[mu]) <= BAR, (mu, rel_l2(got[mu], want[mu]))
        site = np.linalg.norm((got[mu] - want[mu]).reshape(vol, -1), axis=1) / np.linalg.norm(want[mu].reshape(vol, -1), axis=1)
        assert np.median(site) <= SITE_BAR, (mu, float(np.median(site)))


@pytest.mark.parametrize("L,inner", [((8, 8, 8, 8), (1, 1, 1, 1)), ((4, 6, 8, 10), (1, 1, 1, 1)), ((8, 8, 8, 8), (1, 2, 2, 2))])
def test_smeared_force_parity(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat = oracle.Lattice(L)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    g = lat.random_gauge(lat.new_rng(SEED), 4)
    chain = np.random.default_rng(4).standard_normal(g.shape)
    want, wfl = lat.hyp_force(g, ALPHA, chain, want_fl=True)
    dg, fl, f, dc = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), up(qb, lo, chain)
    force = qb.HypCoefs(*ALPHA).smearGetForce(dg, fl)
    gfl = down(lat, lo, fl)
    for mu in range(4):
        assert rel_l2(gfl[mu], wfl[mu]) <= 1e-13          # the smeared links themselves
    force(f, dc)
    check(down(lat, lo, f), want, lat.vol)
    force(dc, dc)                                         # in place, like f.smeared_force(f) (staghmc_spv.nim:745)
    check(down(lat, lo, dc), want, lat.vol)
    force.free()
    for d in (dg, fl, f, dc):
        d.free()


def test_device_force_is_the_gradient_of_the_device_smearing(ctx_factory):
    """sum chain . dV(U)/dU by central differences of the DEVICE smearing = sum f . dU with the DEVICE force."""
    L = (4, 4, 4, 4)
    qb = ctx_factory(L)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    g = lat.random_gauge(lat.new_rng(7), 4)
    rng = np.random.default_rng(0)
    chain, dg_ = rng.standard_normal(g.shape), rng.standard_normal(g.shape)
    coef = qb.HypCoefs(*ALPHA)

    def action(gg):
        a, b = up(qb, lo, gg), qb.DeviceGauge(nlinks=4)
        coef.smear(a, b)
        v = down(lat, lo, b)
        a.free(); b.free()
        return float(np.sum(chain * v))
    eps = 1e-6
    fd = (action(g + eps * dg_) - action(g - eps * dg_)) / (2 * eps)
    a, fl, f, dc = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), up(qb, lo, chain)
    force = coef.smearGetForce(a, fl)
    force(f, dc)
    an = float(np.sum(down(lat, lo, f) * dg_))
    assert abs(fd - an) <= 1e-7 * abs(an), (fd, an)
    force.free()
    for d in (a, fl, f, dc):
        d.free()


def test_matter_force_tail(ctx_factory):
    """fforce's tail (staghmc_spv.nim:826-850 + smeared_one_link_force :713-753): outer products, rephase, odd sites
    negated, smeared force in place, projTAH against the thin links -- every step on the device."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    g = lat.random_gauge(lat.new_rng(SEED), 4)
    psi = lat.gaussian_vector(lat.new_rng(3))
    # oracle
    want = lat.gauge(4)
    lat.outer(want, psi, psi, 0.7, 0.7, accumulate=False)
    lat.setBC(want); lat.stagPhase(want)
    want[:, lat.half:] *= -1
    want = lat.hyp_force(g, ALPHA, want)
    m = (want[..., 0] + 1j * want[..., 1]) @ np.conj(g[..., 0] + 1j * g[..., 1]).transpose(0, 1, 3, 2)
    t = 0.5 * (m - np.conj(m).transpose(0, 1, 3, 2))
    t = t - np.einsum("kxaa->kx", t)[..., None, None] / 3.0 * np.eye(3)
    want = np.stack([t.real, t.imag], axis=-1)
    # device
    dg, fl, f = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    force = qb.HypCoefs(*ALPHA).smearGetForce(dg, fl)
    dpsi = qb.DeviceField(psi)
    f.outer(dpsi, dpsi, 0.7, -0.7, accumulate=False)      # the odd-site sign rides on the weight
    f.rephase()
    force(f, f)
    f.projTAH(dg)
    check(down(lat, lo, f), want, lat.vol)
    force.free()
    for d in (dg, fl, f, dpsi):
        d.free()


def test_fermion_force_step_of_staghmc_spv(ctx_factory):
    """One massive-fermion force of src/stagg_pv_hmc/staghmc_spv.nim on the device, nothing but phi uploaded and the
    force downloaded: smear (:367-386 smearGetForce, setBC, stagPhase) -> solve_fermion (:406-458, D psi = phi) ->
    psi (x) psi(+mu)^+ with rescale (:697-711, :826-850) -> rephase, odd sites negated (:713-728) -> smeared force
    in place (:745) -> projTAH against the thin links (:752).  Against the same chain on the oracle."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, lo = oracle.Lattice(L), oracle.QexLayout(L)
    g = lat.random_gauge(lat.new_rng(SEED), 4)
    phi = lat.gaussian_vector(lat.new_rng(17))
    mass, dt, r2req = 0.1, 0.05, 1e-20
    scale = -0.5 * dt / mass
    # oracle
    sg = lat.hyp_smear(g, ALPHA)
    lat.setBC(sg); lat.stagPhase(sg)
    psi = lat.vector()
    st = lat.solve(sg, psi, phi, mass, r2req)
    want = lat.gauge(4)
    lat.outer(want, psi, psi, scale, scale, accumulate=False)
    lat.setBC(want); lat.stagPhase(want)
    want[:, lat.half:] *= -1
    want = lat.hyp_force(g, ALPHA, want)
    m = (want[..., 0] + 1j * want[..., 1]) @ np.conj(g[..., 0] + 1j * g[..., 1]).transpose(0, 1, 3, 2)
    t = 0.5 * (m - np.conj(m).transpose(0, 1, 3, 2))
    t = t - np.einsum("kxaa->kx", t)[..., None, None] / 3.0 * np.eye(3)
    want = np.stack([t.real, t.imag], axis=-1)
    # device
    dg, sgd, f = up(qb, lo, g), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    force = qb.HypCoefs(*ALPHA).smearGetForce(dg, sgd)
    sgd.rephase()
    s = qb.newStag(sgd)
    dphi, dpsi = qb.DeviceField(phi), qb.DeviceField()
    sp = qb.SolverParams(r2req=r2req)
    s.dev_solve(dpsi, dphi, mass, sp)
    assert abs(sp.iterations - st.iterations) <= 2
    f.outer(dpsi, dpsi, scale, -scale, accumulate=False)
    f.rephase()
    force(f, f)
    f.projTAH(dg)
    got = down(lat, lo, f)
    assert rel_l2(got, want) <= 1e-8, rel_l2(got, want)      # two CG solutions at r2req = 1e-20
    ah = got[..., 0] + 1j * got[..., 1]
    assert np.abs(ah + np.conj(ah).transpose(0, 1, 3, 2)).max() < 1e-13     # anti-Hermitian
    assert np.abs(np.einsum("kxaa->kx", ah)).max() < 1e-13                  # traceless
    force.free()
    for d in (s, dg, sgd, f, dphi, dpsi):
        d.free()
