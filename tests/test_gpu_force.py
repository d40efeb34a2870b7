"""GPU parity of the fermion-force outer products (src/physics/stagD.nim:588-664,
src/stagg_pv_hmc/staghmc_spv.nim:826-850) through the C ABI against the oracle.  Bar: rel-L2 <= 1e-13."""
import numpy as np
import pytest

import oracle
from util import make_inputs, rel_l2

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


CASES = [((8, 8, 8, 8), 4, (1, 1, 1, 1)), ((4, 6, 8, 10), 8, (1, 1, 1, 1)), ((8, 8, 8, 8), 4, (1, 2, 2, 2)),
         ((8, 8, 8, 8), 8, (1, 2, 2, 2))]


def down(lat, lo, dg):
    parts = [lo.scatter(lat.gauge(1)[0])[0] for _ in range(dg.nlinks)]
    dg.download(parts)
    return np.stack([lo.gather([p]) for p in parts])


@pytest.mark.parametrize("L,nlinks,inner", CASES)
def test_stagD2deriv_parity(ctx_factory, L, nlinks, inner):
    qb = ctx_factory(L, inner)
    lat, g, (x, c) = make_inputs(L, nlinks)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    s = qb.Staggered([lo.scatter(g[k])[0] for k in range(nlinks)])
    start = 0.25 * g            # stagD2deriv accumulates: start from something non-zero
    want = start.copy()
    lat.stagD2deriv(want, c, x)
    # host-field entry point
    hf = [lo.scatter(start[k])[0] for k in range(nlinks)]
    qb.stagD2deriv(s, hf, lo.scatter(c)[0], lo.scatter(x)[0])
    for k in range(nlinks):
        assert rel_l2(lo.gather([hf[k]]), want[k]) <= TOL, k
    # device-resident entry point
    gr = qb.DeviceGauge([lo.scatter(start[k])[0] for k in range(nlinks)])
    dc, dx = qb.DeviceField(lo.scatter(c)[0]), qb.DeviceField(lo.scatter(x)[0])
    qb.stagD2deriv(s, gr, dc, dx)
    got = down(lat, lo, gr)
    for k in range(nlinks):
        assert rel_l2(got[k], want[k]) <= TOL, k
    for d in (s, gr, dc, dx):
        d.free()


@pytest.mark.parametrize("L,inner", [((8, 8, 8, 8), (1, 1, 1, 1)), ((4, 6, 8, 10), (1, 1, 1, 1)), ((8, 8, 8, 8), (1, 2, 2, 2))])
def test_stagDeriv_parity(ctx_factory, L, inner):
    qb = ctx_factory(L, inner)
    lat, g, (x, _) = make_inputs(L, 4)
    lo = oracle.QexLayout(L, (1, 1, 1, 1), inner)
    s = qb.Staggered([lo.scatter(g[k])[0] for k in range(4)])
    want = lat.gauge(4)
    lat.stagDeriv(want, x)
    f = qb.DeviceGauge(nlinks=4)
    f.zero()
    dx = qb.DeviceField(lo.scatter(x)[0])
    qb.stagDeriv(s, f, dx)
    got = down(lat, lo, f)
    for k in range(4):
        assert rel_l2(got[k], want[k]) <= TOL, k
    hf = [lo.scatter(lat.gauge(1)[0])[0] for _ in range(4)]
    qb.stagDeriv(s, hf, lo.scatter(x)[0])
    for k in range(4):
        assert rel_l2(lo.gather([hf[k]]), want[k]) <= TOL, k
    for d in (s, f, dx):
        d.free()


def test_stagD2deriv_is_the_gradient_of_the_operator(ctx_factory):
    """(d/dg) redot[(2D) x, c] by central differences of the DEVICE operator = the device stagD2deriv."""
    L = (4, 4, 4, 4)
    qb = ctx_factory(L)
    lat, g, (x, c) = make_inputs(L, 4)
    lo = oracle.QexLayout(L)
    dg = np.random.default_rng(0).standard_normal(g.shape)

    def action(gg):
        s = qb.Staggered([gg[k] for k in range(4)])
        r = lat.vector()
        qb.stagD2(s, 0, r, x, 0.0, 0.0)
        qb.stagD2(s, 1, r, x, 0.0, 0.0)
        s.free()
        return float(np.sum(r * c))
    eps = 1e-6
    fd = (action(g + eps * dg) - action(g - eps * dg)) / (2 * eps)
    s = qb.Staggered([g[k] for k in range(4)])
    grad = [lat.gauge(1)[0] for _ in range(4)]
    qb.stagD2deriv(s, grad, c, x)
    an = float(sum(np.sum(grad[k] * dg[k]) for k in range(4)))
    assert abs(fd - an) <= 1e-7 * abs(an)
    s.free()


def test_force_loop_outer_products(ctx_factory):
    """f[mu] := s0 psi0 (x) psi0(+mu)^+ ; f[mu] += s1 psi1 (x) psi1(+mu)^+ (staghmc_spv.nim:826-850), then the
    rephase + odd-site sign of smeared_one_link_force (:713-728)."""
    L = (8, 8, 8, 8)
    qb = ctx_factory(L)
    lat, g, (p0, p1) = make_inputs(L, 4)
    lo = oracle.QexLayout(L)
    want = lat.gauge(4)
    lat.outer(want, p0, p0, 0.7, 0.7, accumulate=False)
    lat.outer(want, p1, p1, -0.35, -0.35, accumulate=True)
    lat.setBC(want); lat.stagPhase(want)
    want[:, lat.half:] *= -1
    f = qb.DeviceGauge(nlinks=4)
    d0, d1 = qb.DeviceField(p0), qb.DeviceField(p1)
    f.outer(d0, d0, 0.7, 0.7, accumulate=False)          # overwrites whatever f held
    f.outer(d1, d1, -0.35, -0.35, accumulate=True)
    f.rephase()
    got = down(lat, lo, f)
    got[:, lat.half:] *= -1
    for k in range(4):
        assert rel_l2(got[k], want[k]) <= TOL
    # odd-site sign folded into the weights instead: same result
    f.outer(d0, d0, 0.7, -0.7, accumulate=False)
    f.outer(d1, d1, -0.35, 0.35, accumulate=True)
    f.rephase()
    got = down(lat, lo, f)
    for k in range(4):
        assert rel_l2(got[k], want[k]) <= TOL
    for d in (f, d0, d1):
        d.free()
