"""GPU parity against numbers the REFERENCE itself holds (tests/extra/staghmc_sh/ref.{0,1,2}) and against the oracle on
BASELINE.json configs[1] itself (32^4), all through the C ABI."""
import json
import os

import numpy as np
import pytest

from util import make_inputs, make_exact_inputs, rel_l2, staghmc_begin_H, check_begin_H

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture()
def ctx(qb):
    def make(L):
        if qb.is_initialized():
            qb.finalize()
        qb.init(L)
        return qb
    yield make
    if qb.is_initialized():
        qb.finalize()


class DeviceOps:
    """The operations of util.staghmc_begin_H through the C ABI: nHYP smearing + setBC + stagPhase on the device
    (HypCoefs.smear, rephase), Staggered.D / Staggered.solve on host fields, norm2 on a device field."""

    def __init__(self, qb, lat):
        self.qb, self.lat, self.keep = qb, lat, []

    def smear_rephase(self, g, alpha):
        qb = self.qb
        thin = qb.DeviceGauge([g[k] for k in range(4)])
        sg = qb.DeviceGauge(nlinks=4)
        qb.HypCoefs(*alpha).smear(thin, sg)
        sg.rephase()
        self.keep += [thin, sg]
        return qb.newStag(sg)

    def D(self, s, r, x, m):
        s.D(r, x, m)

    def solve(self, s, x, b, m, r2req, maxits):
        sp = self.qb.SolverParams(r2req=r2req, maxits=maxits)
        s.solve(x, b, m, sp)
        return sp.iterations, sp.r2

    def norm2(self, x):
        f = self.qb.DeviceField(x)
        v = f.norm2()
        f.free()
        return v

    def free(self):
        for d in self.keep:
            d.free()


@pytest.mark.parametrize("tid", ["0", "1", "2"])
def test_staghmc_sh_begin_H_on_device(orc, ctx, tid):
    """The first `Begin H` of the reference's golden HMC logs (Sf per pseudofermion; QEX's own bar rel 2e-11,
    tests/extra/staghmc_sh/run:43) reproduced by the device path: smear + rephase, D with negative masses, five or
    ten full `solve`s at r2req = 1e-24 with their reconstruction, norm2.  Iteration counts are held to the
    `Solver[action]` statistics of the same log (+-1) and to the oracle's (+-1 per solve)."""
    fx = json.load(open(os.path.join(GOLD, "staghmc_sh.json")))
    run = fx["runs"][tid]
    qb = ctx(fx["lattice"])
    lat = orc.Lattice(fx["lattice"])
    ops = DeviceOps(qb, lat)
    Sf, T, log = staghmc_begin_H(lat, ops, run["mass"], run["hmasses"], arsq=fx["arsq"], seed=fx["seed"])
    worst = check_begin_H(run, Sf, T, log, fx["rel_tol"])
    assert worst <= 1e-12, worst        # observed: a few ulp
    from util import OracleOps
    _, _, olog = staghmc_begin_H(lat, OracleOps(lat), run["mass"], run["hmasses"], arsq=fx["arsq"], seed=fx["seed"])
    for a, b in zip(log, olog):
        assert a[:3] == b[:3] and abs(a[3] - b[3]) <= 1, (a, b)
        assert a[4] <= fx["arsq"]
    ops.free()


def _config1_case(orc, qb, lat, g, b, recon):
    m, r2req = 0.1, 1e-8
    s = qb.Staggered([g[k] for k in range(4)])
    if recon != 18:
        got, dev = s.recon(recon)
        assert got[0] == recon, (got, dev)
    # Dslash at the full size
    want, out = lat.vector(), lat.vector()
    lat.D(g, want, b, m)
    s.D(out, b, m)
    assert rel_l2(out, want) <= 1e-13
    # the even-even solve that bench.py times
    xe_o = lat.vector()
    it_o, r2_o = lat.solveXX(g, xe_o, b, m, 0, r2req, 10000)
    xe = np.ones_like(b)
    sp = qb.SolverParams(r2req=r2req, maxits=10000)
    s.solveEE(xe, b, m, sp)
    assert abs(sp.iterations - it_o) <= 1, (sp.iterations, it_o)
    assert rel_l2(xe[:lat.half], xe_o[:lat.half]) <= 10 * np.sqrt(r2req)
    Ax = lat.vector()
    lat.stagD2xx(g, Ax, xe, m * m, 0)
    tr = float(((Ax - b)[:lat.half] ** 2).sum() / (b[:lat.half] ** 2).sum())
    assert tr <= r2req, tr
    # the full solve D x = b
    x_o, x = lat.vector(), lat.vector()
    st = lat.solve(g, x_o, b, m, r2req, 10000)
    sp2 = qb.SolverParams(r2req=r2req, maxits=10000)
    s.solve(x, b, m, sp2)
    assert abs(sp2.iterations - st.iterations) <= max(1, st.calls), (sp2.iterations, st.iterations)
    assert rel_l2(x, x_o) <= 10 * np.sqrt(r2req)
    Dx = lat.vector()
    lat.D(g, Dx, x, m)
    tr2 = lat.norm2(Dx - b) / lat.norm2(b)
    assert tr2 <= r2req, tr2
    s.free()
    return sp.iterations, it_o, tr, sp2.iterations, st.iterations, tr2


def test_config1_32x4_parity_links18(orc, ctx):
    """BASELINE.json configs[1] ITSELF -- 32^4 naive, QEX-generated random SU(3) links (RngMilc6 seed 987654321,
    projectSU, setBC, stagPhase), gaussian source, m = 0.1, r2req = 1e-8 -- GPU against the oracle: Dslash rel-L2
    <= 1e-13, CG iterations +-1, solution rel-L2 <= 10 sqrt(r2req), true residual <= r2req.  18-real links (what
    the library selects for QEX's own links)."""
    L = (32, 32, 32, 32)
    qb = ctx(L)
    lat, g, (b, _) = make_inputs(L, 4)
    print(_config1_case(orc, qb, lat, g, b, 18))


def test_config1_32x4_parity_links12(orc, ctx):
    """The same on links that are SU(3) to machine precision, stored as 12 reals."""
    L = (32, 32, 32, 32)
    qb = ctx(L)
    lat, g, (b, _) = make_exact_inputs(L, 4)
    print(_config1_case(orc, qb, lat, g, b, 12))
