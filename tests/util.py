"""Shared helpers of the parity tests: seeded inputs from the oracle, layout conversion."""
import numpy as np

import oracle

SEED = 987654321  # the reference benchmarks' seed (src/physics/stagD.nim:779)


def make_inputs(L, nlinks=4, seed=SEED, nvec=2, long_scale=0.1):
    """Random SU(3) links with antiperiodic-T BC and staggered phases (stagD.nim:72-80), plus
    gaussian sources, exactly as QEX's own RNG would produce them.  nlinks=8 adds long links =
    long_scale * SU(3) from the continuing streams (the synthetic 3-hop operator of
    src/bench/benchStagProp.nim:84-92), interleaved like newStag3 (stagD.nim:552-564)."""
    lat = oracle.Lattice(L)
    rng = lat.new_rng(seed)
    g = lat.random_gauge(rng, 4)
    lat.setBC(g)
    lat.stagPhase(g)
    if nlinks == 8:
        g3 = lat.random_gauge(rng, 4) * long_scale
        gg = np.zeros((8,) + g.shape[1:])
        gg[0::2] = g
        gg[1::2] = g3
        g = gg
    vecs = [lat.gaussian_vector(rng) for _ in range(nvec)]
    return lat, np.ascontiguousarray(g), vecs


def make_exact_inputs(L, nlinks=4, seed=5, nvec=2, long_scale=0.1):
    """Like make_inputs, but the SU(3) links come from a Gram-Schmidt in numpy and are unitary to
    machine precision (QEX's projectSU goes through a closed-form eigen-decomposition and leaves
    ~1e-11, src/maths/matrixFunctions.nim:126-181) -- the inputs the 12-real link storage is for."""
    from qex_b200.synth import random_su3_links, gaussian_vector
    lat = oracle.Lattice(L)
    g = np.ascontiguousarray(np.stack(random_su3_links(lat.vol, 4, seed=seed)))
    lat.setBC(g)
    lat.stagPhase(g)
    if nlinks == 8:
        g3 = np.stack(random_su3_links(lat.vol, 4, seed=seed + 1, scale=[long_scale] * 4))
        gg = np.zeros((8,) + g.shape[1:])
        gg[0::2] = g
        gg[1::2] = g3
        g = gg
    vecs = [gaussian_vector(lat.vol, seed + i) for i in range(nvec)]
    return lat, np.ascontiguousarray(g), vecs


def rel_l2(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def to_qex(lo, f):
    """oracle global field -> native QEX array of rank 0 of layout lo (single rank)."""
    return lo.scatter(f)[0]


def from_qex(lo, a):
    return lo.gather([a])


def hisq_reference(lat, g, coefs):
    """The oracle's HISQ links plus the SAME chain with the projection done exactly (polar factor from numpy's SVD,
    good to ~1e-15).  QEX's projectU takes the closed-form eigenvalue route (matrixFunctions.nim:79-181), which on
    random links is only ~1e-13 from the exact polar factor in L2 with a heavy tail (single sites at 1e-11 where
    U^+U is nearly singular); two correct implementations of it therefore differ by a few times that distance.
    Returns (fl, ll, fl_exact, ll_exact)."""
    fl, ll = lat.hisq_smear(g, coefs)
    t1 = lat.makeImpLinks(g, coefs[:5])
    m = t1[..., 0] + 1j * t1[..., 1]
    u, _, vh = np.linalg.svd(m)
    w = u @ vh
    p = np.ascontiguousarray(np.stack([w.real, w.imag], axis=-1))
    fle, lle = lat.makeImpLinks(p, coefs[5:10], gfLong=p, naik=float(coefs[10]))
    return fl, ll, fle, lle


def hisq_bar(want, exact, base=1e-13, factor=10.0):
    """Parity bar for HISQ links: `factor` times the oracle's own distance from the exact-projection result taken
    over the whole lattice and all four directions (the error sits in a handful of nearly singular sites, so the
    statistic of a single direction or of one rank's block fluctuates by several x), never below `base`."""
    return max(base, factor * rel_l2(want, exact))


# ---- the reference-held pin for D / Ddag conventions / solve: the first "Begin H" of staghmc_sh --------------------
class OracleOps:
    """The operations `staghmc_begin_H` needs, on the oracle."""

    def __init__(self, lat):
        self.lat = lat

    def smear_rephase(self, g, alpha):
        sg = self.lat.hyp_smear(g, alpha)
        self.lat.setBC(sg)
        self.lat.stagPhase(sg)
        return sg

    def D(self, sg, r, x, m):
        self.lat.D(sg, r, x, m)

    def solve(self, sg, x, b, m, r2req, maxits):
        st = self.lat.solve(sg, x, b, m, r2req, maxits)
        return st.iterations, st.r2

    def norm2(self, x):
        return self.lat.norm2(x)


def staghmc_begin_H(lat, ops, mass, hmasses, arsq=1e-24, seed=SEED, maxits=1000000, alpha=(0.4, 0.5, 0.5)):
    """Everything src/examples/staghmc_sh.nim does up to its first `Begin H` line (:716-762), which happens before
    any molecular-dynamics step: per-site RngMilc6 field (:178) -> p.randomTAH r (:722) -> unit gauge (:707),
    nHYP smear + setBC + stagPhase (:728, :309-319) -> psi[k][i].gaussian r in the interleaved order of :735-745
    -> phi = D(m2)^-1 D(m1) psi with the `-m` convention and phi.odd := 0 (:747-757) -> faction (:344-370)
    -> fa = 0.5 |psi|^2, T = 0.5 p^2 - 16 vol (:334-342).
    Inputs (RNG streams) come from the oracle; `ops` supplies smear_rephase / D / solve / norm2, so the same
    sequence runs on the oracle and through the C ABI on the GPU.
    Returns (Sf[k][i], T, log) with log = [(what, mass, iterations, r2)] of every solve in call order."""
    r = lat.new_rng(seed)
    p = lat.random_tah(r)
    p2 = 0.0
    for mu in range(4):
        p2 += float(np.square(p[mu]).sum())
    T = 0.5 * p2 - float(16 * lat.vol)
    sg = ops.smear_rephase(lat.unit_gauge(), alpha)
    nk = len(mass)
    psi = [[None] * (len(hmasses[k]) + 1) for k in range(nk)]
    phi = [[None] * (len(hmasses[k]) + 1) for k in range(nk)]
    i, running = 0, True
    while running:
        running = False
        for k in range(nk):
            if i >= len(psi[k]):
                continue
            psi[k][i] = lat.gaussian_vector(r)
            running = True
        i += 1
    log = []
    for k in range(nk):
        hm, n = hmasses[k], len(hmasses[k]) + 1
        for i in range(n):
            phi[k][i] = lat.vector()
            if i != n - 1:
                ftmp = lat.vector()
                ops.D(sg, ftmp, psi[k][i], -mass[k] if i == 0 else -hm[i - 1])
                it, r2 = ops.solve(sg, phi[k][i], ftmp, -hm[i], arsq, maxits)
                log.append(("init", k, hm[i], it, r2))
            else:
                ops.D(sg, phi[k][i], psi[k][i], -hm[i - 1] if len(hm) > 0 else -mass[k])
            phi[k][i][lat.half:] = 0.0
    Sf = []
    for k in range(nk):
        hm, n = hmasses[k], len(hmasses[k]) + 1
        for i in range(n - 1):
            ftmp = lat.vector()
            ops.D(sg, ftmp, phi[k][i], hm[i])
            m = mass[k] if i == 0 else hm[i - 1]
            it, r2 = ops.solve(sg, psi[k][i], ftmp, m, arsq, maxits)
            log.append(("action", k, m, it, r2))
        m = hm[-1] if len(hm) > 0 else mass[k]
        it, r2 = ops.solve(sg, psi[k][n - 1], phi[k][n - 1], m, arsq, maxits)
        log.append(("action", k, m, it, r2))
        Sf.append([0.5 * ops.norm2(psi[k][i]) for i in range(n)])
    return Sf, T, log


def check_begin_H(run, Sf, T, log, rel_tol):
    """Holds (Sf, T, log) of staghmc_begin_H to one run record of tests/golden/staghmc_sh.json: Sf and T at the
    reference's own relative bar (tests/extra/staghmc_sh/run:43, diffnum 2e-11), and the CG iteration counts to the
    `calls: (iterations div calls):max` statistics the log prints for trajectory 1 (solverBase.nim:84-112): with c
    calls of which c-1 happen before `Begin H`... the last call (at `End H`) is the maximum, so
    its_div_calls == (sum(begin iterations) + its_max) div calls must hold for the begin iterations within +-1."""
    worst = abs(T - run["T"]) / abs(run["T"])
    for k, row in enumerate(run["Sf"]):
        for i, want in enumerate(row):
            worst = max(worst, abs(Sf[k][i] - want) / abs(want))
    assert worst <= rel_tol, (worst, Sf, T)
    # solver statistics: spa[k] serves the action solve at mass[k]; spah[k][i] serves the init solve AND the action solve
    stats = list(run["action_stats"])
    pos = 0
    for k in range(len(run["mass"])):
        ms = [run["mass"][k]] + list(run["hmasses"][k])
        for j, m in enumerate(ms):
            st = stats[pos]
            pos += 1
            assert abs(st["m"] - m) < 1e-12
            mine = [e[3] for e in log if e[1] == k and abs(e[2] - m) < 1e-12]
            assert len(mine) == st["calls"] - 1, (mine, st)
            ok = any((sum(mine) + d + st["its_max"]) // st["calls"] == st["its_div_calls"] for d in (-1, 0, 1))
            assert ok, (mine, st)
    return worst
