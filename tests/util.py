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
