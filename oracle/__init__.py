"""CPU oracle for the QEX staggered Dslash + CG path -- TEST INFRASTRUCTURE ONLY.

ctypes loader for ``oracle/libqex_oracle.so`` (built by ``oracle/Makefile`` from
``qex_oracle.cpp``, a C++/OpenMP restatement of the reference arithmetic; every
function there cites the reference file:line it follows).  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` may import this package.  The product package ``qex_b200`` never
does.

Field layout == QEX single-rank V=1 layout: site ``i = lex(x)//2`` (+vol/2 for
odd sites), colour vector ``float64[vol,3,2]``, link ``float64[vol,3,3,2]``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libqex_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "qex_oracle.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


class OrcLayout(C.Structure):
    _fields_ = [("nd", C.c_int), ("phys", C.c_int * 4), ("rankg", C.c_int * 4), ("inner", C.c_int * 4),
                ("local", C.c_int * 4), ("outer", C.c_int * 4), ("nSites", C.c_int), ("nEven", C.c_int),
                ("nOdd", C.c_int), ("nSitesOuter", C.c_int), ("nEvenOuter", C.c_int), ("nOddOuter", C.c_int),
                ("nSitesInner", C.c_int), ("innerCb", C.c_int), ("innerCbDir", C.c_int)]


class SolveStats(C.Structure):
    _fields_ = [("iterations", C.c_int), ("calls", C.c_int), ("r2", C.c_double), ("b2", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        L = _lib
        dp, vp, ip = C.POINTER(C.c_double), C.c_void_p, C.POINTER(C.c_int)
        L.orc_lat_new.restype = vp
        L.orc_lat_new.argtypes = [ip]
        L.orc_lat_free.argtypes = [vp]
        L.orc_lat_vol.argtypes = [vp]
        L.orc_lat_coords.argtypes = [vp, C.POINTER(C.c_short)]
        L.orc_rng_new.restype = vp
        L.orc_rng_new.argtypes = [vp, C.c_uint64]
        L.orc_rng_free.argtypes = [vp]
        L.orc_rng_next_site.restype = C.c_uint32
        L.orc_rng_next_site.argtypes = [vp, C.c_int]
        L.orc_gaussian.argtypes = [vp, vp, dp, C.c_int]
        L.orc_random_gauge.argtypes = [vp, vp, dp, C.c_int]
        L.orc_random_tah.argtypes = [vp, vp, dp]
        L.orc_plaq.argtypes = [vp, dp, dp]
        L.orc_setbc.argtypes = [vp, dp]
        L.orc_stagphase.argtypes = [vp, dp]
        L.orc_projectU.argtypes = [dp, dp]
        L.orc_projectSU.argtypes = [dp, dp]
        L.orc_stagD2.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_double, C.c_int]
        L.orc_stagD.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_double, C.c_double, C.c_int]
        L.orc_D.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double]
        L.orc_Ddag.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double]
        L.orc_stagD2xx.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_int, dp]
        L.orc_eoReconstruct.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double]
        L.orc_norm2.restype = C.c_double
        L.orc_norm2.argtypes = [vp, dp, C.c_int]
        L.orc_redot.restype = C.c_double
        L.orc_redot.argtypes = [vp, dp, dp, C.c_int]
        L.orc_cg_xx.restype = C.c_int
        L.orc_cg_xx.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_int, C.c_double, C.c_int, dp, dp]
        L.orc_solveXX.restype = C.c_int
        L.orc_solveXX.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_int, C.c_double, C.c_int, dp]
        L.orc_solve.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_double, C.c_int, C.c_int,
                                C.POINTER(SolveStats)]
        L.orc_cgm_xx.restype = C.c_int
        L.orc_cgm_xx.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, dp, C.c_int, C.c_int, C.c_double, C.c_int, dp, dp]
        L.orc_solve_multi.restype = C.c_int
        L.orc_solve_multi.argtypes = [vp, C.c_int, dp, dp, dp, dp, C.c_int, C.c_double, C.c_int, C.POINTER(SolveStats)]
        L.orc_makeImpLinks.argtypes = [vp, dp, dp, dp, dp, C.c_double, dp]
        L.orc_projectU_field.argtypes = [vp, dp, dp, C.c_int]
        L.orc_hisq_smear.argtypes = [vp, dp, dp, dp, dp]
        L.orc_hyp_smear.argtypes = [vp, dp, dp, dp]
        L.orc_fat7l_deriv.argtypes = [vp, dp, dp, dp, dp, dp, dp, dp, C.c_double]
        L.orc_hisq_force.argtypes = [vp, dp, dp, dp, dp, dp]
        L.orc_sym_staple_deriv.argtypes = [vp, dp, dp, dp, C.c_int, dp, C.c_int, dp]
        L.orc_hyp_force.argtypes = [vp, dp, dp, dp, dp, dp]
        L.orc_projectUderiv.argtypes = [dp, dp, dp]
        L.orc_outer.argtypes = [vp, C.c_int, dp, dp, dp, C.c_double, C.c_double, C.c_int]
        L.orc_stagDeriv.argtypes = [vp, C.c_int, dp, dp]
        L.orc_stagD2deriv.argtypes = [vp, C.c_int, dp, dp, dp]
        L.orc_layout_setup.argtypes = [C.POINTER(OrcLayout), ip, ip, ip]
        L.orc_layout_index.argtypes = [C.POINTER(OrcLayout), ip, ip, ip]
        L.orc_layout_coord.argtypes = [C.POINTER(OrcLayout), C.c_int, C.c_int, ip]
        L.orc_layout_table.argtypes = [C.POINTER(OrcLayout), ip, ip]
        L.orc_set_gauss_variant.argtypes = [C.c_int]
        L.orc_set_num_threads.argtypes = [C.c_int]
    return _lib


def _dp(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _i4(v):
    return (C.c_int * 4)(*[int(t) for t in v])


EVEN, ODD, ALL = 0, 1, 2


class Lattice:
    """Single-rank V=1 QEX lattice + the reference operations restated on it."""

    def __init__(self, L):
        self.L = tuple(int(t) for t in L)
        self.h = lib().orc_lat_new(_i4(self.L))
        self.vol = lib().orc_lat_vol(self.h)
        self.half = self.vol // 2
        self._coords = None

    def __del__(self):
        try:
            lib().orc_lat_free(self.h)
        except Exception:
            pass

    @property
    def coords(self):
        """int16[4, vol]: coords[d][i] like Layout.coords (layoutX.nim:160-169)."""
        if self._coords is None:
            c = np.zeros((4, self.vol), dtype=np.int16)
            lib().orc_lat_coords(self.h, c.ctypes.data_as(C.POINTER(C.c_short)))
            self._coords = c
        return self._coords

    # -- inputs ------------------------------------------------------------
    def new_rng(self, seed=17 ** 7):
        return RngField(self, seed)

    def vector(self):
        return np.zeros((self.vol, 3, 2), dtype=np.float64)

    def gauge(self, n=4):
        return np.zeros((n, self.vol, 3, 3, 2), dtype=np.float64)

    def random_gauge(self, rng, n=4):
        g = self.gauge(n)
        lib().orc_random_gauge(self.h, rng.h, _dp(g), n)
        return g

    def random_tah(self, rng, n=4):
        """randomTAH(g, r) (gaugeUtils.nim:1353-1383,1452-1454): field by field, 8 deviates per site each."""
        p = self.gauge(n)
        for mu in range(n):
            lib().orc_random_tah(self.h, rng.h, _dp(p[mu]))
        return p

    def unit_gauge(self, n=4):
        g = self.gauge(n)
        for a in range(3):
            g[:, :, a, a, 0] = 1.0
        return g

    def gaussian_vector(self, rng):
        v = self.vector()
        lib().orc_gaussian(self.h, rng.h, _dp(v), 6)
        return v

    def plaq(self, g):
        pl = np.zeros(6)
        lib().orc_plaq(self.h, _dp(g), _dp(pl))
        return pl

    def setBC(self, g):
        lib().orc_setbc(self.h, _dp(g))

    def stagPhase(self, g):
        lib().orc_stagphase(self.h, _dp(g))

    # -- link smearing -----------------------------------------------------
    def makeImpLinks(self, gf, coef, gfLong=None, naik=0.0):
        """makeImpLinks (fat7l.nim:78-162). coef = (oneLink, threeStaple, fiveStaple, sevenStaple, lepage).
        Returns fl, or (fl, ll) when naik != 0."""
        c = np.ascontiguousarray(coef, dtype=np.float64)
        fl = self.gauge(4)
        if naik != 0.0:
            ll = self.gauge(4)
            lib().orc_makeImpLinks(self.h, _dp(gf), _dp(c), _dp(fl), _dp(gfLong), naik, _dp(ll))
            return fl, ll
        lib().orc_makeImpLinks(self.h, _dp(gf), _dp(c), _dp(fl), None, 0.0, None)
        return fl

    def projectU(self, g):
        out = np.zeros_like(g)
        lib().orc_projectU_field(self.h, _dp(g), _dp(out), g.shape[0])
        return out

    def hisq_smear(self, g, coefs):
        """HisqCoefs.smear (hisqLinks.nim:32-39); coefs = fat7first[5] + fat7second[5] + [naik]."""
        c = np.ascontiguousarray(coefs, dtype=np.float64)
        assert c.shape == (11,)
        fl, ll = self.gauge(4), self.gauge(4)
        lib().orc_hisq_smear(self.h, _dp(g), _dp(c), _dp(fl), _dp(ll))
        return fl, ll

    def hyp_smear(self, g, alpha):
        """HypCoefs.smear (hypsmear.nim:49-162); alpha = (alpha1, alpha2, alpha3)."""
        a = np.ascontiguousarray(alpha, dtype=np.float64)
        fl = self.gauge(4)
        lib().orc_hyp_smear(self.h, _dp(g), _dp(a), _dp(fl))
        return fl

    def fat7lDeriv(self, deriv, gauge, mid, coef, llderiv=None, llgauge=None, llmid=None, naik=0.0):
        """fat7lDeriv (fat7lderiv.nim:122-335); deriv (and llderiv) accumulate."""
        c = np.ascontiguousarray(coef, dtype=np.float64)
        if naik != 0.0:
            lib().orc_fat7l_deriv(self.h, _dp(deriv), _dp(gauge), _dp(mid), _dp(c), _dp(llderiv), _dp(llgauge), _dp(llmid), naik)
        else:
            lib().orc_fat7l_deriv(self.h, _dp(deriv), _dp(gauge), _dp(mid), _dp(c), None, None, None, 0.0)

    def hisq_force(self, u, coefs, dsdsu, dsdsul, dsdu=None):
        """HisqCoefs.smearGetForce -> smearedForce(dsdu, dsdsu, dsdsul) (hisqsmear.nim:51-84); dsdu accumulates."""
        c = np.ascontiguousarray(coefs, dtype=np.float64)
        if dsdu is None:
            dsdu = self.gauge(4)
        lib().orc_hisq_force(self.h, _dp(u), _dp(c), _dp(dsdsu), _dp(dsdsul), _dp(dsdu))
        return dsdu

    def sym_staple_deriv(self, f1, f2, g1, d1, g2, d2, c):
        """symStapleDeriv (smearutil.nim:24-54): f1, f2 accumulate the derivatives wrt g1 (dir d1), g2 (dir d2)."""
        lib().orc_sym_staple_deriv(self.h, _dp(f1), _dp(f2), _dp(g1), d1, _dp(g2), d2, _dp(c))

    def projectUderiv(self, x, chain):
        """projectUderiv(r, x, c) site by site (matrixFunctions.nim:321-357); x, chain: [n, 3, 3, 2]."""
        out = np.zeros_like(x)
        xf, cf, of = x.reshape(-1, 18), chain.reshape(-1, 18), out.reshape(-1, 18)
        for i in range(xf.shape[0]):
            lib().orc_projectUderiv(_dp(xf[i]), _dp(cf[i]), _dp(of[i]))
        return out

    def hyp_force(self, g, alpha, chain, want_fl=False):
        """smearGetForce -> smearedForce(f, chain) (hypsmear.nim:49-262): the chain rule through the three nHYP levels."""
        a = np.ascontiguousarray(alpha, dtype=np.float64)
        f = self.gauge(4)
        fl = self.gauge(4) if want_fl else None
        lib().orc_hyp_force(self.h, _dp(g), _dp(a), _dp(chain), _dp(f), _dp(fl) if want_fl else None)
        return (f, fl) if want_fl else f

    # -- fermion-force outer products --------------------------------------
    def outer(self, f, a, b, ce=1.0, co=1.0, accumulate=True):
        """f[k](x) (+)= (ce|co) a(x) (x) b(x+h_k)^+  (staghmc_spv.nim:833-850)."""
        lib().orc_outer(self.h, f.shape[0], _dp(f), _dp(a), _dp(b), ce, co, int(accumulate))

    def stagDeriv(self, f, x):
        """stagDeriv (stagD.nim:635-664)."""
        lib().orc_stagDeriv(self.h, f.shape[0], _dp(f), _dp(x))

    def stagD2deriv(self, g, c, x):
        """stagD2deriv (stagD.nim:588-633)."""
        lib().orc_stagD2deriv(self.h, g.shape[0], _dp(g), _dp(c), _dp(x))

    # -- operator ----------------------------------------------------------
    def stagD2(self, g, r, x, a, b, parity):
        lib().orc_stagD2(self.h, g.shape[0], _dp(g), _dp(r), _dp(x), a, b, parity)

    def stagD(self, g, r, x, m, parity, sc=1.0, a=0.0):
        lib().orc_stagD(self.h, g.shape[0], _dp(g), _dp(r), _dp(x), m, sc, a, parity)

    def D(self, g, r, x, m):
        lib().orc_D(self.h, g.shape[0], _dp(g), _dp(r), _dp(x), m)

    def Ddag(self, g, r, x, m):
        lib().orc_Ddag(self.h, g.shape[0], _dp(g), _dp(r), _dp(x), m)

    def stagD2xx(self, g, r, x, m2, parity, t=None):
        if t is None:
            t = self.vector()
        lib().orc_stagD2xx(self.h, g.shape[0], _dp(g), _dp(r), _dp(x), m2, parity, _dp(t))

    def eoReconstruct(self, g, r, b, m):
        lib().orc_eoReconstruct(self.h, g.shape[0], _dp(g), _dp(r), _dp(b), m)

    def norm2(self, x, sub=ALL):
        return lib().orc_norm2(self.h, _dp(x), sub)

    def redot(self, x, y, sub=ALL):
        return lib().orc_redot(self.h, _dp(x), _dp(y), sub)

    # -- solvers -----------------------------------------------------------
    def cg_xx(self, g, x, b, m2, parity, r2req, maxits):
        r2 = C.c_double()
        b2 = C.c_double()
        it = lib().orc_cg_xx(self.h, g.shape[0], _dp(g), _dp(x), _dp(b), m2, parity, r2req, maxits,
                             C.byref(r2), C.byref(b2))
        return it, r2.value, b2.value

    def solveXX(self, g, r, x, m, parity, r2req, maxits):
        r2 = C.c_double()
        it = lib().orc_solveXX(self.h, g.shape[0], _dp(g), _dp(r), _dp(x), m, parity, r2req, maxits, C.byref(r2))
        return it, r2.value

    def solve(self, g, x, b, m, r2req, maxits=50000, usePrevSoln=False):
        st = SolveStats()
        lib().orc_solve(self.h, g.shape[0], _dp(g), _dp(x), _dp(b), m, r2req, maxits, int(usePrevSoln), C.byref(st))
        return st

    def cgm_xx(self, g, xs, b, m2, sigmas, parity, r2req, maxits):
        """Multi-shift CG (cgm.nim:84-316): xs float64[nmass, vol, 3, 2]; returns (iterations, r2, b2)."""
        sg = np.ascontiguousarray(sigmas, dtype=np.float64)
        r2, b2 = C.c_double(), C.c_double()
        it = lib().orc_cgm_xx(self.h, g.shape[0], _dp(g), _dp(xs), _dp(b), m2, _dp(sg), len(sg), parity, r2req,
                              maxits, C.byref(r2), C.byref(b2))
        return it, r2.value, b2.value

    def solve_multi(self, g, xs, b, ms, r2req, maxits=50000):
        """solve(s, xs, b, ms, sp) (stagSolve.nim:347-446)."""
        m = np.ascontiguousarray(ms, dtype=np.float64)
        st = SolveStats()
        rc = lib().orc_solve_multi(self.h, g.shape[0], _dp(g), _dp(xs), _dp(b), _dp(m), len(m), r2req, maxits,
                                   C.byref(st))
        if rc != 0:
            raise RuntimeError("solve_multi: no subset qualifies for a solve")
        return st


class RngField:
    def __init__(self, lat, seed):
        self.lat = lat
        self.h = lib().orc_rng_new(lat.h, C.c_uint64(seed))

    def __del__(self):
        try:
            lib().orc_rng_free(self.h)
        except Exception:
            pass


class QexLayout:
    """layoutSetupQ / layoutIndexQ / layoutCoordQ (qlayout.nim:10-185) for any rank/inner geometry."""

    def __init__(self, phys, rankg=(1, 1, 1, 1), inner=(1, 1, 1, 1)):
        self.s = OrcLayout()
        rc = lib().orc_layout_setup(C.byref(self.s), _i4(phys), _i4(rankg), _i4(inner))
        if rc != 0:
            raise ValueError("QEX cannot lay out this geometry (rc=%d)" % rc)

    def index(self, coords):
        r, i = C.c_int(), C.c_int()
        lib().orc_layout_index(C.byref(self.s), _i4(coords), C.byref(r), C.byref(i))
        return r.value, i.value

    def table(self):
        """(rank[vol], index[vol]) for every global site in the oracle's V=1 order."""
        vol = int(np.prod(self.s.phys[:]))
        r = np.zeros(vol, dtype=np.int32)
        i = np.zeros(vol, dtype=np.int32)
        ip = C.POINTER(C.c_int)
        lib().orc_layout_table(C.byref(self.s), r.ctypes.data_as(ip), i.ctypes.data_as(ip))
        return r, i

    @property
    def V(self):
        return self.s.nSitesInner

    @property
    def nSites(self):
        return self.s.nSites

    def scatter(self, f):
        """Global oracle field f[vol, ...] -> list over ranks of native QEX arrays [nOuter, ..., V]."""
        r, i = self.table()
        V, ns = self.V, self.nSites
        nr = int(np.prod(self.s.rankg[:]))
        inner = f.shape[1:]
        out = []
        for rk in range(nr):
            sel = np.nonzero(r == rk)[0]
            loc = np.zeros((ns,) + inner, dtype=f.dtype)
            loc[i[sel]] = f[sel]
            loc = loc.reshape((ns // V, V) + inner)
            out.append(np.ascontiguousarray(np.moveaxis(loc, 1, -1)))
        return out

    def gather(self, parts):
        """Inverse of scatter."""
        r, i = self.table()
        V, ns = self.V, self.nSites
        inner = parts[0].shape[1:-1]
        vol = r.shape[0]
        f = np.zeros((vol,) + inner, dtype=parts[0].dtype)
        for rk, loc in enumerate(parts):
            flat = np.moveaxis(loc, -1, 1).reshape((ns,) + inner)
            sel = np.nonzero(r == rk)[0]
            f[sel] = flat[i[sel]]
        return f

    def coord(self, rank, index):
        c = (C.c_int * 4)()
        lib().orc_layout_coord(C.byref(self.s), rank, index, c)
        return tuple(c)


def hisq_coefs(f7lf=0.0, naik=1.0):
    """HisqCoefs.init + setHisqFat7 (hisqLinks.nim:10-24) as the flat 11-vector hisq_smear takes."""
    def fat7(f, n):
        return [(1.0 + 3.0 * f + n) / 8.0, -1.0 / 16.0, 1.0 / 64.0, -1.0 / 384.0, -f / 16.0]
    return np.array(fat7(f7lf, 0.0) + fat7(2.0 - f7lf, naik) + [-naik / 24.0])


def set_gauss_variant(v):
    lib().orc_set_gauss_variant(v)


def set_num_threads(n):
    lib().orc_set_num_threads(n)


def num_threads():
    return lib().orc_num_threads()
