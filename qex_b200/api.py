"""ctypes binding of include/qex_b200.h plus the QEX-shaped convenience layer.

Host fields are numpy float64 arrays holding NATIVE QEX memory:
colour vector ``[nSitesOuter, 3, 2, V]``, colour matrix ``[nSitesOuter, 3, 3, 2, V]``
(for V=1 that is simply ``[nSites, 3, 2]`` / ``[nSites, 3, 3, 2]``).
"""
import ctypes as C
import os

import numpy as np

EVEN, ODD, ALL = 0, 1, 2

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBPATH = os.environ.get("QEXB200_LIB") or os.path.join(_HERE, "libqex_b200.so")   # override: tuning builds only
_lib = None


class QexB200Error(RuntimeError):
    pass


class _CSolverParams(C.Structure):
    _fields_ = [("r2req", C.c_double), ("maxits", C.c_int), ("use_prev_soln", C.c_int),
                ("verbosity", C.c_int), ("sloppy_solve", C.c_int), ("calls", C.c_int), ("iterations", C.c_int),
                ("iterations_max", C.c_int), ("seconds", C.c_double), ("flops", C.c_double),
                ("r2", C.c_double), ("sloppy_used", C.c_int)]


class HaloDir(C.Structure):
    _fields_ = [("mu", C.c_int), ("depth", C.c_int), ("faceh", C.c_int), ("rank_plus", C.c_int),
                ("rank_minus", C.c_int), ("off_f", C.c_longlong), ("n_f", C.c_longlong),
                ("off_b", C.c_longlong), ("n_b", C.c_longlong)]


class Fat7lCoefs(C.Structure):
    """Fat7lCoefs (src/gauge/fat7l.nim:5-10)."""
    _fields_ = [("oneLink", C.c_double), ("threeStaple", C.c_double), ("fiveStaple", C.c_double),
                ("sevenStaple", C.c_double), ("lepage", C.c_double)]


class HisqCoefs(C.Structure):
    """HisqCoefs (src/physics/hisqLinks.nim:4-43): init() sets the HISQ defaults, smear(g, fl, ll)."""
    _fields_ = [("fat7first", Fat7lCoefs), ("fat7second", Fat7lCoefs), ("naik", C.c_double)]

    def init(self):
        lib().qexb200_hisq_coefs_init(C.byref(self))
        return self

    def smear(self, g, fl, ll):
        _ck(lib().qexb200_gauge_smear_hisq(fl.h, ll.h, g.h, C.byref(self)))

    def smearGetForce(self, u, su, sul):
        """smearGetForce(u, su, sul) (src/gauge/hisqsmear.nim:51-84): returns smearedForce(dsdu, dsdsu, dsdsul)
        (dsdu accumulates); its .free() releases the kept fields."""
        h = C.c_int()
        _ck(lib().qexb200_hisq_smear_get_force(su.h, sul.h, u.h, C.byref(self), C.byref(h)))

        def smearedForce(dsdu, dsdsu, dsdsul):
            _ck(lib().qexb200_hisq_smeared_force(h.value, dsdu.h, dsdsu.h, dsdsul.h))

        def free():
            if is_initialized():
                lib().qexb200_hisq_destroy(h.value)
        smearedForce.free = free
        return smearedForce


class HypCoefs(C.Structure):
    """HypCoefs (src/gauge/hypsmear.nim:16-19,269-273): smear(g, fl)."""
    _fields_ = [("alpha1", C.c_double), ("alpha2", C.c_double), ("alpha3", C.c_double)]

    def smear(self, g, fl):
        _ck(lib().qexb200_gauge_smear_hyp(fl.h, g.h, C.byref(self)))

    def smearGetForce(self, g, fl):
        """smearGetForce(gf, fl) (hypsmear.nim:49-162): smears and returns smearedForce(f, chain) as a callable
        (its .free() releases the kept fields; g must stay unchanged meanwhile)."""
        h = C.c_int()
        _ck(lib().qexb200_hyp_smear_get_force(fl.h, g.h, C.byref(self), C.byref(h)))

        def smearedForce(f, chain):
            _ck(lib().qexb200_hyp_smeared_force(h.value, f.h, chain.h))

        def free():
            if is_initialized():
                lib().qexb200_hyp_destroy(h.value)
        smearedForce.free = free
        return smearedForce


def lib_path():
    return _LIBPATH


def lib():
    """Loads libqex_b200.so; fails loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIBPATH):
            raise QexB200Error("%s not found: build it with `python -m qex_b200.build` "
                               "(or __graft_entry__.build()); there is no CPU fallback" % _LIBPATH)
        L = C.CDLL(_LIBPATH)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        spp = C.POINTER(_CSolverParams)
        L.qexb200_last_error.restype = C.c_char_p
        L.qexb200_init.argtypes = [ip, ip, C.c_int, C.c_int, C.c_void_p]
        L.qexb200_nccl_unique_id.argtypes = [C.c_void_p]
        L.qexb200_p2p_export.argtypes = [C.c_void_p]
        L.qexb200_p2p_attach.argtypes = [C.c_void_p]
        L.qexb200_set_host_layout.argtypes = [ip]
        L.qexb200_local_sites.argtypes = [ip, ip]
        L.qexb200_gauge_create.argtypes = [C.c_int, ip]
        L.qexb200_gauge_load.argtypes = [C.c_int, C.POINTER(dp)]
        L.qexb200_gauge_destroy.argtypes = [C.c_int]
        L.qexb200_gauge_rephase.argtypes = [C.c_int]
        L.qexb200_gauge_recon.argtypes = [C.c_int, C.c_int, ip, dp]
        L.qexb200_hisq_coefs_init.restype = None
        L.qexb200_hisq_coefs_init.argtypes = [C.POINTER(HisqCoefs)]
        L.qexb200_gauge_makeImpLinks.argtypes = [C.c_int, C.c_int, C.POINTER(Fat7lCoefs), C.c_int, C.c_int, C.c_double]
        L.qexb200_gauge_smear_hisq.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(HisqCoefs)]
        L.qexb200_gauge_smear_hyp.argtypes = [C.c_int, C.c_int, C.POINTER(HypCoefs)]
        L.qexb200_gauge_projectU.argtypes = [C.c_int, C.c_int]
        L.qexb200_gauge_fat7lDeriv.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(Fat7lCoefs), C.c_int, C.c_int, C.c_int, C.c_double]
        L.qexb200_hisq_smear_get_force.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(HisqCoefs), ip]
        L.qexb200_hisq_smeared_force.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
        L.qexb200_hisq_destroy.argtypes = [C.c_int]
        L.qexb200_hyp_smear_get_force.argtypes = [C.c_int, C.c_int, C.POINTER(HypCoefs), ip]
        L.qexb200_hyp_smeared_force.argtypes = [C.c_int, C.c_int, C.c_int]
        L.qexb200_hyp_destroy.argtypes = [C.c_int]
        L.qexb200_gauge_projTAH.argtypes = [C.c_int, C.c_int, C.c_int]
        L.qexb200_gauge_plaq.argtypes = [C.c_int, dp]
        L.qexb200_gauge_join3.argtypes = [C.c_int, C.c_int, C.c_int]
        L.qexb200_gauge_split3.argtypes = [C.c_int, C.c_int, C.c_int]
        L.qexb200_gauge_download.argtypes = [C.c_int, C.POINTER(dp)]
        L.qexb200_gauge_zero.argtypes = [C.c_int]
        L.qexb200_gauge_scale.argtypes = [C.c_int, C.c_double]
        L.qexb200_dev_stagDeriv.argtypes = [C.c_int, C.c_int, C.c_int]
        L.qexb200_stagDeriv.argtypes = [C.c_int, C.POINTER(dp), dp]
        L.qexb200_dev_stagD2deriv.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
        L.qexb200_stagD2deriv.argtypes = [C.c_int, C.POINTER(dp), dp, dp]
        L.qexb200_dev_outer.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int]
        L.qexb200_scidac_lattice.argtypes = [C.c_char_p, ip]
        L.qexb200_gauge_read_scidac.argtypes = [C.c_int, C.c_char_p, C.POINTER(C.c_uint)]
        L.qexb200_gauge_load_sitemajor.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.qexb200_stagD2.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double, C.c_double]
        L.qexb200_stagD.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double, C.c_double, C.c_double]
        L.qexb200_D.argtypes = [C.c_int, dp, dp, C.c_double]
        L.qexb200_Ddag.argtypes = [C.c_int, dp, dp, C.c_double]
        L.qexb200_stagD2xx.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double]
        L.qexb200_peqDdag.argtypes = [C.c_int, dp, dp, C.c_double]
        L.qexb200_eoReduce.argtypes = [C.c_int, dp, dp, C.c_double]
        L.qexb200_stagDb.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double, C.c_double]
        L.qexb200_stagD1.argtypes = [C.c_int, C.c_int, dp, dp, C.c_int]
        L.qexb200_dev_peqDdag.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
        L.qexb200_dev_eoReduce.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
        L.qexb200_eoReconstruct.argtypes = [C.c_int, dp, dp, C.c_double]
        L.qexb200_solveXX.argtypes = [C.c_int, C.c_int, dp, dp, C.c_double, spp]
        L.qexb200_solve.argtypes = [C.c_int, dp, dp, C.c_double, spp]
        L.qexb200_solveXX_multi.argtypes = [C.c_int, C.c_int, C.POINTER(dp), dp, dp, C.c_int, spp]
        L.qexb200_solve_multi.argtypes = [C.c_int, C.POINTER(dp), dp, dp, C.c_int, spp]
        L.qexb200_dev_solveXX_multi.argtypes = [C.c_int, C.c_int, ip, C.c_int, dp, C.c_int, spp]
        L.qexb200_dev_solve_multi.argtypes = [C.c_int, ip, C.c_int, dp, C.c_int, spp]
        L.qexb200_rng_create.argtypes = [C.c_ulonglong, ip]
        L.qexb200_rng_destroy.argtypes = [C.c_int]
        L.qexb200_gauge_random.argtypes = [C.c_int, C.c_int]
        L.qexb200_gauge_randomTAH.argtypes = [C.c_int, C.c_int]
        L.qexb200_gauge_unit.argtypes = [C.c_int]
        L.qexb200_gauge_reunit.argtypes = [C.c_int, dp]
        L.qexb200_field_gaussian.argtypes = [C.c_int, C.c_int]
        L.qexb200_field_create.argtypes = [ip]
        L.qexb200_field_destroy.argtypes = [C.c_int]
        L.qexb200_field_upload.argtypes = [C.c_int, dp]
        L.qexb200_field_download.argtypes = [C.c_int, dp]
        L.qexb200_field_zero.argtypes = [C.c_int]
        L.qexb200_field_norm2.argtypes = [C.c_int, C.c_int, dp]
        L.qexb200_field_redot.argtypes = [C.c_int, C.c_int, C.c_int, dp]
        L.qexb200_dev_stagD2.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]
        L.qexb200_dev_stagD.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
        L.qexb200_dev_D.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
        L.qexb200_dev_Ddag.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double]
        L.qexb200_dev_stagD2xx.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double]
        L.qexb200_dev_solveXX.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, spp]
        L.qexb200_dev_solve.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, spp]
        L.qexb200_plan_peers.argtypes = [ip, ip, C.c_int, C.POINTER(C.c_ubyte)]
        L.qexb200_gauge_cached.argtypes = [C.POINTER(C.POINTER(C.c_double)), C.c_int, C.c_ulonglong, ip, ip]
        L.qexb200_gauge_invalidate.argtypes = [C.POINTER(C.POINTER(C.c_double)), C.c_int]
        L.qexb200_field_axpby.argtypes = [C.c_int, C.c_double, C.c_int, C.c_double, C.c_int]
        L.qexb200_solveEO.argtypes = [C.c_int, dp, dp, C.c_double, spp]
        L.qexb200_dev_solveEO.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, spp]
        L.qexb200_cg_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, ip]
        L.qexb200_cg_reset.argtypes = [C.c_int]
        L.qexb200_cg_solve.argtypes = [C.c_int, C.c_double, spp]
        L.qexb200_cg_destroy.argtypes = [C.c_int]
        L.qexb200_plan_tiles.argtypes = [ip, ip, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint), ip]
        L.qexb200_debug_role_times.argtypes = [C.c_int, C.POINTER(C.c_ulonglong), C.c_int]
        L.qexb200_plan_halo.argtypes = [ip, ip, C.c_int, C.c_int, C.POINTER(HaloDir), ip, C.POINTER(C.c_longlong)]
        L.qexb200_timer_stop.argtypes = [C.POINTER(C.c_float)]
        L.qexb200_norm2_noalloc.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
        L.qexb200_launch_count.restype = C.c_longlong
        L.qexb200_launch_count.argtypes = [C.c_int]
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise QexB200Error(lib().qexb200_last_error().decode())


def _dp(a):
    if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]):
        raise QexB200Error("host fields must be C-contiguous float64 numpy arrays")
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _i4(v):
    return (C.c_int * 4)(*[int(t) for t in v])


# ---- lifetime ----------------------------------------------------------------
def nccl_unique_id():
    try:
        import torch  # noqa: F401  (see init)
    except ImportError:
        pass
    buf = C.create_string_buffer(128)
    _ck(lib().qexb200_nccl_unique_id(buf))
    return buf.raw


def init(latsize, rankgeom=(1, 1, 1, 1), myrank=0, device=0, nccl_id=None):
    """qexInit + qudaInit analogue (src/quda/qudaWrapperImpl.nim:71-127)."""
    if nccl_id is not None:
        # the process must hold ONE libnccl.so.2; when torch is the launcher its copy has to be
        # loaded before ours is bound (csrc/capi.cu nccl_load)
        try:
            import torch  # noqa: F401
        except ImportError:
            pass
    idbuf = C.create_string_buffer(nccl_id, 128) if nccl_id is not None else None
    _ck(lib().qexb200_init(_i4(latsize), _i4(rankgeom), int(myrank), int(device), idbuf))


def p2p_export():
    """64-byte handle of this rank's peer-memory window (to be all-gathered in rank order)."""
    buf = C.create_string_buffer(64)
    _ck(lib().qexb200_p2p_export(buf))
    return buf.raw


def p2p_attach(handles):
    """handles: the concatenated 64-byte handles of all ranks, in rank order."""
    buf = C.create_string_buffer(bytes(handles), len(handles))
    _ck(lib().qexb200_p2p_attach(buf))


def p2p_active():
    return _lib is not None and bool(_lib.qexb200_p2p_active())


def p2p_attach_torch(dist, device="cuda"):
    """Convenience for torch.distributed launchers: all-gather the window handles and attach."""
    import torch
    mine = torch.tensor(list(p2p_export()), dtype=torch.uint8, device=device)
    allh = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(allh, mine)
    p2p_attach(b"".join(bytes(t.cpu().tolist()) for t in allh))
    dist.barrier()


def plan_halo(latsize, rankgeom, myrank, nlinks):
    """Halo plan of one rank (pure host arithmetic, works without a GPU)."""
    out = (HaloDir * 4)()
    n, tot = C.c_int(), C.c_longlong()
    _ck(lib().qexb200_plan_halo(_i4(latsize), _i4(rankgeom), int(myrank), int(nlinks), out, C.byref(n), C.byref(tot)))
    return [dict((f[0], getattr(out[i], f[0])) for f in HaloDir._fields_) for i in range(n.value)], tot.value


def plan_tiles(latsize, rankgeom, myrank, nlinks, parity):
    """Block -> tile table of the multi-rank sweep (host arithmetic): list of (tile, need bits)."""
    n = C.c_int(0)
    _ck(lib().qexb200_plan_tiles(_i4(latsize), _i4(rankgeom), int(myrank), int(nlinks), int(parity), None, C.byref(n)))
    out = (C.c_uint * n.value)()
    _ck(lib().qexb200_plan_tiles(_i4(latsize), _i4(rankgeom), int(myrank), int(nlinks), int(parity), out, C.byref(n)))
    return [(int(e) & 0xffffff, int(e) >> 24) for e in out]


def role_times(on):
    """Arms (on=True) or reads and disarms the per-launch role timestamps of the multi-rank sweep; returns
    (ring[1024][8], xrank[8]) as numpy uint64 when reading."""
    import numpy as np
    if on:
        _ck(lib().qexb200_debug_role_times(1, None, 0))
        return None
    n = 1024 * 8 + 8
    buf = (C.c_ulonglong * n)()
    _ck(lib().qexb200_debug_role_times(0, buf, n))
    a = np.frombuffer(buf, dtype=np.uint64).copy()
    return a[:1024 * 8].reshape(1024, 8), a[1024 * 8:]


def plan_peers(latsize, rankgeom, myrank):
    """peer_of[81] of one rank (pure host arithmetic, works without a GPU)."""
    out = (C.c_ubyte * 81)()
    _ck(lib().qexb200_plan_peers(_i4(latsize), _i4(rankgeom), int(myrank), out))
    return list(out)


def finalize():
    if _lib is not None:
        _ck(_lib.qexb200_finalize())


def is_initialized():
    return _lib is not None and bool(_lib.qexb200_is_initialized())


def getFileLattice(fn):
    """getFileLattice (src/io/readerQiolite.nim:11-17)."""
    d = (C.c_int * 4)()
    _ck(lib().qexb200_scidac_lattice(os.fsencode(fn), d))
    return list(d)


def set_host_layout(innergeom):
    _ck(lib().qexb200_set_host_layout(_i4(innergeom)))


def local_sites():
    n, e = C.c_int(), C.c_int()
    _ck(lib().qexb200_local_sites(C.byref(n), C.byref(e)))
    return n.value, e.value


def sync():
    _ck(lib().qexb200_sync())


def timer_start():
    _ck(lib().qexb200_timer_start())


def timer_stop():
    ms = C.c_float()
    _ck(lib().qexb200_timer_stop(C.byref(ms)))
    return ms.value


def launch_count(reset=False):
    return int(lib().qexb200_launch_count(int(reset)))


# ---- SolverParams (src/solvers/solverBase.nim:10-58) ----------------------------
class SolverParams:
    def __init__(self, r2req=1e-6, maxits=50000, usePrevSoln=False, verbosity=1, sloppySolve=0):
        self.sloppySolve = sloppySolve   # SloppyNone/Single/Half = 0/1/2; accepted, the solve is fp64 (sloppyUsed == 0)
        self.sloppyUsed = 0
        self.r2req = r2req
        self.maxits = maxits
        self.usePrevSoln = usePrevSoln
        self.verbosity = verbosity
        self.resetStats()

    def resetStats(self):
        self.calls = 0
        self.iterations = 0
        self.iterationsMax = 0
        self.seconds = 0.0
        self.flops = 0.0
        self.r2 = 0.0

    def _c(self):
        c = _CSolverParams()
        c.r2req, c.maxits, c.use_prev_soln, c.verbosity = self.r2req, int(self.maxits), int(self.usePrevSoln), self.verbosity
        c.sloppy_solve = int(self.sloppySolve)
        return c

    def _addStats(self, c):  # addStats, solverBase.nim:75-83
        self.calls += c.calls
        self.iterations += c.iterations
        self.iterationsMax = max(self.iterationsMax, c.iterations_max)
        self.seconds += c.seconds
        self.flops += c.flops
        self.r2 = c.r2
        self.sloppyUsed = c.sloppy_used

    def getStats(self):  # getStats typ 3, solverBase.nim:84-112
        s = max(1e-12, self.seconds)
        return "%d  %.4gs  %.4gGf/s  %.4g" % (self.iterations // max(1, self.calls), s / max(1, self.calls),
                                            1e-9 * self.flops / s, self.r2)


# ---- per-site random number field (newRNGField, src/rng/distributionUtils.nim:306-331) ------------
class RngField:
    """newRNGField(RngMilc6, lo, seed) on the device; g.random(r), p.randomTAH(r) and v.gaussian(r) draw from it."""

    def __init__(self, seed=17 ** 7):
        h = C.c_int()
        _ck(lib().qexb200_rng_create(C.c_ulonglong(int(seed)), C.byref(h)))
        self.h = h.value

    def free(self):
        if self.h is not None and is_initialized():
            lib().qexb200_rng_destroy(self.h)
        self.h = None


def newRNGField(seed=17 ** 7):
    return RngField(seed)


# ---- device-resident field -------------------------------------------------------
class DeviceField:
    def __init__(self, host=None):
        h = C.c_int()
        _ck(lib().qexb200_field_create(C.byref(h)))
        self.h = h.value
        if host is not None:
            self.upload(host)

    def upload(self, host):
        _ck(lib().qexb200_field_upload(self.h, _dp(host)))

    def download(self, out):
        _ck(lib().qexb200_field_download(self.h, _dp(out)))
        return out

    def zero(self):
        _ck(lib().qexb200_field_zero(self.h))

    def gaussian(self, rng):
        """v.gaussian r (src/rng/distributionUtils.nim:65-97)."""
        _ck(lib().qexb200_field_gaussian(self.h, rng.h))

    def norm2(self, subset=ALL):
        o = C.c_double()
        _ck(lib().qexb200_field_norm2(self.h, subset, C.byref(o)))
        return o.value

    def axpby(self, a, x, b, subset=ALL):
        """self = a x + b self on a subset (`:=`, `+=` of src/field/fieldET.nim:547-599)."""
        _ck(lib().qexb200_field_axpby(self.h, float(a), x.h, float(b), subset))

    def assign(self, x, subset=ALL):
        self.axpby(1.0, x, 0.0, subset)

    def redot(self, other, subset=ALL):
        o = C.c_double()
        _ck(lib().qexb200_field_redot(self.h, other.h, subset, C.byref(o)))
        return o.value

    def free(self):
        if self.h is not None and is_initialized():
            lib().qexb200_field_destroy(self.h)
        self.h = None


# ---- device link sets ---------------------------------------------------------------
class DeviceGauge:
    """A device-resident link set (4 or 8 link fields): lo.newGauge() + the link pipeline of
    src/gauge/fat7l.nim, src/physics/hisqLinks.nim, src/gauge/hypsmear.nim on the GPU."""

    def __init__(self, links=None, nlinks=4):
        self.nlinks = len(links) if links is not None else int(nlinks)
        h = C.c_int()
        _ck(lib().qexb200_gauge_create(self.nlinks, C.byref(h)))
        self.h = h.value
        self._owner = True
        if links is not None:
            self.load(links)

    def load(self, links):
        arr = (C.POINTER(C.c_double) * self.nlinks)(*[_dp(l) for l in links])
        _ck(lib().qexb200_gauge_load(self.h, arr))

    def download(self, links):
        arr = (C.POINTER(C.c_double) * self.nlinks)(*[_dp(l) for l in links])
        _ck(lib().qexb200_gauge_download(self.h, arr))
        return links

    def rephase(self):
        _ck(lib().qexb200_gauge_rephase(self.h))

    def random(self, rng):
        """g.random r (src/gauge/gaugeUtils.nim:1424-1429)."""
        _ck(lib().qexb200_gauge_random(self.h, rng.h))

    def randomTAH(self, rng):
        """p.randomTAH r (gaugeUtils.nim:1353-1383,1452-1454)."""
        _ck(lib().qexb200_gauge_randomTAH(self.h, rng.h))

    def unit(self):
        """g.unit (gaugeUtils.nim:1448-1450)."""
        _ck(lib().qexb200_gauge_unit(self.h))

    def reunit(self):
        """reunit(g) (src/examples/staghmc_sh.nim:247-257): projectSU on every link; returns
        ((avg, max) before, (avg, max) after) of checkSU."""
        d = np.zeros(4)
        _ck(lib().qexb200_gauge_reunit(self.h, _dp(d)))
        return (float(d[0]), float(d[1])), (float(d[2]), float(d[3]))

    def loadGauge(self, fn):
        """loadGauge(g, fn) (src/gauge/gaugeUtils.nim:87-106): first lattice record of a SciDAC/ILDG file.
        Returns (file suma, file sumb, this rank's suma, this rank's sumb)."""
        sums = (C.c_uint * 4)()
        _ck(lib().qexb200_gauge_read_scidac(self.h, os.fsencode(fn), sums))
        return tuple(int(t) for t in sums)

    def load_sitemajor(self, data, big_endian=False):
        """Site-major array [site lex][mu][3][3][2] (float32 or float64), e.g. MILC order."""
        a = np.ascontiguousarray(data)
        _ck(lib().qexb200_gauge_load_sitemajor(self.h, a.ctypes.data_as(C.c_void_p), a.dtype.itemsize, int(big_endian)))

    def plaq(self):
        """plaq (src/gauge/gaugeUtils.nim:213-284) of the one-hop links."""
        pl = np.zeros(6)
        _ck(lib().qexb200_gauge_plaq(self.h, _dp(pl)))
        return pl

    def projectU(self, src):
        _ck(lib().qexb200_gauge_projectU(self.h, src.h))

    def projTAH(self, g, adj="no_adj"):
        """projTAH(f, g, adj) (src/stagg_pv_hmc/staghmc_spv_gforce.nim:256-291)."""
        _ck(lib().qexb200_gauge_projTAH(self.h, g.h, int(adj == "adj")))

    def split3(self, fl, ll):
        """fl[mu] = self[2 mu], ll[mu] = self[2 mu + 1] (the inverse of newStag3's interleaving)."""
        _ck(lib().qexb200_gauge_split3(fl.h, ll.h, self.h))

    def zero(self):
        _ck(lib().qexb200_gauge_zero(self.h))

    def scale(self, c):
        _ck(lib().qexb200_gauge_scale(self.h, float(c)))

    def outer(self, a, b, scale_even=1.0, scale_odd=1.0, accumulate=True):
        """f[mu](x) (+)= scale * a(x) (x) b(x+mu)^+ on DeviceFields (staghmc_spv.nim:826-850)."""
        _ck(lib().qexb200_dev_outer(self.h, a.h, b.h, float(scale_even), float(scale_odd), int(accumulate)))

    def recon(self, want=12):
        """Ask for 12-real (rows 0,1 + rebuilt third row; +-s*SU(3) links), 14-real (rows 0,1 + one complex factor
        per link; c*U(3) links: projected smeared links, HISQ Naik links) or 18-real link storage; four digits
        address the one-hop and three-hop class separately (1814 = fat links 18, Naik links 14).  Returns
        ([storage of the one-hop class, of the three-hop class], [measured deviation per class])."""
        got, dev = (C.c_int * 2)(), (C.c_double * 2)()
        _ck(lib().qexb200_gauge_recon(self.h, int(want), got, dev))
        return [got[0], got[1]], [dev[0], dev[1]]

    def free(self):
        if self.h is not None and self._owner and is_initialized():
            lib().qexb200_gauge_destroy(self.h)
        self.h = None


def makeImpLinks(fl, gf, coef, ll=None, gfLong=None, naik=0.0):
    """makeImpLinks(fl, gf, coef[, ll, gfLong, naik]) (src/gauge/fat7l.nim:78-165) on DeviceGauges."""
    _ck(lib().qexb200_gauge_makeImpLinks(fl.h, gf.h, C.byref(coef), ll.h if ll is not None else -1,
                                         gfLong.h if gfLong is not None else -1, float(naik)))


# ---- Staggered (src/physics/stagD.nim:15-22,522-586; stagSolve.nim) ---------------
class Staggered(DeviceGauge):
    """Staggered[G,T]: holds the link set (device resident) and offers D/Ddag/solve.  Built from host
    link fields (uploaded) or from a DeviceGauge (shared by reference, like newStag keeps `g`)."""

    def __init__(self, links):
        if isinstance(links, DeviceGauge):
            self.nlinks, self.h, self._owner = links.nlinks, links.h, False
        else:
            DeviceGauge.__init__(self, links)

    # host-field operators
    def D(self, r, x, m):
        _ck(lib().qexb200_D(self.h, _dp(r), _dp(x), m))

    def Ddag(self, r, x, m):
        _ck(lib().qexb200_Ddag(self.h, _dp(r), _dp(x), m))

    def eoReconstruct(self, r, b, m):
        _ck(lib().qexb200_eoReconstruct(self.h, _dp(r), _dp(b), m))

    def peqDdag(self, r, x, m):
        """peqDdag (stagD.nim:572-574): r += Ddag x (host arrays, or DeviceFields)."""
        if isinstance(r, DeviceField):
            _ck(lib().qexb200_dev_peqDdag(self.h, r.h, x.h, m))
        else:
            _ck(lib().qexb200_peqDdag(self.h, _dp(r), _dp(x), m))

    def eoReduce(self, r, b, m):
        """eoReduce (stagD.nim:575-580): r.even = (Ddag b).even."""
        if isinstance(r, DeviceField):
            _ck(lib().qexb200_dev_eoReduce(self.h, r.h, b.h, m))
        else:
            _ck(lib().qexb200_eoReduce(self.h, _dp(r), _dp(b), m))

    def solveXX(self, r, x, m, sp, parEven=True):
        c = sp._c()
        _ck(lib().qexb200_solveXX(self.h, EVEN if parEven else ODD, _dp(r), _dp(x), m, C.byref(c)))
        sp._addStats(c)

    def solveEE(self, r, x, m, sp):
        self.solveXX(r, x, m, sp, True)

    def solveEO(self, r, x, m, sp):
        """solveEO (stagSolve.nim:17-38): CG on the even subset starting from the content of r (host arrays or
        DeviceFields)."""
        c = sp._c()
        if isinstance(r, DeviceField):
            _ck(lib().qexb200_dev_solveEO(self.h, r.h, x.h, m, C.byref(c)))
        else:
            _ck(lib().qexb200_solveEO(self.h, _dp(r), _dp(x), m, C.byref(c)))
        sp._addStats(c)

    def solveOO(self, r, x, m, sp):
        self.solveXX(r, x, m, sp, False)

    def solve(self, x, b, m, sp):
        """solve(s, x, b, m, sp0) (stagSolve.nim:224-294); `sp` may be a SolverParams or an r2req float
        like the convenience overload at stagSolve.nim:462-472."""
        if not isinstance(sp, SolverParams):
            sp = SolverParams(r2req=float(sp))
        c = sp._c()
        _ck(lib().qexb200_solve(self.h, _dp(x), _dp(b), m, C.byref(c)))
        sp._addStats(c)
        return sp

    # multi-shift (solveXX(s, xs, b, ms, ...) / solve(s, xs, b, ms, sp), stagSolve.nim:296-446)
    def solveXX_multi(self, xs, b, shifts, sp, parEven=True):
        arr = (C.POINTER(C.c_double) * len(xs))(*[_dp(x) for x in xs])
        sh = (C.c_double * len(xs))(*[float(t) for t in shifts])
        c = sp._c()
        _ck(lib().qexb200_solveXX_multi(self.h, EVEN if parEven else ODD, arr, _dp(b), sh, len(xs), C.byref(c)))
        sp._addStats(c)

    def solve_multi(self, xs, b, ms, sp):
        arr = (C.POINTER(C.c_double) * len(xs))(*[_dp(x) for x in xs])
        mm = (C.c_double * len(xs))(*[float(t) for t in ms])
        c = sp._c()
        _ck(lib().qexb200_solve_multi(self.h, arr, _dp(b), mm, len(xs), C.byref(c)))
        sp._addStats(c)
        return sp

    def dev_solveXX_multi(self, xs, b, shifts, sp, parEven=True):
        hs = (C.c_int * len(xs))(*[x.h for x in xs])
        sh = (C.c_double * len(xs))(*[float(t) for t in shifts])
        c = sp._c()
        _ck(lib().qexb200_dev_solveXX_multi(self.h, EVEN if parEven else ODD, hs, b.h, sh, len(xs), C.byref(c)))
        sp._addStats(c)

    def dev_solve_multi(self, xs, b, ms, sp):
        hs = (C.c_int * len(xs))(*[x.h for x in xs])
        mm = (C.c_double * len(xs))(*[float(t) for t in ms])
        c = sp._c()
        _ck(lib().qexb200_dev_solve_multi(self.h, hs, b.h, mm, len(xs), C.byref(c)))
        sp._addStats(c)

    # device-field operators
    def dev_D(self, r, x, m):
        _ck(lib().qexb200_dev_D(self.h, r.h, x.h, m))

    def dev_Ddag(self, r, x, m):
        _ck(lib().qexb200_dev_Ddag(self.h, r.h, x.h, m))

    def dev_stagD2(self, parity, r, x, a, b):
        _ck(lib().qexb200_dev_stagD2(self.h, parity, r.h, x.h, a, b))

    def dev_stagD(self, parity, r, x, m, sc=1.0, a=0.0):
        _ck(lib().qexb200_dev_stagD(self.h, parity, r.h, x.h, m, sc, a))

    def dev_stagD2xx(self, parity, r, x, m2):
        _ck(lib().qexb200_dev_stagD2xx(self.h, parity, r.h, x.h, m2))

    def dev_solveXX(self, r, x, m, sp, parEven=True):
        c = sp._c()
        _ck(lib().qexb200_dev_solveXX(self.h, EVEN if parEven else ODD, r.h, x.h, m, C.byref(c)))
        sp._addStats(c)

    def dev_solve(self, x, b, m, sp):
        c = sp._c()
        _ck(lib().qexb200_dev_solve(self.h, x.h, b.h, m, C.byref(c)))
        sp._addStats(c)


class CgState:
    """CgState[T] (src/solvers/cg.nim:13-53): newCgState(x, b) bound to A = stagD2ee/oo(m^2) of `stag`; solve() can
    be called again to continue the same Krylov sequence, reset() starts over from the current x."""

    def __init__(self, stag, x, b, parEven=True):
        h = C.c_int()
        _ck(lib().qexb200_cg_create(stag.h, EVEN if parEven else ODD, x.h, b.h, C.byref(h)))
        self.h = h.value

    def reset(self):
        _ck(lib().qexb200_cg_reset(self.h))

    def solve(self, m, sp):
        c = sp._c()
        _ck(lib().qexb200_cg_solve(self.h, m, C.byref(c)))
        sp._addStats(c)

    def free(self):
        if self.h is not None and _lib is not None and _lib.qexb200_is_initialized():
            _lib.qexb200_cg_destroy(self.h)
        self.h = None


def cachedStag(links, version):
    """The glue's handle cache (qexb200_gauge_cached): a Staggered for HOST link fields, uploaded only when the
    (pointers, version) pair is new.  Returns (Staggered that does not own its handle, was_cached)."""
    n = len(links)
    arr = (C.POINTER(C.c_double) * n)(*[_dp(l) for l in links])
    h, hit = C.c_int(), C.c_int()
    _ck(lib().qexb200_gauge_cached(arr, n, C.c_ulonglong(int(version)), C.byref(h), C.byref(hit)))
    g = DeviceGauge.__new__(DeviceGauge)
    g.nlinks, g.h, g._owner = n, h.value, False
    return Staggered(g), bool(hit.value)


def invalidateStag(links):
    """qexb200_gauge_invalidate: forget the device copy cached against these host link arrays (call before freeing them)."""
    n = len(links)
    arr = (C.POINTER(C.c_double) * n)(*[_dp(l) for l in links])
    _ck(lib().qexb200_gauge_invalidate(arr, n))


def newStag(g):
    """newStag(g) (stagD.nim:533-542): naive one-hop operator from 4 link fields."""
    if isinstance(g, DeviceGauge):
        if g.nlinks != 4:
            raise QexB200Error("newStag needs 4 link fields")
        return Staggered(g)
    if len(g) != 4:
        raise QexB200Error("newStag needs 4 link fields")
    return Staggered(list(g))


def newStag3(g, g3=None):
    """newStag3(g, g3) (stagD.nim:552-564): links interleaved as [fat0, long0, fat1, long1, ...];
    newStag3(g) with 8 fields takes them already interleaved (stagD.nim:543-551)."""
    if isinstance(g, DeviceGauge) and isinstance(g3, DeviceGauge):
        out = DeviceGauge(nlinks=8)
        _ck(lib().qexb200_gauge_join3(out.h, g.h, g3.h))
        s = Staggered(out)
        s._owner, out._owner = True, False
        return s
    if g3 is None:
        if len(g) != 8:
            raise QexB200Error("newStag3(g) needs 8 interleaved link fields")
        return Staggered(list(g))
    gg = []
    for a, b in zip(g, g3):
        gg += [a, b]
    return Staggered(gg)


def stagD2(s, parity, r, x, a, b):
    """stagD2(sd, r, g, x, a, b) (stagD.nim:348-395) on the subset `parity`."""
    _ck(lib().qexb200_stagD2(s.h, parity, _dp(r), _dp(x), a, b))


def stagD(s, parity, r, x, m, sc=1.0, a=0.0):
    """stagD(sd, r, g, x, m, sc, a) (stagD.nim:405-409)."""
    _ck(lib().qexb200_stagD(s.h, parity, _dp(r), _dp(x), m, sc, a))


def stagD2ee(s, r, x, m2):
    """stagD2ee (stagD.nim:463-465)."""
    _ck(lib().qexb200_stagD2xx(s.h, EVEN, _dp(r), _dp(x), m2))


def stagD2oo(s, r, x, m2):
    """stagD2oo (stagD.nim:467-469)."""
    _ck(lib().qexb200_stagD2xx(s.h, ODD, _dp(r), _dp(x), m2))


def fat7lDeriv(deriv, gauge, mid, coef, llderiv=None, llgauge=None, llmid=None, naik=0.0):
    """fat7lDeriv (src/gauge/fat7lderiv.nim:122-335) on DeviceGauges; deriv / llderiv accumulate."""
    _ck(lib().qexb200_gauge_fat7lDeriv(deriv.h, gauge.h, mid.h, C.byref(coef), llderiv.h if llderiv is not None else -1,
                                       llgauge.h if llgauge is not None else -1, llmid.h if llmid is not None else -1,
                                       float(naik)))


def stagDeriv(s, f, x):
    """stagDeriv(s, f, x) (stagD.nim:635-664).  f: DeviceGauge + x: DeviceField, or host link fields + host vector."""
    if isinstance(f, DeviceGauge):
        _ck(lib().qexb200_dev_stagDeriv(s.h, f.h, x.h))
    else:
        arr = (C.POINTER(C.c_double) * len(f))(*[_dp(t) for t in f])
        _ck(lib().qexb200_stagDeriv(s.h, arr, _dp(x)))


def stagD2deriv(s, g, c, x):
    """stagD2deriv(s, g, c, x) (stagD.nim:588-633): g += (d/dg) redot[(2D) x, c]."""
    if isinstance(g, DeviceGauge):
        _ck(lib().qexb200_dev_stagD2deriv(s.h, g.h, c.h, x.h))
    else:
        arr = (C.POINTER(C.c_double) * len(g))(*[_dp(t) for t in g])
        _ck(lib().qexb200_stagD2deriv(s.h, arr, _dp(c), _dp(x)))


def stagDb(s, parity, r, x, m, sc=1.0):
    """stagDb(sd, r, g, x, m, sc) (stagD.nim:425-427)."""
    _ck(lib().qexb200_stagDb(s.h, parity, _dp(r), _dp(x), m, sc))


def stagD1(s, parity, r, x):
    """stagD1 (stagD.nim:414-417)."""
    _ck(lib().qexb200_stagD1(s.h, parity, _dp(r), _dp(x), 0))


def stagD1x(s, parity, r, x):
    """stagD1x (stagD.nim:419-422)."""
    _ck(lib().qexb200_stagD1(s.h, parity, _dp(r), _dp(x), 1))
