"""qex_b200 -- B200-native staggered Dslash + CG backend behind QEX's operator/solver API.

The product is ``libqex_b200.so`` (CUDA, sm_100a, C ABI in ``include/qex_b200.h``).
This package is the thin host-side mirror of the reference interface used by the
tests and the benchmark: same names and argument meaning as QEX
(``newStag``/``newStag3``, ``Staggered.D/Ddag/solve/solveEE``, ``stagD2``,
``SolverParams``; src/physics/stagD.nim, src/physics/stagSolve.nim,
src/solvers/solverBase.nim), calling straight through ctypes into the C ABI.

There is no CPU fallback: every compute call raises ``QexB200Error`` if the
library is missing or no CUDA device is present.
"""
from .api import (  # noqa: F401
    QexB200Error, SolverParams, Staggered, CgState, DeviceField, DeviceGauge, RngField, newRNGField, Fat7lCoefs, HisqCoefs, HypCoefs, makeImpLinks, fat7lDeriv, getFileLattice, stagDeriv, stagD2deriv, init, finalize, is_initialized,
    newStag, newStag3, cachedStag, invalidateStag, stagD2, stagD, stagD2ee, stagD2oo, stagDb, stagD1, stagD1x, set_host_layout, local_sites,
    nccl_unique_id, plan_halo, plan_peers, plan_tiles, role_times, p2p_export, p2p_attach, p2p_active, p2p_attach_torch, sync, timer_start, timer_stop, launch_count, lib, lib_path,
    EVEN, ODD, ALL,
)
