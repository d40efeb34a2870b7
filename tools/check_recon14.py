"""One-shot check of the 14-real link storage on the GPU box: solve on nHYP links at 32^4 with 18- and 14-real sweeps."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402
from qex_b200.synth import random_su3_links, gaussian_vector  # noqa: E402

L = tuple(int(t) for t in os.environ.get("LAT", "32,32,32,32").split(","))
qb.init(L)
vol = int(np.prod(L))
g, fl = qb.DeviceGauge(random_su3_links(vol, 4, seed=1)), qb.DeviceGauge(nlinks=4)
qb.HypCoefs(0.4, 0.5, 0.5).smear(g, fl)
fl.rephase()
s = qb.newStag(fl)
b, x = qb.DeviceField(gaussian_vector(vol, 7)), qb.DeviceField()
for mode in (18, 14, 18, 14):
    got, dev = s.recon(mode)
    sp = qb.SolverParams(r2req=1e-12, maxits=20000)
    qb.timer_start()
    s.dev_solve(x, b, 0.1, sp)
    ms = qb.timer_stop()
    print("recon", got, "dev %.2e" % dev[0], "solve ms %.3f" % ms, "its", sp.iterations, "r2 %.3e" % sp.r2, "x2 %.15e" % x.norm2(), flush=True)
qb.finalize()
