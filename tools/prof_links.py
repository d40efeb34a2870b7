"""Short run for ncu: one HISQ smear, one nHYP smear and the force outer products at 32^4 on resident links."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402
from qex_b200.synth import random_su3_links, gaussian_vector  # noqa: E402

L = tuple(int(t) for t in os.environ.get("LAT", "32,32,32,32").split(","))
qb.init(L)
vol = int(np.prod(L))
g = qb.DeviceGauge(random_su3_links(vol, 4, seed=1))
fl, ll = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
which = os.environ.get("WHICH", "hisq,nhyp,force").split(",")
reps = int(os.environ.get("REPS", "1"))
for _ in range(reps):
    if "hisq" in which:
        qb.timer_start(); qb.HisqCoefs().init().smear(g, fl, ll); print("hisq ms", qb.timer_stop())
    if "nhyp" in which:
        qb.timer_start(); qb.HypCoefs(0.4, 0.5, 0.5).smear(g, fl); print("nhyp ms", qb.timer_stop())
    if "hypforce" in which:
        f = qb.DeviceGauge(nlinks=4); f.zero()
        force = qb.HypCoefs(0.4, 0.5, 0.5).smearGetForce(g, fl)
        qb.timer_start(); force(f, ll if "hisq" in which else fl); print("nhyp smearedForce ms", qb.timer_stop())
        force.free(); f.free()
    if "force" in which:
        x, c = qb.DeviceField(gaussian_vector(vol, 1)), qb.DeviceField(gaussian_vector(vol, 2))
        s = qb.newStag(g)
        fl.zero()
        qb.timer_start(); qb.stagD2deriv(s, fl, c, x); print("stagD2deriv ms", qb.timer_stop())
        qb.timer_start(); fl.outer(x, x, 1.0, -1.0, accumulate=False); print("outer ms", qb.timer_stop())
        x.free(); c.free()
print("launches", qb.launch_count())
qb.finalize()
