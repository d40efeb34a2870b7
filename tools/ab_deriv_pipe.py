"""A/B of the staple-derivative kernels on the GPU box (32^4): nHYP smearedForce with the register-gather kernel and
with the shared-memory ring (QEXB200_DERIV_PIPE=1), results compared bit for bit, both timed.
    python tools/ab_deriv_pipe.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402
from qex_b200.synth import random_su3_links  # noqa: E402

L = tuple(int(t) for t in os.environ.get("LAT", "32,32,32,32").split(","))
qb.init(L)
vol = int(np.prod(L))
g, fl = qb.DeviceGauge(random_su3_links(vol, 4, seed=1)), qb.DeviceGauge(nlinks=4)
chain = qb.DeviceGauge(random_su3_links(vol, 4, seed=2))
res = {}
for mode in ("0", "1", "0", "1"):
    os.environ["QEXB200_DERIV_PIPE"] = mode
    force = qb.HypCoefs(0.4, 0.5, 0.5).smearGetForce(g, fl)
    f = qb.DeviceGauge(nlinks=4)
    f.zero()
    qb.timer_start()
    force(f, chain)
    ms = qb.timer_stop()
    out = [np.zeros((vol, 3, 3, 2)) for _ in range(4)]
    f.download(out)
    res[mode] = np.stack(out)
    print("deriv pipe", mode, "nHYP smearedForce ms %.3f" % ms, flush=True)
    force.free(); f.free()
print("bit-identical:", bool(np.array_equal(res["0"], res["1"])), "finite:", bool(np.isfinite(res["1"]).all()))
qb.finalize()
