"""Short run for ncu: a few parity sweeps (naive and fat+Naik) and a 10-iteration CG at 32^4."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402
from qex_b200.synth import random_su3_links, gaussian_vector  # noqa: E402

L = (32, 32, 32, 32)
qb.init(L)
vol = int(np.prod(L))
for nlinks in (4, 8):
    links = random_su3_links(vol, nlinks, seed=1)
    s = qb.Staggered(links)
    del links
    fb, fx = qb.DeviceField(gaussian_vector(vol)), qb.DeviceField()
    for mode in (18, 12 if nlinks == 4 else 1812):
        rc, dev = s.recon(mode)
        for _ in range(4):
            s.dev_stagD2(0, fx, fb, 0.0, 0.0)
            s.dev_stagD2(1, fx, fb, 0.0, 0.0)
        sp = qb.SolverParams(r2req=0.0, maxits=10)
        s.dev_solveXX(fx, fb, 0.1, sp)
        qb.sync()
        print("nlinks", nlinks, "links", rc, "iterations", sp.iterations, "launches", qb.launch_count())
    fb.free(); fx.free(); s.free()
qb.finalize()
