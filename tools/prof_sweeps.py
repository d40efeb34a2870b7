"""Short run for `ncu --set full`: three parity sweeps per link-storage mode at 32^4
(naive 18 / 12 real, fat+Naik 18/18 and 18/12 real)."""
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
    links = random_su3_links(vol, nlinks, seed=1, scale=None if nlinks == 4 else [1, .1, 1, .1, 1, .1, 1, .1])
    s = qb.Staggered(links)
    del links
    fb, fx = qb.DeviceField(gaussian_vector(vol)), qb.DeviceField()
    for mode in (18, 12 if nlinks == 4 else 1812):
        rc, dev = s.recon(mode)
        for i in range(3):
            s.dev_stagD2(i & 1, fx, fb, 0.0, 0.0)
        qb.sync()
        print("nlinks", nlinks, "links", rc, "launches", qb.launch_count())
    fb.free(); fx.free(); s.free()
qb.finalize()
