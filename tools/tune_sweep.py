"""Launch-shape sweep of the parity-sweep kernel on one GPU (QEXB200_SWEEP_TUNE variants in csrc/dslash.cu).
The variants are compiled only with -DQEXB200_TUNE: build with `QEXB200_EXTRA_FLAGS=-DQEXB200_TUNE python -m qex_b200.build --force`."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402
from qex_b200.synth import random_su3_links, gaussian_vector  # noqa: E402

L = (32, 32, 32, 32)
qb.init(L)
vol = int(np.prod(L))
BYTES = {4: 1248, 8: 2400}
for nlinks in (4, 8):
    s = qb.Staggered(random_su3_links(vol, nlinks, seed=1))
    fb, fx = qb.DeviceField(gaussian_vector(vol)), qb.DeviceField()
    for tune in range(0, 13):
        os.environ["QEXB200_SWEEP_TUNE"] = str(tune)
        for _ in range(10):
            s.dev_stagD2(0, fx, fb, 0.0, 0.0); s.dev_stagD2(1, fx, fb, 0.0, 0.0)
        best = 1e9
        for rep in range(3):
            qb.sync(); qb.timer_start()
            for _ in range(100):
                s.dev_stagD2(0, fx, fb, 0.0, 0.0); s.dev_stagD2(1, fx, fb, 0.0, 0.0)
            best = min(best, qb.timer_stop() / 200)
        print("nlinks %d tune %2d: %.4f ms  %.0f GB/s" % (nlinks, tune, best, vol / 2 * BYTES[nlinks] / best * 1e-6), flush=True)
    fb.free(); fx.free(); s.free()
qb.finalize()
