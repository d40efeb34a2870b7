"""Residency sweep of the fused multi-rank kernel (QEXB200_FUSED_MINB), 32^4 per GPU."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402
from qex_b200.synth import random_su3_links, gaussian_vector  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
geoms = {2: (1, 1, 1, 2), 4: (1, 1, 2, 2), 8: (1, 2, 2, 2)}[world]
lat = tuple(32 * g for g in geoms)
qb.init(lat, geoms, rank, lr, None)
qb.p2p_attach_torch(dist)
nsites, neven = qb.local_sites()
for nlinks in (4, 8):
    s = qb.Staggered(random_su3_links(nsites, nlinks, seed=rank, scale=None if nlinks == 4 else [1, .1] * 4))
    fb, fx = qb.DeviceField(gaussian_vector(nsites, rank)), qb.DeviceField()
    for mode in (18, 12 if nlinks == 4 else 1812):
        rc, _ = s.recon(mode)
        bytes_site = sum(8 * (96 if rc[j] == 12 else 144) for j in range(nlinks // 4)) + 96
        for minb in (6,):   # other residencies need k_sweep_fused instantiated with that MINB (csrc/dslash.cu)
            os.environ["QEXB200_FUSED_MINB"] = str(minb)
            for _ in range(10):
                s.dev_stagD2(0, fx, fb, 0.0, 0.0); s.dev_stagD2(1, fx, fb, 0.0, 0.0)
            best = 1e9
            for rep in range(3):
                qb.sync(); dist.barrier(); qb.timer_start()
                for _ in range(100):
                    s.dev_stagD2(0, fx, fb, 0.0, 0.0); s.dev_stagD2(1, fx, fb, 0.0, 0.0)
                best = min(best, qb.timer_stop() / 200)
            sp = qb.SolverParams(r2req=0.0, maxits=100)
            s.dev_solveXX(fx, fb, 0.1, sp)
            qb.sync(); dist.barrier(); qb.timer_start()
            s.dev_solveXX(fx, fb, 0.1, sp)
            cg = qb.timer_stop() / 100
            if rank == 0:
                print("nlinks %d links %s minb %d: sweep %.4f ms %.0f GB/s | CG iteration %.4f ms" %
                      (nlinks, rc, minb, best, neven * bytes_site / best * 1e-6, cg), flush=True)
    fb.free(); fx.free(); s.free()
qb.finalize()
dist.destroy_process_group()
