"""Diagnostic for the peer-memory sweep: growing local volumes, back-to-back sweeps, short timeout.
Launch: QEXB200_P2P_TIMEOUT_S=3 torchrun --nproc-per-node 2 tools/p2p_diag.py"""
import os
import sys
import time

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
for L in (8, 16, 24, 32):
    lat = tuple(L * g for g in geoms)
    qb.init(lat, geoms, rank, lr, None)
    qb.p2p_attach_torch(dist)
    nsites, neven = qb.local_sites()
    for nlinks in (4, 8):
        s = qb.Staggered(random_su3_links(nsites, nlinks, seed=rank))
        fb, fx = qb.DeviceField(gaussian_vector(nsites, rank)), qb.DeviceField()
        for name, fn in (("stagD2", lambda: (s.dev_stagD2(0, fx, fb, 0.0, 0.0), s.dev_stagD2(1, fx, fb, 0.0, 0.0))),
                         ("stagD2ee", lambda: s.dev_stagD2xx(0, fx, fb, 0.01)),
                         ("cg20", lambda: s.dev_solveXX(fx, fb, 0.1, qb.SolverParams(r2req=0.0, maxits=20)))):
            t0 = time.time()
            try:
                for _ in range(20 if name != "cg20" else 2):
                    fn()
                qb.sync()
                msg = "ok"
            except qb.QexB200Error as e:
                msg = "FAIL " + str(e)
            print("rank %d L=%d nlinks=%d %s: %s (%.2fs)" % (rank, L, nlinks, name, msg, time.time() - t0), flush=True)
            if msg != "ok":
                dist.barrier()
                os._exit(1)
        fb.free(); fx.free(); s.free()
    qb.finalize()
    dist.barrier()
dist.destroy_process_group()
