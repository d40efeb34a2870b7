"""Per-role timing of the one-kernel multi-rank sweep: where do the microseconds of
k_sweep_mr go on N GPUs -- packing, interior tiles, waiting for the neighbours' flags, face tiles, cross-rank sums?

  python -m torch.distributed.run --nproc-per-node N tools/role_times.py --lat 32,32,64,128 --rankgeom 1,1,2,4 [--nlinks 8]

The library stamps %globaltimer inside the kernel (qexb200_debug_role_times); this tool runs back-to-back parity
sweeps and then a fixed number of CG iterations with the stamps armed, and prints one JSON line: per rank the mean
over launches of each role's end relative to the first block's start, the flag waits, and the time inside
xrank_sum, next to the CUDA-event times of the same loops WITHOUT the stamps."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402

NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def summarise(ring, xr):
    used = ring[(ring[:, 0] != NONE) & (ring[:, 1] > 0)]
    if len(used) == 0:
        return {"launches": 0}
    t0 = used[:, 0].astype(np.int64)
    rel = lambda col: (used[:, col].astype(np.int64) - t0) * 1e-3          # us
    face = used[:, 3] != NONE
    out = {"launches": int(len(used)),
           "pack_end_us": float(rel(1).mean()),
           "interior_end_us": float(rel(2)[used[:, 2] > 0].mean()) if (used[:, 2] > 0).any() else None,
           "face_start_us": float(rel(3)[face].mean()) if face.any() else None,
           "face_end_us": float(rel(4)[face].mean()) if face.any() else None,
           "kernel_us": float(np.maximum(rel(2), rel(4)).mean()),
           "kernel_us_p90": float(np.percentile(np.maximum(rel(2), rel(4)), 90)),
           "wait_max_us": float(used[:, 5].max() * 1e-3),
           "wait_mean_of_max_us": float(used[:, 5].mean() * 1e-3),
           "face_tiles_that_waited_per_launch": float(used[:, 7].mean()),
           "wait_sum_per_launch_us": float(used[:, 6].mean() * 1e-3)}
    if xr[1] > 0:
        out["xrank_sum_mean_us"] = float(xr[0]) / float(xr[1]) * 1e-3
        out["xrank_sum_max_us"] = float(xr[2]) * 1e-3
        out["xrank_sums"] = int(xr[1])
    return out


def main():
    import torch
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--lat", default="32,32,64,128")
    ap.add_argument("--rankgeom", default="1,1,2,4")
    ap.add_argument("--nlinks", type=int, default=4)
    ap.add_argument("--recon", type=int, default=18)
    ap.add_argument("--sweeps", type=int, default=200)
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--short", action="store_true", help="six sweeps only: the run ncu wraps on rank 0")
    a = ap.parse_args()
    rank, local_rank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = tuple(int(t) for t in a.lat.split(","))
    G = tuple(int(t) for t in a.rankgeom.split(","))
    qb.init(L, G, rank, local_rank, None)
    qb.p2p_attach_torch(dist)
    r = qb.newRNGField(987654321)
    g = qb.DeviceGauge(nlinks=4)
    g.random(r)
    if a.nlinks == 8:
        g3 = qb.DeviceGauge(nlinks=4)
        g3.random(r)
        g3.scale(0.1)
        g.rephase()
        s = qb.newStag3(g, g3)
        g3.free()
    else:
        g.rephase()
        s = qb.newStag(g)
    if a.recon != 18:
        s.recon(a.recon)
    b, x = qb.DeviceField(), qb.DeviceField()
    b.gaussian(r)

    def tmax(v):
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sweeps(n):
        for i in range(n):
            s.dev_stagD2(i & 1, x, b, 0.0, 0.0)

    def cg(n):
        sp = qb.SolverParams(r2req=1e-30, maxits=n)
        s.dev_solveXX(x, b, 0.1, sp)
        return sp.iterations

    out = {"lat": L, "rankgeom": G, "nlinks": a.nlinks, "recon": a.recon, "world": world}
    if a.short:
        sweeps(6); qb.sync(); dist.barrier()
        if rank == 0:
            print(json.dumps(dict(out, short=True)))
        qb.finalize()
        dist.destroy_process_group()
        return
    sweeps(20); qb.sync(); dist.barrier()
    qb.timer_start(); sweeps(a.sweeps); out["sweep_ms_events"] = tmax(qb.timer_stop()) / a.sweeps
    cg(10); qb.sync(); dist.barrier()
    qb.timer_start(); it = cg(a.iters); out["cg_iter_ms_events"] = tmax(qb.timer_stop()) / max(1, it)
    # the same loops with the stamps armed
    dist.barrier()
    qb.role_times(True)
    sweeps(a.sweeps); qb.sync()
    ring, xr = qb.role_times(False)
    mine_sw = summarise(ring, xr)
    dist.barrier()
    qb.role_times(True)
    cg(a.iters); qb.sync()
    ring, xr = qb.role_times(False)
    mine_cg = summarise(ring, xr)
    allr = [None] * world
    dist.all_gather_object(allr, {"rank": rank, "back_to_back_sweeps": mine_sw, "cg": mine_cg})
    out["ranks"] = allr
    if rank == 0:
        print(json.dumps(out))
    qb.finalize()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
