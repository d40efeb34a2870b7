"""Parity sweep and CG iteration at the sizes BASELINE.json quotes the metric on (one GPU): 32^4, 48^3x96, 64^3x128,
naive and fat+Naik, links generated ON THE DEVICE by the reference's recipe (per-site RngMilc6, seed 987654321,
projectSU; long links = 0.1 * SU(3) from the continued streams, benchStagProp.nim:84-92).  Prints one JSON line per
(lattice, link set): sweep time by CUDA events, bytes/site, fraction of the measured HBM peak, CG iteration time.

  python tools/big_sweep.py [--lat 48,48,48,96] [--nlinks 8] [--sweeps 100] [--iters 50] [--short]
--short: three sweeps per case only (the run ncu wraps)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qex_b200 as qb  # noqa: E402


def peak():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def make_stag(nlinks, seed=987654321):
    r = qb.newRNGField(seed)
    g = qb.DeviceGauge(nlinks=4)
    g.random(r)
    if nlinks == 4:
        g.rephase()
        s = qb.newStag(g)
        keep = [g]
    else:
        g3 = qb.DeviceGauge(nlinks=4)
        g3.random(r)
        g3.scale(0.1)
        g.rephase()
        s = qb.newStag3(g, g3)
        g.free(); g3.free()
        keep = []
    b = qb.DeviceField()
    b.gaussian(r)
    r.free()
    return s, b, keep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lat", default="32,32,32,32;48,48,48,96;64,64,64,128")
    ap.add_argument("--nlinks", default="4,8")
    ap.add_argument("--sweeps", type=int, default=100)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--mass", type=float, default=0.1)
    ap.add_argument("--short", action="store_true")
    a = ap.parse_args()
    pk = peak()
    for lat in a.lat.split(";"):
        L = tuple(int(t) for t in lat.split(","))
        qb.init(L)
        nsites, neven = qb.local_sites()
        for nlinks in (int(t) for t in a.nlinks.split(",")):
            s, b, keep = make_stag(nlinks)
            x = qb.DeviceField()
            bytes_site = nlinks * 2 * 144 + 96
            if a.short:
                for i in range(3):
                    s.dev_stagD2(i & 1, x, b, 0.0, 0.0)
                qb.sync()
                print(json.dumps({"lat": L, "nlinks": nlinks, "short": True}))
            else:
                for i in range(6):
                    s.dev_stagD2(i & 1, x, b, 0.0, 0.0)
                qb.sync()
                qb.timer_start()
                for i in range(a.sweeps):
                    s.dev_stagD2(i & 1, x, b, 0.0, 0.0)
                ms = qb.timer_stop() / a.sweeps
                gbs = 1e-9 * neven * bytes_site / (1e-3 * ms)
                # CG: a fixed number of iterations (maxits) -> time per iteration
                sp = qb.SolverParams(r2req=1e-30, maxits=5)
                s.dev_solveXX(x, b, a.mass, sp)
                qb.sync()
                sp = qb.SolverParams(r2req=1e-30, maxits=a.iters)
                qb.timer_start()
                s.dev_solveXX(x, b, a.mass, sp)
                cg_ms = qb.timer_stop() / max(1, sp.iterations)
                cg_bytes = 2 * bytes_site + 48 + 432
                print(json.dumps({"lat": L, "nlinks": nlinks, "links": "18-real", "sites_per_parity": neven,
                                  "bytes_per_site": bytes_site, "sweep_ms": ms, "sweep_gbs": gbs, "frac": gbs / pk,
                                  "sweep_gflops": 1e-9 * neven * (570 if nlinks == 4 else 1146) / (1e-3 * ms),
                                  "cg_iter_ms": cg_ms, "cg_iters_timed": sp.iterations,
                                  "cg_iter_gbs": 1e-9 * neven * cg_bytes / (1e-3 * cg_ms),
                                  "cg_iter_frac": 1e-9 * neven * cg_bytes / (1e-3 * cg_ms) / pk}), flush=True)
            x.free(); b.free(); s.free()
            for k in keep:
                k.free()
        qb.finalize()


if __name__ == "__main__":
    main()
