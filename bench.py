#!/usr/bin/env python
"""Benchmark of the staggered Dslash + CG hot path (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N --steps K --warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): 32^4 naive staggered, random SU(3) links with antiperiodic-T
BC + staggered phases, m = 0.1, even-even CG to r2req = 1e-8, fp64; weak scaling keeps 32^4 per GPU
(32^4, 32^3x64, 32^2x64^2, 32x64^3 for 1/2/4/8 GPUs).

A "step" is one solveEE call (stagSolve.nim:57-138) = one CG solve on a device-resident source.
value  = QEX's solver Gflop/s, (g.len*4*72+60)*nEven*iterations / seconds (stagSolve.nim:92), whole job.
e2e    = the same through qexb200_solveXX with HOST (pinned) buffers: H2D of the source and D2H of
         the solution inside the timed region.
roofline = the dominant kernel (one parity sweep of the Dslash) timed live with CUDA events on the
         library's stream over back-to-back launches, against the measured HBM peak.
The reference arm (--impl reference) times the CPU oracle (the reference itself cannot be built in
this image: no Nim/MPI/QMP) on all host cores, on a bounded sample of the same solve.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MASS = 0.1
R2REQ = 1e-8
BYTES_PER_SITE = {4: 1248, 8: 2400}     # compulsory bytes per output site of one parity sweep
FLOPS_PER_SITE = {4: 570, 8: 1146}      # algorithmic flops per output site of one sweep
WEAK = {1: ((32, 32, 32, 32), (1, 1, 1, 1)), 2: ((32, 32, 32, 64), (1, 1, 1, 2)),
        4: ((32, 32, 64, 64), (1, 1, 2, 2)), 8: ((32, 64, 64, 64), (1, 2, 2, 2))}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: NVML polled every ~5 ms in a thread
    (nvidia-smi subprocesses are too slow for a 0.1 s region); falls back to nvidia-smi."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, dev):
        self.dev, self.sm, self.mx, self.reasons, self.stop, self.t = dev, [], 0.0, set(), False, None
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(dev)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _sample_nvml(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
            else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                          ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        o = subprocess.run(["nvidia-smi", "-i", str(self.dev), "--query-gpu=" + self.Q,
                            "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        r = [t.strip() for t in o.strip().split(",")]
        self.sm.append(float(r[0]))
        self.mx = max(self.mx, float(r[1]))
        for n, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
            if v.lower().startswith("active"):
                self.reasons.add(n)

    def _run(self):
        while not self.stop:
            try:
                if self.nv:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            time.sleep(0.005 if self.nv else 0.1)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nv else "nvidia-smi"}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def cpu_reference_leg(lat_shape, nlinks, nsolves, threads=None):
    """Times the CPU oracle's CG (oracle/qex_oracle.cpp orc_cg_xx) on a bounded sample: `nsolves`
    complete even-even solves of the same system to the same r2req.  Returns Gflop/s in QEX's flop model."""
    import oracle
    from qex_b200.synth import random_su3_links, gaussian_vector
    oracle.build()
    if threads:
        oracle.set_num_threads(threads)
    lat = oracle.Lattice(lat_shape)
    g = np.ascontiguousarray(np.stack(random_su3_links(lat.vol, 4, seed=1)))
    lat.setBC(g)
    lat.stagPhase(g)
    if nlinks == 8:
        g3 = np.stack(random_su3_links(lat.vol, 4, seed=2, scale=[0.1] * 4))
        gg = np.zeros((8,) + g.shape[1:])
        gg[0::2], gg[1::2] = g, g3
        g = gg
    b = gaussian_vector(lat.vol, 0)
    x = lat.vector()
    lat.cg_xx(g, x, b, MASS * MASS, 0, 0.0, 2)      # warm the caches / thread pool
    its = 0
    t0 = time.perf_counter()
    for _ in range(nsolves):
        x[:] = 0
        it, r2, b2 = lat.cg_xx(g, x, b, MASS * MASS, 0, R2REQ, 20000)
        its += it
    dt = time.perf_counter() - t0
    flops = (nlinks * 4 * 72 + 60) * lat.half * its
    return {"value": 1e-9 * flops / dt, "unit": "GFLOP/s", "cores": oracle.num_threads(), "kind": "port",
            "sample": "%d complete solves (%d CG iterations) of the %s %s even-even system to r2req=1e-8 "
                      "(oracle C++/OpenMP port, %.1f s)"
                      % (nsolves, its, "x".join(map(str, lat_shape)), "naive" if nlinks == 4 else "fat+Naik", dt),
            "seconds": dt, "iterations": its, "ms_per_solve": 1e3 * dt / nsolves}


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path.  QEX itself (Nim + MPI + QMP)
    cannot be built in this image, so this is the oracle port, with all host threads; one step = one
    complete solve of the same system."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    lat_shape = WEAK[1][0]
    if args.warmup > 0:
        cpu_reference_leg(lat_shape, 4, 1)
    res = cpu_reference_leg(lat_shape, 4, max(1, args.steps))
    v = res["value"]
    line = {"impl": "reference", "metric": "staggered CG solver throughput (Dslash GFLOP/s, QEX flop model)",
            "value": v, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": res["ms_per_solve"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "32x32x32x32 naive staggered even-even CG to r2req=1e-8, m=0.1, fp64, random SU(3) "
                                   "(BASELINE configs[1])", "cg_iterations": res["iterations"] // max(1, args.steps),
                       "time_to_solution_ms": res["ms_per_solve"]},
            "cpu_baseline": res,
            "e2e": {"value": v, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    import qex_b200 as qb
    from qex_b200.synth import random_su3_links, gaussian_vector

    rank, local_rank, world = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d" % args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        if args.transport == "nccl":
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt = torch.tensor(list(qb.nccl_unique_id()), dtype=torch.uint8, device="cuda")
            dist.broadcast(idt, 0)
            nccl_id = bytes(idt.cpu().tolist())
    lat_shape, rankgeom = WEAK[world]
    if world > 1:   # room in the peer arena for the nHYP force state (52 fields) and its work fields (28)
        os.environ.setdefault("QEXB200_ARENA_FIELDS", "128")
    qb.init(lat_shape, rankgeom, rank, local_rank, nccl_id)
    if world > 1 and args.transport == "p2p":
        qb.p2p_attach_torch(dist)     # halos and scalar sums over NVLink stores (csrc/comm.cu)
    nsites, neven = qb.local_sites()
    peak, peak_src = measured_peak()

    def barrier():
        qb.sync()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxreduce(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    results = {}
    for nlinks in ((4, 8) if not args.naive_only else (4,)):
        links = random_su3_links(nsites, 4, seed=1 + 17 * rank)
        if nlinks == 4:
            s = qb.Staggered(links)
            s.rephase()                      # setBC + stagPhase on the device copy
        else:
            # synthetic 3-hop operator of the reference's own bench (benchStagProp.nim:84-92):
            # fat = phased SU(3), long = 0.1 * SU(3)
            long_links = random_su3_links(nsites, 4, seed=2 + 17 * rank, scale=[0.1] * 4)
            s = qb.newStag3(phase_on_host(links, lat_shape, rankgeom, rank), long_links)
            del long_links
        del links
        # source in pinned host memory (e2e leg copies from it every step)
        hb = torch.from_numpy(gaussian_vector(nsites, seed=rank)).contiguous().pin_memory()
        hx = torch.zeros_like(hb).pin_memory()
        b_np, x_np = hb.numpy(), hx.numpy()
        fb, fx = qb.DeviceField(b_np), qb.DeviceField()
        # link storage modes: the requested one first (headline), then the 18-real form.  For the
        # fat+Naik set only the Naik class is compressed (1812): real HISQ fat links are not unitary.
        modes = [18] if args.recon == 18 else ([12, 18] if nlinks == 4 else [1812, 18])
        for mode in modes:
            got_rc, dev = s.recon(mode)
            nclass = nlinks // 4
            bytes_site = sum(8 * (96 if got_rc[j] == 12 else 144) for j in range(nclass)) + 96
            key = (nlinks, mode)

            # ---- value: K solves, source resident in HBM --------------------------------
            def one_solve():
                sp = qb.SolverParams(r2req=R2REQ, maxits=20000)
                s.dev_solveXX(fx, fb, MASS, sp, parEven=True)
                return sp
            for _ in range(args.warmup):
                sp = one_solve()
            barrier()
            qb.launch_count(reset=True)
            with ClockSampler(local_rank) as cs:
                qb.timer_start()
                for _ in range(args.steps):
                    sp = one_solve()
                ms = qb.timer_stop()
            barrier()
            launches = qb.launch_count()
            ms = maxreduce(ms)
            iters = sp.iterations
            neven_global = neven * world
            flops_step = (nlinks * 4 * 72 + 60) * neven_global * iters
            gflops = 1e-9 * flops_step * args.steps / (1e-3 * ms)

            # ---- e2e: same solve through the host-buffer C ABI -------------------------
            for _ in range(max(1, args.warmup // 2)):
                spe = qb.SolverParams(r2req=R2REQ, maxits=20000)
                s.solveXX(x_np, b_np, MASS, spe, parEven=True)
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                spe = qb.SolverParams(r2req=R2REQ, maxits=20000)
                s.solveXX(x_np, b_np, MASS, spe, parEven=True)
            barrier()
            e2e_s = maxreduce(time.perf_counter() - t0)
            e2e_gflops = 1e-9 * (nlinks * 4 * 72 + 60) * neven_global * spe.iterations * args.steps / e2e_s

            # ---- roofline of the dominant kernel: one parity sweep ---------------------
            nrep = args.sweeps
            for _ in range(10):
                s.dev_stagD2(0, fx, fb, 0.0, 0.0)
                s.dev_stagD2(1, fx, fb, 0.0, 0.0)
            barrier()
            qb.timer_start()
            for _ in range(nrep // 2):
                s.dev_stagD2(0, fx, fb, 0.0, 0.0)
                s.dev_stagD2(1, fx, fb, 0.0, 0.0)
            sw_ms = maxreduce(qb.timer_stop()) / (2 * (nrep // 2))
            barrier()
            achieved = neven * bytes_site / (1e-3 * sw_ms) * 1e-9
            cg_bytes = 2 * bytes_site + 48 + 432      # 2 sweeps + 4m^2 x read + fused BLAS-1
            results[key] = {
                "gflops": gflops, "ms_per_step": ms / args.steps, "iterations": iters, "r2": sp.r2,
                "launches": launches, "clocks": cs.summary(), "recon": got_rc[:nclass], "recon_dev": dev[:nclass],
                "e2e_gflops": e2e_gflops, "e2e_ms_per_step": 1e3 * e2e_s / args.steps,
                "bytes_site": bytes_site, "sweep_ms": sw_ms, "sweep_gbs": achieved, "sweep_frac": achieved / peak,
                "sweep_gflops": 1e-9 * neven * FLOPS_PER_SITE[nlinks] / (1e-3 * sw_ms),
                "cg_iter_ms": ms / args.steps / max(1, iters),
                "cg_iter_gbs": 1e-9 * neven * cg_bytes / (1e-3 * ms / args.steps / max(1, iters)),
            }
        fb.free()
        fx.free()
        s.free()

    def kernel_name(nlinks, rc):
        return "k_sweep%s<NL=%d, links %s real>" % ("_fused" if world > 1 and args.transport == "p2p" else "", nlinks // 4,
                                                     "/".join(str(t) for t in rc))

    def roof(rr, nlinks):
        tr = ncu_traffic(nlinks, rr["recon"]) if world == 1 else None
        return {"bound": "hbm", "kernel": kernel_name(nlinks, rr["recon"]) + " (one parity sweep of the Dslash)",
                "achieved": rr["sweep_gbs"], "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": rr["sweep_frac"], "traffic": tr, "bytes_per_site": rr["bytes_site"],
                "algorithmic_bytes_per_launch": rr["bytes_site"] * neven, "ms_per_launch": rr["sweep_ms"],
                "gflops": rr["sweep_gflops"], "cg_iteration_gbs": rr["cg_iter_gbs"],
                "cg_iteration_frac": rr["cg_iter_gbs"] / peak}

    head_mode = 18 if args.recon == 18 else 12
    r = results[(4, head_mode)]
    line = {
        "metric": "staggered CG solver throughput (Dslash GFLOP/s, QEX flop model)",
        "value": r["gflops"], "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s naive staggered even-even CG to r2req=1e-8, m=0.1, fp64, random SU(3) "
                               "(BASELINE configs[1]; 32^4 per GPU)" % "x".join(map(str, lat_shape)),
                   "rankgeom": list(rankgeom), "transport": (args.transport if world > 1 else "none"),
                   "link_storage": "%s-real" % "/".join(str(t) for t in r["recon"]), "cg_iterations": r["iterations"],
                   "l2": "links (403-604 MB per sweep) exceed the 126 MB L2; CG vectors (25 MB each) are L2-resident by design",
                   "time_to_solution_ms": r["ms_per_step"]},
        "e2e": {"value": r["e2e_gflops"], "unit": "GFLOP/s", "h2d_bytes_per_step": int(nsites * 48),
                "d2h_bytes_per_step": int(nsites * 48), "ms_per_step": r["e2e_ms_per_step"]},
        "gpu_launches": r["launches"],
        "clocks": r["clocks"],
        "roofline": roof(r, 4),
    }
    variants = {}
    for (nlinks, mode), rr in results.items():
        if (nlinks, mode) == (4, head_mode):
            continue
        name = ("naive" if nlinks == 4 else "hisq_fat_naik") + "_links%s" % "_".join(str(t) for t in rr["recon"])
        variants[name] = {
            "workload": "same lattice, " + ("naive one-hop links" if nlinks == 4 else
                                           "fat + 0.1*Naik links (benchStagProp.nim:84-92)"),
            "cg_gflops": rr["gflops"], "time_to_solution_ms": rr["ms_per_step"], "cg_iterations": rr["iterations"],
            "e2e_gflops": rr["e2e_gflops"], "roofline": roof(rr, nlinks)}
    line["variants"] = variants
    if world == 1 or args.transport == "p2p":
        # on several GPUs: link-shaped fields in the peer-mapped arena, neighbours read over NVLink (csrc/smear.cu)
        try:
            lp = link_pipeline_leg(qb, nsites, peak, rank, cpu=(world == 1 and not args.no_cpu), maxreduce_fn=maxreduce)
            for k in ("hisq", "nhyp", "nhyp_force", "hisq_force"):
                lp[k]["ms"] = maxreduce(lp[k]["ms"])
                gbs = 1e-9 * nsites * lp[k]["algorithmic_bytes_per_site"] / (1e-3 * lp[k]["ms"])
                lp[k]["achieved_gbs"], lp[k]["frac_of_hbm_peak"] = gbs, gbs / peak
            lp["sites"] = nsites * world
            lp["unit"] = "ms per smear of the whole lattice (max over ranks); GB/s per GPU"
            line["link_pipeline"] = lp
        except Exception as e:   # never lose the headline over the side leg
            line["link_pipeline"] = {"error": str(e)[:300]}
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_reference_leg(lat_shape, 4, args.ref_solves)
    else:
        line["cpu_baseline"] = None
    qb.finalize()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))


# Compulsory HBM traffic of the device link pipeline in matrix passes (144 B per site each): every
# distinct field a launch reads counted once, every field it writes once.
HYP_PASSES = 12 * 3 + 12 * 6 + 4 * 8
HYP_FORCE_PASSES = 24 + 4 * 4 + 12 * 7 + 12 * 5 + 24 * 7 + 12 * 5 + 12 * 7     # smearedForce alone
# fat7lDeriv, passes per direction sig: 3 staples (3) + 3 zeroings; per rho: stpl5 (6) + zero + deriv (7) + axpy (3) +
# axpby (3) + 2 derivs (14); 3-staple part: zero + 3 x (scale 2 + axpy 3 + deriv 7) + 2 axpy (6)
FAT7_DERIV_PASSES = 4 * (12 + 3 * 34 + 1 + 3 * 12 + 6)
HISQ_FORCE_PASSES = 2 * FAT7_DERIV_PASSES - 4 * 3 * 3 + 16 + 8 + 4 * 3 + 4 * 3
HISQ_PASSES = 4 * (3 * (5 + 10 + 6) - 1) + 4 * (3 * (5 + 10 + 8) - 1) + 4 * 2 + 4 * 2


def link_pipeline_leg(qb, nsites, peak, rank, cpu=True, reps=3, maxreduce_fn=lambda v: v):
    """HisqCoefs.smear and HypCoefs.smear on resident links (what an HMC force step does before
    its solves), timed with CUDA events on the library's stream; the oracle's CPU time per site beside."""
    from qex_b200.synth import random_su3_links
    g = qb.DeviceGauge(random_su3_links(nsites, 4, seed=5 + rank))
    fl, ll = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    hc, yc = qb.HisqCoefs().init(), qb.HypCoefs(0.4, 0.5, 0.5)
    out = {"unit": "ms per smear of the whole lattice", "sites": nsites}
    for name, fn, passes in (("hisq", lambda: hc.smear(g, fl, ll), HISQ_PASSES), ("nhyp", lambda: yc.smear(g, fl), HYP_PASSES)):
        fn()
        qb.sync()
        qb.launch_count(reset=True)
        qb.timer_start()
        for _ in range(reps):
            fn()
        ms = qb.timer_stop() / reps
        gbs = 1e-9 * nsites * passes * 144 / (1e-3 * ms)
        out[name] = {"ms": ms, "launches": qb.launch_count() // reps, "algorithmic_bytes_per_site": passes * 144,
                     "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / peak}
    # smearGetForce + smearedForce (hypsmear.nim:49-262): what every nHYP force step adds after the solves
    force = yc.smearGetForce(g, fl)
    ll.zero()
    for it in range(reps + 1):
        if it == 1:
            qb.sync(); qb.launch_count(reset=True); qb.timer_start()
        force(ll, fl)
    ms = qb.timer_stop() / reps
    gbs = 1e-9 * nsites * HYP_FORCE_PASSES * 144 / (1e-3 * ms)
    out["nhyp_force"] = {"ms": ms, "launches": qb.launch_count() // reps, "algorithmic_bytes_per_site": HYP_FORCE_PASSES * 144,
                         "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / peak}
    force.free()
    # HISQ: smearGetForce + smearedForce (hisqsmear.nim:51-84): second fat7 + Naik derivative, projectUderiv, first fat7
    hforce = hc.smearGetForce(g, fl, ll)
    du, c1, c2 = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    du.zero(); c1.projectU(fl); c2.projectU(fl)          # any non-trivial chain fields will do for timing
    for it in range(reps + 1):
        if it == 1:
            qb.sync(); qb.launch_count(reset=True); qb.timer_start()
        hforce(du, c1, c2)
    ms = qb.timer_stop() / reps
    gbs = 1e-9 * nsites * HISQ_FORCE_PASSES * 144 / (1e-3 * ms)
    out["hisq_force"] = {"ms": ms, "launches": qb.launch_count() // reps, "algorithmic_bytes_per_site": HISQ_FORCE_PASSES * 144,
                         "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / peak}
    hforce.free()
    for d in (du, c1, c2):
        d.free()
    # One massive-fermion force of staghmc_spv.nim on resident fields (the sequence BASELINE configs[3] repeats per
    # force step): smearGetForce -> setBC/stagPhase -> solve D psi = phi (f_tol 1e-12, m = 0.1) -> psi (x) psi(+mu)^+ ->
    # rephase -> smearedForce in place -> projTAH.  Stage times with CUDA events (each stage synchronised).
    from qex_b200.synth import gaussian_vector
    dphi, dpsi, frc = qb.DeviceField(gaussian_vector(nsites, 7 + rank)), qb.DeviceField(), qb.DeviceGauge(nlinks=4)
    stages = {}

    def timed(name, fn):
        qb.timer_start()
        r = fn()
        stages[name] = stages.get(name, 0.0) + qb.timer_stop()
        return r
    # nHYP links are U(3): rows 0,1 + the determinant factor (14 reals) instead of 18 for the solve's sweeps;
    # sticky, re-derived (and re-verified on the device) inside smearGetForce and rephase below
    fl.recon(14)
    for it in range(2):
        stages.clear()
        force = timed("smearGetForce", lambda: yc.smearGetForce(g, fl))
        timed("rephase", fl.rephase)
        sop = qb.newStag(fl)
        sp = qb.SolverParams(r2req=1e-12, maxits=20000)
        timed("solve", lambda: sop.dev_solve(dpsi, dphi, MASS, sp))
        timed("outer_products", lambda: (frc.outer(dpsi, dpsi, -0.25, 0.25, accumulate=False), frc.rephase()))
        timed("smearedForce", lambda: force(frc, frc))
        timed("projTAH", lambda: frc.projTAH(g))
        force.free()
    stages = {k: maxreduce_fn(v) for k, v in stages.items()}
    storage = fl.recon(14)[0][0]      # what the solve above read (18 on a split lattice: not enabled there yet)
    out["hmc_force_step"] = {"workload": "one massive-fermion force step of staghmc_spv on nHYP links, all on the device",
                             "ms": sum(stages.values()), "stages_ms": stages, "cg_iterations": sp.iterations, "r2": sp.r2,
                             "link_storage": "%d-real" % storage}
    fl.recon(18)
    for d in (dphi, dpsi, frc):
        d.free()
    yc.smear(g, fl)
    out["plaq_smeared"] = [float(t) for t in fl.plaq()]
    for d in (g, fl, ll):
        d.free()
    if cpu:
        import oracle
        Ls = (16, 16, 16, 16)
        lat = oracle.Lattice(Ls)
        gg = lat.random_gauge(lat.new_rng(1))
        t0 = time.perf_counter(); lat.hisq_smear(gg, oracle.hisq_coefs()); t1 = time.perf_counter()
        lat.hyp_smear(gg, (0.4, 0.5, 0.5)); t2 = time.perf_counter()
        scale = nsites / lat.vol
        out["cpu_port"] = {"sample": "oracle (C++/OpenMP restatement) on 16^4, scaled by sites to this lattice",
                           "cores": oracle.num_threads(), "hisq_ms": 1e3 * (t1 - t0) * scale,
                           "nhyp_ms": 1e3 * (t2 - t1) * scale}
    return out


def ncu_traffic(nlinks, recon):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the 32^4 parity sweep, from the
    committed `ncu --set full` capture (profiles/ncu_traffic.json); None if that variant was not captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    key = "nl%d_%s" % (nlinks // 4, "_".join(str(t) for t in recon))
    return json.load(open(p)).get(key)


def phase_on_host(links, lat_shape, rankgeom, rank):
    """setBC + stagPhase (stagD.nim:72-80,509-522) with numpy on V=1 host links of this rank."""
    G, lat = rankgeom, lat_shape
    X = [lat[d] // G[d] for d in range(4)]
    rc, r = [0] * 4, rank
    for d in (3, 2, 1, 0):
        rc[d] = r % G[d]
        r //= G[d]
    vol = int(np.prod(X))
    lex = np.arange(vol)
    x0 = lex % X[0]
    x1 = (lex // X[0]) % X[1]
    x2 = (lex // (X[0] * X[1])) % X[2]
    x3 = lex // (X[0] * X[1] * X[2])
    par = (x0 + x1 + x2 + x3) & 1
    idx = np.where(par == 1, (lex + vol) // 2, lex // 2)
    gx = [x0 + rc[0] * X[0], x1 + rc[1] * X[1], x2 + rc[2] * X[2], x3 + rc[3] * X[3]]
    flips = [gx[3] & 1, (gx[0] + gx[3]) & 1, (gx[0] + gx[1] + gx[3]) & 1, (gx[3] == lat[3] - 1).astype(int)]
    out = []
    for mu in range(4):
        sgn = np.ones(vol)
        sgn[idx] = 1.0 - 2.0 * flips[mu]
        out.append(np.ascontiguousarray(links[mu] * sgn[:, None, None, None]))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sweeps", type=int, default=200, help="back-to-back parity sweeps for the roofline leg")
    ap.add_argument("--ref-solves", type=int, default=8, help="complete CPU solves in the bounded cpu_baseline sample")
    ap.add_argument("--naive-only", action="store_true")
    ap.add_argument("--recon", type=int, default=12, choices=[12, 18],
                    help="link storage of the headline run: 12 = rows 0,1 + rebuilt third row where the links allow it")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU halo/sum transport: peer-memory NVLink stores (default) or NCCL send/recv + allreduce")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
