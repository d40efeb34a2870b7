#!/usr/bin/env python
"""Benchmark of the staggered Dslash + CG hot path (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--config c2|c3|c4|c5] [--overlap off] [--transport nccl]

Default workload c2 = BASELINE.json configs[1]: 32^4 naive staggered per GPU (weak series 32^4, 32^3x64 on 1.1.1.2,
32^2x64^2 on 1.1.2.2, 32^2x64x128 on 1.1.2.4; --alt8: 32x64^3 on 1.2.2.2), inputs by the reference's own recipe
generated on the device (per-site RngMilc6 seed 987654321, g.random = gaussian + projectSU, setBC + stagPhase, gaussian
source), m = 0.1, even-even CG to r2req = 1e-8, fp64.  c3 / c5 = 48^3x96 / 64^3x128 fat+Naik (strong scaling),
c4 = the solve sequence of one staghmc_spv trajectory at 64^4 on nHYP links.

A "step" is one solveEE call (stagSolve.nim:57-138) = one CG solve on a device-resident source.
value  = QEX's solver Gflop/s, (g.len*4*72+60)*nEven*iterations / seconds (stagSolve.nim:92), whole job, with the link
         storage the library selects for these links (18-real); `variants` = after reunit with 12-real storage, fat+Naik.
e2e    = the same through qexb200_solveXX with HOST (pinned) buffers: H2D of the source and D2H of the solution inside the
         timed region (gauge handle resident); e2e_cold adds create + upload + re-layout of the links per step.
roofline = the dominant kernel (one parity sweep of the Dslash) timed live with CUDA events on the library's stream over
         back-to-back launches, against the measured HBM peak; roofline_target = the north_star config (fat+Naik 18/18).
parity = preflight before anything is timed: a small lattice on the same rank grid against the oracle (checker only);
true_residual = |b - A x|^2 / |b|^2 of the timed solve from two public stagD2 calls.
The reference arm (--impl reference) times the CPU oracle (the reference itself cannot be built in this image: no
Nim/MPI/QMP) with all host threads on the same global lattice and inputs the GPU arm uses at that N.
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


# ---- workloads: SURVEY.md 8d numbering (config 2 = BASELINE.json configs[1], ...) -----------------------------
GEOM = {1: (1, 1, 1, 1), 2: (1, 1, 1, 2), 4: (1, 1, 2, 2), 8: (1, 1, 2, 4)}   # split the slowest dims
GEOM_ALT8 = (1, 2, 2, 2)                                                       # the 32x64^3 series of SURVEY 8d


def workload(cfg, world, alt8=False):
    """(global lattice, rank geometry, link classes, scaling, description) of a config at `world` GPUs."""
    G = GEOM_ALT8 if (alt8 and world == 8) else GEOM[world]
    if cfg == "c2":      # weak series, 32^4 per GPU
        lat = tuple(32 * G[d] for d in range(4))
        return lat, G, 4, "weak", ("%s naive staggered even-even CG to r2req=1e-8, m=0.1, fp64, QEX-recipe random SU(3) links "
                                   "(RngMilc6 seed 987654321, projectSU, setBC, stagPhase; BASELINE configs[1], 32^4 per GPU)"
                                   % "x".join(map(str, lat)))
    if cfg == "c3":
        lat = (48, 48, 48, 96)
        return lat, G, 8, "strong", ("48x48x48x96 HISQ-shaped fat+Naik staggered even-even CG to r2req=1e-8, m=0.1, fp64, fat = phased "
                                     "random SU(3), long = 0.1*SU(3) (benchStagProp.nim:84-92; BASELINE configs[2])")
    if cfg == "c5":
        lat = (64, 64, 64, 128)
        return lat, G, 8, "strong", ("64x64x64x128 HISQ-shaped fat+Naik staggered even-even CG to r2req=1e-8, m=0.1, fp64, fat = phased "
                                     "random SU(3), long = 0.1*SU(3) (benchStagProp.nim:84-92; BASELINE configs[4])")
    if cfg == "c4":
        lat = (64, 64, 64, 64)
        return lat, G, 4, "strong", ("64x64x64x64 nHYP-smeared staggered: the solve sequence of one staghmc_spv trajectory "
                                     "(Nf=2 fields m=0.1, 2 Pauli-Villars fields m_pv=0.75, 10 force steps; a_tol=1e-20, "
                                     "f_tol=1e-12, input_hmc.xml:86-88; BASELINE configs[3])")
    raise SystemExit("unknown config " + cfg)


def cpu_reference_leg(lat_shape, nlinks, nsolves, threads=None, qex_links=True, maxits=20000):
    """Times the CPU oracle's CG (oracle/qex_oracle.cpp orc_cg_xx; the reference itself cannot be built here) on
    complete even-even solves of the system the GPU arm solves: same lattice, same QEX-recipe links and source,
    same r2req.  Returns Gflop/s in QEX's flop model."""
    import oracle
    oracle.build()
    oracle.set_num_threads(threads or os.cpu_count())     # torchrun exports OMP_NUM_THREADS=1: undo it
    lat = oracle.Lattice(lat_shape)
    t0 = time.perf_counter()
    if qex_links:
        rng = lat.new_rng(987654321)
        g = lat.random_gauge(rng, 4)
    else:
        from qex_b200.synth import random_su3_links
        g = np.ascontiguousarray(np.stack(random_su3_links(lat.vol, 4, seed=1)))
    lat.setBC(g)
    lat.stagPhase(g)
    if nlinks == 8:
        if qex_links:
            g3 = lat.random_gauge(rng, 4) * 0.1
        else:
            g3 = np.stack(random_su3_links(lat.vol, 4, seed=2, scale=[0.1] * 4))
        gg = np.zeros((8,) + g.shape[1:])
        gg[0::2], gg[1::2] = g, g3
        g = gg
    if qex_links:
        b = lat.gaussian_vector(rng)
    else:
        from qex_b200.synth import gaussian_vector
        b = gaussian_vector(lat.vol, 0)
    setup = time.perf_counter() - t0
    x = lat.vector()
    lat.cg_xx(g, x, b, MASS * MASS, 0, 0.0, 2)      # warm the caches / thread pool
    its = 0
    t0 = time.perf_counter()
    for _ in range(nsolves):
        x[:] = 0
        it, r2, b2 = lat.cg_xx(g, x, b, MASS * MASS, 0, R2REQ, maxits)
        its += it
    dt = time.perf_counter() - t0
    flops = (nlinks * 4 * 72 + 60) * lat.half * its
    return {"value": 1e-9 * flops / dt, "unit": "GFLOP/s", "cores": oracle.num_threads(), "kind": "port",
            "sample": "%d even-even CG solve(s) (%d iterations%s) of the %s %s system to r2req=1e-8 (oracle C++/OpenMP "
                      "port, %.1f s; inputs generated in %.1f s, untimed)"
                      % (nsolves, its, "" if maxits >= 20000 else ", capped at %d per solve" % maxits,
                         "x".join(map(str, lat_shape)), "naive" if nlinks == 4 else "fat+Naik", dt, setup),
            "seconds": dt, "iterations": its, "ms_per_solve": 1e3 * dt / nsolves}


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path.  QEX itself (Nim + MPI + QMP) cannot be
    built in this image, so this is the oracle port with all host threads, on the SAME global lattice the GPU arm
    solves at this N (config.workload strings are identical); one step = one complete solve.  For the large
    lattices (c3/c5, and c2 at 8 GPUs) a step is a solve capped at --ref-maxits iterations: GFLOP/s per iteration
    is what the metric compares, and a complete 48^3x96 CPU solve would take minutes."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    world = max(world, args.gpus)
    cfg = args.config
    lat_shape, G, nlinks, scaling, wl = workload(cfg, world, args.alt8)
    vol = int(np.prod(lat_shape))
    cap = 20000 if vol <= 2 * 32 ** 4 else args.ref_maxits
    if cfg == "c4":
        print(json.dumps({"impl": "reference", "unavailable": "config c4 (trajectory solve sequence) has no CPU arm: run c2/c3/c5"}))
        return
    nsteps = max(1, args.steps) if vol <= 2 * 32 ** 4 else max(1, min(args.steps, 3))
    if args.warmup > 0 and vol <= 32 ** 4:
        cpu_reference_leg(lat_shape, nlinks, 1, maxits=cap)
    res = cpu_reference_leg(lat_shape, nlinks, nsteps, maxits=cap)
    v = res["value"]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": res["ms_per_solve"], "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(wl, lat_shape, G, nlinks),
            "solve": {"cg_iterations": res["iterations"] // nsteps, "time_to_solution_ms": res["ms_per_solve"], "solves_timed": nsteps},
            "cpu_baseline": res,
            "e2e": {"value": v, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            # what was actually timed: on the large lattices `steps` is the requested label, the
            # sample is `solves_timed` solves each capped at `capped_at` iterations (GFLOP/s per iteration compares)
            "solves_timed": nsteps, "capped_at": (cap if cap < 20000 else None)}
    print(json.dumps(line))


METRIC = "staggered CG solver throughput (Dslash GFLOP/s, QEX flop model)"


def config_dict(wl, lat_shape, rankgeom, nlinks):
    """The same dict in both arms (the driver compares them): purely descriptive, nothing measured."""
    vol_rank = int(np.prod(lat_shape)) // int(np.prod(rankgeom))
    return {"workload": wl, "lattice": list(lat_shape), "rankgeom": list(rankgeom), "mass": MASS, "r2req": R2REQ,
            "l2": "no flush needed: the links of one parity sweep (%.0f MB per GPU, 18-real) exceed the 126 MB L2"
                  % (1e-6 * (vol_rank // 2) * nlinks * 2 * 144)}


def preflight_parity(qb, dist, rank, local_rank, world, G, transport, nccl_id_fn):
    """Before the timed region: a small lattice split over the SAME rank geometry, every rank against its block of
    the oracle's global result (the oracle is the checker here, never the thing measured): Dslash rel-L2, CG
    iterations, true residual, for the naive and the fat+Naik link sets."""
    import oracle
    L = tuple(8 * G[d] for d in range(4))
    lat = oracle.Lattice(L)
    lo = oracle.QexLayout(L, G, (1, 1, 1, 1))
    mine = lambda f: lo.scatter(f)[rank]
    rng = lat.new_rng(987654321)
    g = lat.random_gauge(rng, 4)
    lat.setBC(g); lat.stagPhase(g)
    g3 = lat.random_gauge(rng, 4) * 0.1
    gg = np.zeros((8,) + g.shape[1:]); gg[0::2], gg[1::2] = g, g3
    b = lat.gaussian_vector(rng)
    qb.init(L, G, rank, local_rank, nccl_id_fn() if (world > 1 and transport == "nccl") else None)
    if world > 1 and transport == "p2p":
        qb.p2p_attach_torch(dist)
    out = {"lattice": list(L), "rankgeom": list(G), "checker": "oracle (CPU restatement), every rank its own block"}
    worst = 0.0
    for name, links in (("naive", g), ("fat_naik", gg)):
        nl = links.shape[0]
        s = qb.Staggered([mine(links[k]) for k in range(nl)])
        want = lat.vector()
        lat.D(links, want, b, MASS)
        got = np.zeros_like(mine(b))
        s.D(got, mine(b), MASS)
        e = float(np.linalg.norm((got - mine(want)).ravel()) / np.linalg.norm(mine(want).ravel()))
        xo = lat.vector()
        it_o, _ = lat.solveXX(links, xo, b, MASS, 0, 1e-10, 10000)
        x = np.zeros_like(mine(b))
        sp = qb.SolverParams(r2req=1e-10, maxits=10000)
        s.solveEE(x, mine(b), MASS, sp)
        # true residual of the device solution with the ORACLE's operator needs the global x: compare blocks instead
        ex = float(np.linalg.norm((x - mine(xo)).ravel()) / max(np.linalg.norm(mine(xo).ravel()), 1e-300))
        out[name] = {"dslash_rel_l2": e, "cg_iters": [int(sp.iterations), int(it_o)], "solution_rel_l2": ex, "r2": sp.r2}
        worst = max(worst, e)
        s.free()
    qb.finalize()
    if world > 1:
        import torch
        t = torch.tensor([worst] + [out[n][k] for n in ("naive", "fat_naik") for k in ("dslash_rel_l2", "solution_rel_l2")],
                         dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        v = t.tolist()
        out["naive"]["dslash_rel_l2"], out["naive"]["solution_rel_l2"] = v[1], v[2]
        out["fat_naik"]["dslash_rel_l2"], out["fat_naik"]["solution_rel_l2"] = v[3], v[4]
        worst = v[0]
    out["dslash_rel_l2"] = worst
    out["cg_iters"] = out["naive"]["cg_iters"]
    out["ok"] = bool(worst <= 1e-13 and all(abs(out[n]["cg_iters"][0] - out[n]["cg_iters"][1]) <= 1 and
                                            out[n]["solution_rel_l2"] <= 1e-4 for n in ("naive", "fat_naik")))
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    import qex_b200 as qb

    rank, local_rank, world = dist_env()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d" % args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def nccl_id():
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.tensor(list(qb.nccl_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
        return bytes(idt.cpu().tolist())

    cfg = args.config
    lat_shape, rankgeom, nl_cfg, scaling, wl = workload(cfg, world, args.alt8)
    if args.overlap == "off":
        os.environ["QEXB200_OVERLAP"] = "0"
    parity = None
    if not args.no_parity:
        parity = preflight_parity(qb, dist if world > 1 else None, rank, local_rank, world, rankgeom, args.transport, nccl_id)
    if world > 1:   # room in the peer arena for the nHYP force state (52 fields) and its work fields (28)
        os.environ.setdefault("QEXB200_ARENA_FIELDS", "128" if cfg == "c2" else ("48" if cfg == "c4" else "24"))
    qb.init(lat_shape, rankgeom, rank, local_rank, nccl_id() if (world > 1 and args.transport == "nccl") else None)
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

    if cfg == "c4":
        line = trajectory_leg(args, qb, wl, lat_shape, rankgeom, world, rank, local_rank, nsites, neven, peak, peak_src,
                              barrier, maxreduce)
        line["parity"] = parity
        qb.finalize()
        if world > 1:
            dist.destroy_process_group()
        if rank == 0:
            print(json.dumps(line), flush=True)
        bad = verdict_of_checks(parity, [], args.no_parity)
        if bad:
            sys.stderr.write("bench.py: FAILED checks: %s\n" % "; ".join(bad))
            sys.exit(3)
        return

    # ---- inputs by the reference's recipe, generated on the device: per-site RngMilc6 seeded with the
    # global lexicographic index, g.random = gaussian + projectSU, setBC + stagPhase, long links = 0.1 * SU(3) from the
    # continued streams (benchStagProp.nim:84-92), then the gaussian source (stagSolve.nim:542)
    def make_operator(nlinks, reunit):
        r = qb.newRNGField(987654321)
        g = qb.DeviceGauge(nlinks=4)
        g.random(r)
        dev = None
        if nlinks == 4:
            if reunit:
                dev = g.reunit()
            g.rephase()
            s = qb.newStag(g)
            keep = [g]
        else:
            g3 = qb.DeviceGauge(nlinks=4)
            g3.random(r)
            if reunit:
                dev = g3.reunit()
            g3.scale(0.1)
            g.rephase()
            s = qb.newStag3(g, g3)
            g.free(); g3.free()
            keep = []
        b = qb.DeviceField()
        b.gaussian(r)
        r.free()
        return s, b, keep, dev

    results = {}
    big = cfg in ("c3", "c5")
    link_sets = [8] if big else ([4] if args.naive_only else [4, 8])
    for nlinks in link_sets:
        # (key, storage request, reunit first?)
        if nlinks == 4:
            variants = [("links18", 18, False)] + ([] if args.recon == 18 else [("reunit_links12", 12, True)])
        else:
            variants = [("links18_18", 18, False)] + ([] if (args.recon == 18 or big) else [("reunit_links18_12", 1812, True)])
        for vi, (vname, mode, reunit) in enumerate(variants):
            s, fb, keep, unit_dev = make_operator(nlinks, reunit)
            got_rc, dev = s.recon(mode)
            nclass = nlinks // 4
            bytes_site = sum(8 * {12: 96, 14: 112, 18: 144}[got_rc[j]] for j in range(nclass)) + 96
            fx = qb.DeviceField()
            # source in pinned host memory (the e2e leg copies from it every step)
            hb = torch.zeros((nsites, 3, 2), dtype=torch.float64).pin_memory()
            hx = torch.zeros_like(hb).pin_memory()
            b_np, x_np = hb.numpy(), hx.numpy()
            fb.download(b_np)

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
            gflops = 1e-9 * (nlinks * 4 * 72 + 60) * neven_global * iters * args.steps / (1e-3 * ms)
            # true residual |b - A x|^2 / |b|^2 of the TIMED solve with an independent application of the operator:
            # two public stagD2 calls (a, b != 0), not the solver's fused apply
            fx.axpby(0.0, fx, 0.0, qb.ODD)                      # x.odd := 0
            s.dev_stagD2(1, fx, fx, 0.0, 0.0)                   # x.odd = (2D_oe) x.even
            res = qb.DeviceField()
            res.assign(fb)
            s.dev_stagD2(0, res, fx, 1.0, -4.0 * MASS * MASS)   # res.even = b - 4 m^2 x + (2D_eo)(2D_oe) x
            true_r2 = res.norm2(qb.EVEN) / fb.norm2(qb.EVEN)
            res.free()

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
            nrep = args.sweeps if not big else max(20, args.sweeps // 4)
            for _ in range(5):
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
            rr = {
                "gflops": gflops, "ms_per_step": ms / args.steps, "iterations": iters, "r2": sp.r2, "true_r2": true_r2,
                "launches": launches, "clocks": cs.summary(), "recon": got_rc[:nclass], "recon_dev": dev[:nclass],
                "e2e_gflops": e2e_gflops, "e2e_ms_per_step": 1e3 * e2e_s / args.steps,
                "bytes_site": bytes_site, "sweep_ms": sw_ms, "sweep_gbs": achieved, "sweep_frac": achieved / peak,
                "sweep_gflops": 1e-9 * neven * FLOPS_PER_SITE[nlinks] / (1e-3 * sw_ms),
                "cg_iter_ms": ms / args.steps / max(1, iters),
                "cg_iter_gbs": 1e-9 * neven * cg_bytes / (1e-3 * ms / args.steps / max(1, iters)),
                "reunit_checkSU": unit_dev,
            }
            # ---- cold e2e (first variant only): what a caller without a cached handle pays, like qudaSolveXX
            # (qudaWrapperImpl.nim:198-261): create + upload + re-layout of the links, then the host-buffer solve
            if vi == 0 and not args.no_cold and not big and world == 1:
                host_links = [np.zeros((nsites, 3, 3, 2)) for _ in range(nlinks)]
                s.download(host_links)
                barrier()
                t0 = time.perf_counter()
                for _ in range(2):
                    sc = qb.Staggered(host_links)
                    spc = qb.SolverParams(r2req=R2REQ, maxits=20000)
                    sc.solveXX(x_np, b_np, MASS, spc, parEven=True)
                    sc.free()
                barrier()
                cold_s = maxreduce(time.perf_counter() - t0) / 2
                rr["e2e_cold_gflops"] = 1e-9 * (nlinks * 4 * 72 + 60) * neven_global * spc.iterations / cold_s
                rr["e2e_cold_ms"] = 1e3 * cold_s
                rr["e2e_cold_h2d_bytes"] = int(nsites * (48 + nlinks * 144))
                del host_links
            results[(nlinks, vname)] = rr
            fb.free(); fx.free(); s.free()
            for k in keep:
                k.free()
            del hb, hx

    def kernel_name(nlinks, rc):
        return "k_sweep%s<NL=%d, links %s real>" % ("_mr" if world > 1 and args.transport == "p2p" else "", nlinks // 4,
                                                     "/".join(str(t) for t in rc))

    def roof(rr, nlinks):
        tr = ncu_traffic(nlinks, rr["recon"], lat_shape) if world == 1 else None
        return {"bound": "hbm", "kernel": kernel_name(nlinks, rr["recon"]) + " (one parity sweep of the Dslash)",
                "achieved": rr["sweep_gbs"], "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": rr["sweep_frac"], "traffic": tr, "bytes_per_site": rr["bytes_site"],
                "algorithmic_bytes_per_launch": rr["bytes_site"] * neven, "ms_per_launch": rr["sweep_ms"],
                "gflops": rr["sweep_gflops"], "cg_iteration_gbs": rr["cg_iter_gbs"],
                "cg_iteration_frac": rr["cg_iter_gbs"] / peak}

    head_key = (8, "links18_18") if big else (4, "links18")
    r = results[head_key]
    hl = head_key[0]
    line = {
        "metric": METRIC,
        "value": r["gflops"], "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_dict(wl, lat_shape, rankgeom, hl),
        "solve": {"cg_iterations": r["iterations"], "time_to_solution_ms": r["ms_per_step"], "r2": r["r2"]},
        "setup": {"transport": (args.transport if world > 1 else "none"), "overlap": args.overlap,
                  "link_storage": "%s-real (what the library selects for these links)" % "/".join(str(t) for t in r["recon"]),
                  "cg_vectors_mb": 1e-6 * neven * 48},
        "e2e": {"value": r["e2e_gflops"], "unit": "GFLOP/s", "h2d_bytes_per_step": int(nsites * 48),
                "d2h_bytes_per_step": int(nsites * 48), "ms_per_step": r["e2e_ms_per_step"],
                "note": "gauge handle resident (cached against the links, INTEGRATION.md section 2); e2e_cold pays the upload"},
        "gpu_launches": r["launches"],
        "clocks": r["clocks"],
        "roofline": roof(r, hl),
        "true_residual": r["true_r2"],
        "parity": parity,
    }
    if "e2e_cold_gflops" in r:
        line["e2e_cold"] = {"value": r["e2e_cold_gflops"], "unit": "GFLOP/s", "ms_per_step": r["e2e_cold_ms"],
                            "h2d_bytes_per_step": r["e2e_cold_h2d_bytes"], "d2h_bytes_per_step": int(nsites * 48),
                            "what": "gauge_create + gauge_load (H2D + re-layout) + solveXX with host buffers + gauge_destroy, per step"}
    if (8, "links18_18") in results and not big:
        # the north_star target config: 32^4 fat+Naik ("HISQ") fp64, all 18-real, >= 0.70 of the HBM roofline
        line["roofline_target"] = dict(roof(results[(8, "links18_18")], 8), target_frac=0.70,
                                       workload="parity sweep of the fat+Naik operator, 18/18-real links, same lattice")
    variants = {}
    for (nlinks, vname), rr in results.items():
        if (nlinks, vname) == head_key:
            continue
        name = ("naive_" if nlinks == 4 else "hisq_fat_naik_") + vname
        variants[name] = {
            "workload": "same lattice, " + ("naive one-hop links" if nlinks == 4 else "fat + 0.1*Naik links (benchStagProp.nim:84-92)")
                        + (", after reunit (projectSU again, staghmc_sh.nim:247-257: links unitary to 1e-15, 12-real storage accepted)"
                           if "reunit" in vname else ""),
            "link_storage": "/".join(str(t) for t in rr["recon"]), "reunit_checkSU_before_after": rr["reunit_checkSU"],
            "cg_gflops": rr["gflops"], "time_to_solution_ms": rr["ms_per_step"], "cg_iterations": rr["iterations"],
            "true_residual": rr["true_r2"], "e2e_gflops": rr["e2e_gflops"], "roofline": roof(rr, nlinks)}
    line["variants"] = variants
    # the driver keeps `roofline` whole: the parity summary, the north_star target sweep and the
    # fat+Naik solve of this lattice ride inside it
    line["roofline"]["parity"] = parity_summary(parity)
    line["roofline"]["true_residual"] = r["true_r2"]
    if "roofline_target" in line:
        t = line["roofline_target"]
        line["roofline"]["target"] = {"kernel": t["kernel"], "ms_per_launch": t["ms_per_launch"], "frac": t["frac"],
                                      "achieved": t["achieved"], "traffic": t["traffic"], "target_frac": 0.70}
    if (8, "links18_18") in results and not big:
        hr = results[(8, "links18_18")]
        line["roofline"]["hisq_weak"] = {"solve_ms": hr["ms_per_step"], "iterations": hr["iterations"],
                                         "true_residual": hr["true_r2"], "cg_iteration_frac": hr["cg_iter_gbs"] / peak}
    if cfg == "c2" and not args.no_links and (world == 1 or args.transport == "p2p"):
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
    qb.finalize()
    if world == 1 and not args.no_cpu:
        if big:    # bounded sample: the same operator on a 24^3x48 lattice (GFLOP/s per iteration is what is compared)
            line["cpu_baseline"] = cpu_reference_leg((24, 24, 24, 48), 8, 1, qex_links=False)
            line["cpu_baseline"]["sample"] += "; SAMPLE lattice 24x24x24x48 instead of the full %s" % "x".join(map(str, lat_shape))
        else:
            line["cpu_baseline"] = cpu_reference_leg(lat_shape, hl, args.ref_solves)
    else:
        line["cpu_baseline"] = None
    if world > 1:
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
    # a line whose preflight parity or true residual failed is printed (for the record) and the run FAILS: rc 0 at
    # every N then means Dslash <= 1e-13 and CG iterations within +-1 of the oracle on that rank grid
    bad = verdict_of_checks(parity, [rr["true_r2"] for rr in results.values()], args.no_parity)
    if bad:
        sys.stderr.write("bench.py: FAILED checks: %s\n" % "; ".join(bad))
        sys.exit(3)


TRUE_RESIDUAL_SLACK = 1.01   # the CG stops on the RECURSIVE r2 <= r2req b2; the true one agrees to rounding


def parity_summary(parity):
    if parity is None:
        return None
    return {"ok": parity["ok"], "dslash_rel_l2": parity["dslash_rel_l2"], "cg_iters": parity["cg_iters"],
            "cg_iters_fat_naik": parity["fat_naik"]["cg_iters"], "lattice": parity["lattice"], "rankgeom": parity["rankgeom"]}


def verdict_of_checks(parity, true_residuals, no_parity=False):
    """Reasons this run must exit non-zero (empty list = pass).  Pure host logic (tests/test_bench_host.py)."""
    bad = []
    if parity is None:
        if not no_parity:
            bad.append("preflight parity did not run")
    elif not parity.get("ok", False):
        bad.append("preflight parity against the oracle failed: dslash rel-L2 %.3e, cg iters %s"
                   % (parity.get("dslash_rel_l2", float("nan")), parity.get("cg_iters")))
    for t in true_residuals:
        if not (t <= TRUE_RESIDUAL_SLACK * R2REQ):
            bad.append("true residual %.3e of a timed solve exceeds r2req = %.1e" % (t, R2REQ))
    return bad


def trajectory_leg(args, qb, wl, lat_shape, rankgeom, world, rank, local_rank, nsites, neven, peak, peak_src, barrier, maxreduce):
    """BASELINE configs[3]: the solve sequence of ONE staghmc_spv trajectory on nHYP-smeared links, replayed from the
    call trace of the driver (staghmc_spv.nim): h0 = generate_pseudoferms (:521-554: Nf x Ddag, num_pv x solve(m_pv,
    a_tol)) + calc_action (:640-668: Nf x solve(-m, a_tol), num_pv x D); evolve = f_steps x fforce (:780-830: Nf x
    solve(m, f_tol), num_pv x stagD2 on the odd subset); h1 = calc_action.  Links: nHYP smear (alpha 0.4, 0.5, 0.5) of
    QEX-recipe random links ON THE DEVICE, then setBC + stagPhase; they stay fixed over the trace (no gauge update:
    the MD integrator and the gauge force are out of scope), so iteration counts are those of one configuration.
    A step = the whole sequence; value = QEX solver GFLOP/s over its CG iterations."""
    Nf, NPV, FSTEPS, M, MPV, ATOL, FTOL = 2, 2, 10, 0.1, 0.75, 1e-20, 1e-12
    rng = qb.newRNGField(987654321)
    g = qb.DeviceGauge(nlinks=4)
    g.random(rng)
    fl = qb.DeviceGauge(nlinks=4)
    fl.recon(14)                      # nHYP links are U(3): rows 0,1 + one complex factor per link (sticky request)
    qb.timer_start()
    qb.HypCoefs(0.4, 0.5, 0.5).smear(g, fl)
    fl.rephase()
    smear_ms = maxreduce(qb.timer_stop())
    s = qb.newStag(fl)
    storage = fl.recon(14)[0][0]
    bytes_site = 8 * {12: 96, 14: 112, 18: 144}[storage] + 96
    phi = [qb.DeviceField() for _ in range(Nf + NPV)]
    psi, tmp = qb.DeviceField(), qb.DeviceField()
    for f in phi:
        f.gaussian(rng)

    def sequence(record):
        its, napp = 0, 0

        def solve(x, b, m, tol, tag):
            nonlocal its
            sp = qb.SolverParams(r2req=tol, maxits=100000)
            qb.timer_start()
            s.dev_solve(x, b, m, sp)
            ms = qb.timer_stop()
            its += sp.iterations
            if record is not None:
                record.append({"call": tag, "mass": m, "r2req": tol, "iterations": sp.iterations, "ms": ms, "r2": sp.r2})
        for k in range(Nf + NPV):                         # generate_pseudoferms
            psi.gaussian(rng)
            if k < Nf:
                s.dev_Ddag(phi[k], psi, M); napp += 1
            else:
                solve(phi[k], psi, MPV, ATOL, "h0 pseudofermion PV solve")
            phi[k].axpby(0.0, phi[k], 0.0, qb.ODD)        # phi.odd := 0
        for tag in ("h0 action solve",):
            for k in range(Nf + NPV):
                if k < Nf:
                    solve(psi, phi[k], -M, ATOL, tag)
                else:
                    s.dev_D(psi, phi[k], MPV); napp += 1
        for step in range(FSTEPS):                        # fforce per force step
            for k in range(Nf + NPV):
                if k < Nf:
                    solve(psi, phi[k], M, FTOL, "force solve, step %d" % step)
                else:
                    s.dev_stagD2(1, tmp, phi[k], 0.0, 0.0); napp += 0.5
        for k in range(Nf + NPV):                         # calc_action h1
            if k < Nf:
                solve(psi, phi[k], -M, ATOL, "h1 action solve")
            else:
                s.dev_D(psi, phi[k], MPV); napp += 1
        return its, napp

    sequence(None)                                         # warm-up (one full trace)
    barrier()
    qb.launch_count(reset=True)
    trace = []
    with ClockSampler(local_rank) as cs:
        t0 = time.perf_counter()
        for i in range(args.steps):
            its, napp = sequence(trace if i == 0 else None)
        barrier()
        wall = maxreduce(time.perf_counter() - t0)
    launches = qb.launch_count()
    neven_global = neven * world
    gflops = 1e-9 * (4 * 4 * 72 + 60) * neven_global * its * args.steps / wall
    # the sweep of this link set, for the roofline
    for _ in range(6):
        s.dev_stagD2(0, tmp, psi, 0.0, 0.0); s.dev_stagD2(1, tmp, psi, 0.0, 0.0)
    barrier()
    qb.timer_start()
    for _ in range(50):
        s.dev_stagD2(0, tmp, psi, 0.0, 0.0); s.dev_stagD2(1, tmp, psi, 0.0, 0.0)
    sw_ms = maxreduce(qb.timer_stop()) / 100
    achieved = 1e-9 * neven * bytes_site / (1e-3 * sw_ms)
    solve_ms = sum(t["ms"] for t in trace)
    per = {}
    for t in trace:
        k = t["call"].split(",")[0]
        d = per.setdefault(k, {"calls": 0, "iterations": 0, "ms": 0.0})
        d["calls"] += 1; d["iterations"] += t["iterations"]; d["ms"] += t["ms"]
    cg_bytes = 2 * bytes_site + 48 + 432
    return {
        "metric": METRIC, "value": gflops, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps, "warmup": 1,
        "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_dict(wl, lat_shape, rankgeom, 4),
        "solve": {"cg_iterations": its, "operator_applications_outside_cg": napp, "nhyp_smear_rephase_ms": smear_ms,
                  "solve_ms_of_the_trace": solve_ms},
        "setup": {"transport": (args.transport if world > 1 else "none"), "link_storage": "%d-real" % storage},
        "trace_summary": per, "trace": trace,
        "e2e": {"value": gflops, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "an HMC trajectory keeps every field on the device; there is no per-step host traffic to time"},
        "gpu_launches": launches, "clocks": cs.summary(),
        "roofline": {"bound": "hbm", "kernel": "k_sweep%s<NL=1, links %d real>" % ("_mr" if world > 1 else "", storage),
                     "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "bytes_per_site": bytes_site, "ms_per_launch": sw_ms,
                     "cg_iteration_gbs": 1e-9 * neven * cg_bytes * its / (1e-3 * solve_ms),
                     "cg_iteration_frac": 1e-9 * neven * cg_bytes * its / (1e-3 * solve_ms) / peak},
        "cpu_baseline": None,
    }


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


def ncu_traffic(nlinks, recon, lat_shape=(32, 32, 32, 32)):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the parity sweep on this lattice, from the committed
    ncu captures (profiles/ncu_traffic.json); None if that variant was not captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    key = "nl%d_%s" % (nlinks // 4, "_".join(str(t) for t in recon))
    d = json.load(open(p))
    if tuple(lat_shape) != (32, 32, 32, 32):
        key = "x".join(map(str, lat_shape)) + "_" + key
    return d.get(key)


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
    ap.add_argument("--config", default="c2", choices=["c2", "c3", "c4", "c5"],
                    help="SURVEY 8d numbering: c2 = 32^4 naive per GPU (BASELINE configs[1], default, weak scaling); c3 = 48^3x96 "
                         "HISQ strong scaling; c4 = 64^4 nHYP trajectory solve sequence; c5 = 64^3x128 HISQ strong scaling")
    ap.add_argument("--sweeps", type=int, default=200, help="back-to-back parity sweeps for the roofline leg")
    ap.add_argument("--ref-solves", type=int, default=8, help="complete CPU solves in the bounded cpu_baseline sample")
    ap.add_argument("--ref-maxits", type=int, default=20, help="iteration cap per CPU solve on the large lattices")
    ap.add_argument("--naive-only", action="store_true")
    ap.add_argument("--recon", type=int, default=12, choices=[12, 18],
                    help="12: also run the variants on re-unitarised links stored as 12 reals; 18: 18-real runs only")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU halo/sum transport: peer-memory NVLink stores (default) or NCCL send/recv + allreduce")
    ap.add_argument("--overlap", default="on", choices=["on", "off"],
                    help="off: the halo is waited for BEFORE any site is computed (face tiles first / exchange before the interior kernel)")
    ap.add_argument("--alt8", action="store_true", help="8 GPUs as 1.2.2.2 (global 32x64^3) instead of 1.1.2.4")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-links", action="store_true", help="skip the link-pipeline side leg")
    ap.add_argument("--no-cold", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the preflight parity check against the oracle")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.config in ("c3", "c5", "c4") and args.impl == "ours":
        args.steps = min(args.steps, 2)
        args.warmup = 3 if args.config != "c4" else 1
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
