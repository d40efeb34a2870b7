"""One rank of a multi-rank parity run (launched by torch.distributed.run; see test_gpu_multi.py and
test_dist_cpu.py).  Every rank builds the same global inputs with the oracle, keeps its own block in
native QEX layout, runs the product and compares ITS block with the oracle's global result."""
import argparse
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def rank_coords(rank, G):
    rc = [0] * 4
    for d in (3, 2, 1, 0):
        rc[d] = rank % G[d]
        rank //= G[d]
    return rc


def cpu_plan_check(L, G, nlinks, rank, world, dist):
    """Host logic of the N>1 path on CPU (gloo): the halo plans of all ranks pair up, and an exchange
    driven by the plan moves exactly the planned byte counts between the planned peers."""
    import torch
    import qex_b200 as qb
    plan, total = qb.plan_halo(L, G, rank, nlinks)
    allp = [None] * world
    dist.all_gather_object(allp, (plan, total))
    rc = rank_coords(rank, G)
    for h in plan:
        mu = h["mu"]
        up = list(rc); up[mu] = (rc[mu] + 1) % G[mu]
        dn = list(rc); dn[mu] = (rc[mu] - 1) % G[mu]
        lex = lambda c: ((c[0] * G[1] + c[1]) * G[2] + c[2]) * G[3] + c[3]   # dimension 0 slowest
        assert h["rank_plus"] == lex(up) and h["rank_minus"] == lex(dn), (h, rc)
        peer = [q for q in allp[h["rank_plus"]][0] if q["mu"] == mu][0]
        assert peer["rank_minus"] == rank and peer["n_f"] == h["n_f"] and peer["n_b"] == h["n_b"]
        X = [L[d] // G[d] for d in range(4)]
        assert h["faceh"] == int(np.prod(X)) // X[mu] // 2
        assert h["depth"] == (3 if nlinks == 8 else 1)
        assert h["n_f"] == 3 * h["depth"] * h["faceh"] and h["n_b"] == 3 * (4 if nlinks == 8 else 1) * h["faceh"]
    # block -> tile table of the one-kernel multi-rank sweep (csrc/dslash.cu plan_tiles): every tile once, interior tiles
    # first, and a face tile waits exactly for the neighbours its sites hop to
    X = [L[d] // G[d] for d in range(4)]
    Vh = int(np.prod(X)) // 2
    dep = 3 if nlinks == 8 else 1
    for parity in (0, 1):
        tiles = qb.plan_tiles(L, G, rank, nlinks, parity)
        ntiles = (Vh + 127) // 128
        assert sorted(t for t, _ in tiles) == list(range(ntiles))
        needs = [n for _, n in tiles]
        first_face = next((i for i, n in enumerate(needs) if n), len(needs))
        assert all(n == 0 for n in needs[:first_face]) and all(n != 0 for n in needs[first_face:])
        idx = np.arange(Vh)
        xh = X[0] // 2
        x1 = (idx // xh) % X[1]; x2 = (idx // (xh * X[1])) % X[2]; x3 = idx // (xh * X[1] * X[2])
        x0 = 2 * (idx % xh) + ((parity + x1 + x2 + x3) & 1)
        xs = [x0, x1, x2, x3]
        want = np.zeros(ntiles, dtype=np.int64)
        for mu in range(4):
            if G[mu] == 1:
                continue
            np.bitwise_or.at(want, idx // 128, np.where(xs[mu] + dep >= X[mu], 1 << (2 * mu), 0))
            np.bitwise_or.at(want, idx // 128, np.where(xs[mu] - dep < 0, 1 << (2 * mu + 1), 0))
        for t, n in tiles:
            assert n == want[t], (parity, t, n, want[t])
        split = {h["mu"] for h in plan}
        assert all(((n >> (2 * mu)) & 3) == 0 for n in needs for mu in range(4) if mu not in split)
    # emulate one sweep's exchange with gloo: same order as csrc/comm.cu halo_start
    send = torch.arange(2 * total, dtype=torch.float64) + 1e6 * rank
    recv = torch.zeros(2 * total, dtype=torch.float64)
    ops = []
    for h in plan:
        f0, f1, b0, b1 = 2 * h["off_f"], 2 * (h["off_f"] + h["n_f"]), 2 * h["off_b"], 2 * (h["off_b"] + h["n_b"])
        ops += [dist.P2POp(dist.isend, send[f0:f1], h["rank_minus"]), dist.P2POp(dist.isend, send[b0:b1], h["rank_plus"]),
                dist.P2POp(dist.irecv, recv[f0:f1], h["rank_plus"]), dist.P2POp(dist.irecv, recv[b0:b1], h["rank_minus"])]
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for h in plan:
        f0, b0 = 2 * h["off_f"], 2 * h["off_b"]
        assert recv[f0] == f0 + 1e6 * h["rank_plus"], "psi faces must come from the + neighbour"
        assert recv[b0] == b0 + 1e6 * h["rank_minus"], "U^+ psi faces must come from the - neighbour"
    return True


def cpu_sweep_check(L, G, nlinks, rank, world, dist):
    """One PROCESS per rank on the CPU (gloo): the site functions the CUDA kernels wrap (csrc/sweep_site.cuh, compiled
    with g++ into tests/emu/_emu_sweep.so) pack this rank's faces, the send buffers travel through torch.distributed
    with the library's own halo plan (same pairing and order as csrc/comm.cu halo_start), then the interior sweep and the
    boundary terms run on the received data -- and every rank compares its block with the oracle's global stagD2."""
    import ctypes as C
    import torch
    import oracle
    import qex_b200 as qb
    from util import make_inputs, rel_l2
    from test_sweep_emu import EmuHalo, SO, block_index, links_to_dev, vec_to_dev, vec_from_dev
    lib = C.CDLL(SO)     # built by the launching test (tests/test_dist_cpu.py) before the ranks start
    dp_, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    lib.emu_rank_pack.argtypes = [C.c_int, ip, C.c_int, dp_, dp_, C.POINTER(EmuHalo), C.c_int, dp_]
    lib.emu_rank_sweep.argtypes = [C.c_int, ip, C.c_int, C.c_int, dp_, dp_, dp_, C.c_double, C.c_double, C.c_double, C.c_double,
                                   C.POINTER(EmuHalo), C.c_int, dp_]
    lat, g, (x, r0) = make_inputs(L, nlinks)
    X = [L[d] // G[d] for d in range(4)]
    Vh = int(np.prod(X)) // 2
    bk = block_index(L, G, rank)
    plan, total = qb.plan_halo(L, G, rank, nlinks)
    halo = (EmuHalo * 4)()
    for i, h in enumerate(plan):
        halo[i] = EmuHalo(h["mu"], h["depth"], h["faceh"], h["rank_plus"], h["rank_minus"], h["off_f"], h["off_b"])
    comm_mask = sum(1 << d for d in range(4) if G[d] > 1)
    dl, dx = links_to_dev(g[:, bk]), vec_to_dev(x[bk])
    dp = lambda a: a.ctypes.data_as(dp_)
    Xc = (C.c_int * 4)(*X)
    worst = 0.0
    for parity in (0, 1):
        for a, b in ((0.0, 0.0), (0.7, -0.3)):
            send = torch.full((2 * total,), float("nan"), dtype=torch.float64)
            recv = torch.full((2 * total,), float("nan"), dtype=torch.float64)
            assert lib.emu_rank_pack(nlinks // 4, Xc, parity, dp(dl), dp(dx), halo, len(plan), dp(send.numpy())) == 0
            ops = []
            for h in plan:      # csrc/comm.cu halo_start: psi faces to the - neighbour, products to the + neighbour
                f0, f1, b0, b1 = 2 * h["off_f"], 2 * (h["off_f"] + h["n_f"]), 2 * h["off_b"], 2 * (h["off_b"] + h["n_b"])
                ops += [dist.P2POp(dist.isend, send[f0:f1], h["rank_minus"]), dist.P2POp(dist.isend, send[b0:b1], h["rank_plus"]),
                        dist.P2POp(dist.irecv, recv[f0:f1], h["rank_plus"]), dist.P2POp(dist.irecv, recv[b0:b1], h["rank_minus"])]
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            dr = vec_to_dev(r0[bk])
            assert lib.emu_rank_sweep(nlinks // 4, Xc, comm_mask, parity, dp(dl), dp(dx), dp(dr), a, b, 1.0, 1.0, halo, len(plan),
                                      dp(recv.numpy())) == 0
            want = r0.copy()
            lat.stagD2(g, want, x, a, b, parity)
            got, w = vec_from_dev(dr), want[bk]
            sl = slice(0, Vh) if parity == 0 else slice(Vh, 2 * Vh)
            assert np.isfinite(got).all()
            e = rel_l2(got[sl], w[sl])
            assert e <= 1e-13, (rank, parity, a, b, e)
            worst = max(worst, e)
    return worst


def gpu_case(L, G, nlinks, inner, rank, world, local_rank, dist, transport="nccl", recon=18):
    import torch
    import oracle
    import qex_b200 as qb
    from util import make_inputs, make_exact_inputs, rel_l2
    torch.cuda.set_device(local_rank)
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if transport == "nccl":
        if rank == 0:
            idt = torch.tensor(list(qb.nccl_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(idt, 0)
    qb.init(L, G, rank, local_rank, bytes(idt.cpu().tolist()) if transport == "nccl" else None)
    if transport == "p2p":
        qb.p2p_attach_torch(dist)
        assert qb.p2p_active()
    qb.set_host_layout(inner)
    lat, g, (x, b) = (make_exact_inputs if recon in (12, 14) else make_inputs)(L, nlinks)
    if recon == 14:
        # U(3) links whose determinant phase changes from site to site (what projectU leaves: nHYP links, the Naik
        # products of the HISQ W links): exact SU(3) times a per-link phase
        th = np.random.default_rng(99).uniform(0, 2 * np.pi, g.shape[:2])
        gc = (g[..., 0] + 1j * g[..., 1]) * np.exp(1j * th)[:, :, None, None]
        g = np.ascontiguousarray(np.stack([gc.real, gc.imag], axis=-1))
    lo = oracle.QexLayout(L, G, inner)
    mine = lambda f: lo.scatter(f)[rank]
    s = qb.Staggered([mine(g[k]) for k in range(nlinks)])
    res = {}
    if recon in (12, 14):
        got_rc, dev = s.recon(recon)
        assert got_rc == ([recon, recon] if nlinks == 8 else [recon, 18]), (got_rc, dev)
    # Dslash, all flavours
    for parity in (0, 1):
        want = b.copy()
        lat.stagD2(g, want, x, 0.7, -0.3, parity)
        got = mine(b)
        qb.stagD2(s, parity, got, mine(x), 0.7, -0.3)
        res["stagD2_%d" % parity] = rel_l2(got, mine(want))
    for name in ("D", "Ddag"):
        want = lat.vector()
        getattr(lat, name)(g, want, x, 0.1)
        got = np.zeros_like(mine(x))
        getattr(s, name)(got, mine(x), 0.1)
        res[name] = rel_l2(got, mine(want))
    want = lat.vector()
    lat.stagD2xx(g, want, x, 0.01, 0)
    got = np.zeros_like(mine(x))
    qb.stagD2ee(s, got, mine(x), 0.01)
    we = want.copy(); we[lat.half:] = 0
    res["stagD2ee"] = rel_l2(got, mine(we))
    for k, v in res.items():
        assert v <= 1e-13, (k, v)
    # reductions over ranks
    fx = qb.DeviceField(mine(x))
    n2 = fx.norm2()
    assert abs(n2 - lat.norm2(x)) <= 1e-12 * n2
    fx.free()
    # device RNG field: seeded with the GLOBAL lexicographic site index, so every rank draws its block of the same
    # global fields (distributionUtils.nim:319-324)
    if inner == (1, 1, 1, 1):
        ro = lat.new_rng(12345)
        go, vo = lat.random_gauge(ro, 4), lat.gaussian_vector(ro)
        rd = qb.newRNGField(12345)
        dg, dv = qb.DeviceGauge(nlinks=4), qb.DeviceField()
        dg.random(rd)
        dv.gaussian(rd)
        gd = [np.zeros_like(mine(go[0])) for _ in range(4)]
        dg.download(gd)
        vd = dv.download(np.zeros_like(mine(vo)))
        assert np.count_nonzero(vd != mine(vo)) <= 1, "device gaussian field differs from the oracle's block"
        res["rng_links"] = max(rel_l2(gd[k], mine(go[k])) for k in range(4))
        assert res["rng_links"] <= 1e-9, res
        # reunit on the split lattice: deviations are sums / maxima over ALL ranks (four cross-rank sums in one kernel)
        before, after = dg.reunit()
        assert 1e-13 < before[1] < 1e-6 and after[1] < 5e-15 and after[0] <= after[1], (before, after)
        res["reunit"] = (before, after)
        for d in (dg, dv, rd):
            d.free()
    # CG and the full solve
    m, r2req = 0.1, 1e-10
    want = lat.vector()
    it_o, _ = lat.solveXX(g, want, b, m, 0, r2req, 10000)
    got = np.zeros_like(mine(b))
    sp = qb.SolverParams(r2req=r2req, maxits=10000)
    s.solveEE(got, mine(b), m, sp)
    assert abs(sp.iterations - it_o) <= 1, (sp.iterations, it_o)
    assert rel_l2(got, mine(want)) <= 10 * np.sqrt(r2req)
    res["solveEE_its"] = (sp.iterations, it_o)
    want = lat.vector()
    st = lat.solve(g, want, b, m, r2req, 10000)
    got = np.zeros_like(mine(b))
    sp = qb.SolverParams(r2req=r2req, maxits=10000)
    s.solve(got, mine(b), m, sp)
    assert abs(sp.iterations - st.iterations) <= 1, (sp.iterations, st.iterations)
    assert sp.r2 <= r2req
    assert rel_l2(got, mine(want)) <= 10 * np.sqrt(r2req)
    res["solve_its"] = (sp.iterations, st.iterations)
    s.free()
    qb.finalize()
    return res


def links_case(L, G, inner, rank, world, local_rank, dist):
    """Device link pipeline, force outer products and plaq on a lattice split over ranks (peer arena + NVLink
    reads of the neighbours' fields): every rank compares its block with the oracle's global result."""
    import torch
    import oracle
    import qex_b200 as qb
    from util import make_inputs, rel_l2, SEED
    torch.cuda.set_device(local_rank)
    qb.init(L, G, rank, local_rank, None)
    qb.p2p_attach_torch(dist)
    qb.set_host_layout(inner)
    lat = oracle.Lattice(L)
    g = lat.random_gauge(lat.new_rng(SEED), 4)
    lo = oracle.QexLayout(L, G, inner)
    mine = lambda f: lo.scatter(f)[rank]

    def down(dg):
        parts = [np.zeros_like(mine(g[0])) for _ in range(dg.nlinks)]
        dg.download(parts)
        return parts
    res = {}
    dg = qb.DeviceGauge([mine(g[k]) for k in range(4)])
    res["plaq"] = float(np.abs(dg.plaq() - lat.plaq(g)).max())
    assert res["plaq"] < 1e-14
    # nHYP, then rephase, then a solve on the smeared links (staghmc_spv.nim:367-386)
    want = lat.hyp_smear(g, (0.4, 0.5, 0.5))
    fl = qb.DeviceGauge(nlinks=4)
    qb.HypCoefs(0.4, 0.5, 0.5).smear(dg, fl)
    got = down(fl)
    res["nhyp"] = max(rel_l2(got[mu], mine(want[mu])) for mu in range(4))
    assert res["nhyp"] <= 1e-13, res
    assert np.abs(fl.plaq() - lat.plaq(want)).max() < 1e-13
    if int(os.environ.get("QEXB200_ARENA_FIELDS", "64")) >= 100:
        # smearGetForce / smearedForce on the split lattice (hypsmear.nim:49-262), in place like staghmc_spv.nim:745
        chain = np.random.default_rng(4).standard_normal(g.shape)
        wantf = lat.hyp_force(g, (0.4, 0.5, 0.5), chain)
        fl2 = qb.DeviceGauge(nlinks=4)
        dc = qb.DeviceGauge([mine(chain[k]) for k in range(4)])
        force = qb.HypCoefs(0.4, 0.5, 0.5).smearGetForce(dg, fl2)
        force(dc, dc)
        gotf = down(dc)
        res["nhyp_force"] = max(rel_l2(gotf[mu], mine(wantf[mu])) for mu in range(4))
        assert res["nhyp_force"] <= 1e-11, res
        dc.projTAH(dg)
        force.free(); fl2.free(); dc.free()
    fl.rephase()
    lat.setBC(want); lat.stagPhase(want)
    s = qb.newStag(fl)
    x = lat.gaussian_vector(lat.new_rng(3))
    ref, sol = lat.vector(), np.zeros_like(mine(x))
    st = lat.solve(want, ref, x, 0.1, 1e-10)
    sp = qb.SolverParams(r2req=1e-10)
    s.solve(sol, mine(x), 0.1, sp)
    assert abs(sp.iterations - st.iterations) <= 1
    assert rel_l2(sol, mine(ref)) < 1e-4
    # HISQ on phased thin links (testStagProp.nim:24-36); bar follows the oracle's conditioning
    gp = g.copy(); lat.setBC(gp); lat.stagPhase(gp)
    from util import hisq_reference, hisq_bar
    wfl, wll, efl, ell = hisq_reference(lat, gp, oracle.hisq_coefs())
    dgp = qb.DeviceGauge([mine(gp[k]) for k in range(4)])
    hfl, hll = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
    qb.HisqCoefs().init().smear(dgp, hfl, hll)
    gfl, gll = down(hfl), down(hll)
    res["hisq_fat"] = max(rel_l2(gfl[mu], mine(wfl[mu])) for mu in range(4))
    res["hisq_long"] = max(rel_l2(gll[mu], mine(wll[mu])) for mu in range(4))
    bar_f, bar_l = hisq_bar(wfl, efl), hisq_bar(wll, ell)   # 10x the oracle's own distance from the exact-projection chain
    assert res["hisq_fat"] <= bar_f and res["hisq_long"] <= bar_l, (res, bar_f, bar_l)
    for mu in range(4):   # site by site the typical deviation is at rounding level
        w = mine(wfl[mu])
        site = np.linalg.norm((gfl[mu] - w).reshape(w.shape[0], -1), axis=1) / np.linalg.norm(w.reshape(w.shape[0], -1), axis=1)
        # bar from an independent quantity: the ORACLE's own median per-site distance from the
        # exact-projection chain (numpy SVD polar factor) on this rank's block, 4.6-4.8e-15 at these sizes.  Two correct
        # implementations of QEX's closed-form projectU, each that far from the exact result, differ by up to twice it;
        # another factor 2 covers the block-to-block spread.  Observed device-vs-oracle medians: 0.6-1.01e-14.
        e = mine(efl[mu])
        own = np.linalg.norm((w - e).reshape(w.shape[0], -1), axis=1) / np.linalg.norm(e.reshape(e.shape[0], -1), axis=1)
        bar_site = 4.0 * float(np.median(own))
        res.setdefault("hisq_site_median", []).append([float(np.median(site)), bar_site])
        assert np.median(site) < bar_site, (mu, float(np.median(site)), bar_site)
    if int(os.environ.get("QEXB200_ARENA_FIELDS", "64")) >= 100:
        # fat7lDeriv and the HISQ smearedForce on the split lattice (fat7lderiv.nim:122-335, hisqsmear.nim:51-84), on
        # links near the unit gauge so that the chain is well conditioned and the bar can be tight
        gs = lat.projectU(lat.unit_gauge() + 0.15 * np.random.default_rng(8).standard_normal(g.shape))
        lat.setBC(gs); lat.stagPhase(gs)
        rr = np.random.default_rng(3)
        ch, chl = rr.standard_normal(g.shape), rr.standard_normal(g.shape)
        wantf = lat.hisq_force(gs, oracle.hisq_coefs(), ch, chl)
        dgs = qb.DeviceGauge([mine(gs[k]) for k in range(4)])
        su, sul, du = qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4), qb.DeviceGauge(nlinks=4)
        dch, dchl = qb.DeviceGauge([mine(ch[k]) for k in range(4)]), qb.DeviceGauge([mine(chl[k]) for k in range(4)])
        hforce = qb.HisqCoefs().init().smearGetForce(dgs, su, sul)
        du.zero()
        hforce(du, dch, dchl)
        gotf = down(du)
        res["hisq_force"] = max(rel_l2(gotf[mu], mine(wantf[mu])) for mu in range(4))
        assert res["hisq_force"] <= 1e-12, res
        hforce.free()
        for d in (dgs, su, sul, du, dch, dchl):
            d.free()
    # fat + Lepage only (all five coefficients), tight bar: no projection in between
    c5 = (1.0, -1 / 16., 1 / 64., -1 / 384., -1 / 8.)
    w5, w5l = lat.makeImpLinks(gp, c5, gfLong=gp, naik=-1 / 24.)
    qb.makeImpLinks(hfl, dgp, qb.Fat7lCoefs(*c5), hll, dgp, -1 / 24.)
    gfl, gll = down(hfl), down(hll)
    res["fat7"] = max(rel_l2(gfl[mu], mine(w5[mu])) for mu in range(4))
    res["naik"] = max(rel_l2(gll[mu], mine(w5l[mu])) for mu in range(4))
    assert res["fat7"] <= 1e-13 and res["naik"] <= 1e-13, res
    for d in (s, dg, fl, dgp, hfl, hll):
        d.free()
    # force outer products, one-hop and one+three-hop
    for nl in (4, 8):
        _, gg, (xv, cv) = make_inputs(L, nl)
        sop = qb.Staggered([mine(gg[k]) for k in range(nl)])
        start = 0.25 * gg
        wantf = start.copy()
        lat.stagD2deriv(wantf, cv, xv)
        gr = qb.DeviceGauge([mine(start[k]) for k in range(nl)])
        dc, dx = qb.DeviceField(mine(cv)), qb.DeviceField(mine(xv))
        qb.stagD2deriv(sop, gr, dc, dx)
        gotf = down(gr)
        res["stagD2deriv_%d" % nl] = max(rel_l2(gotf[k], mine(wantf[k])) for k in range(nl))
        assert res["stagD2deriv_%d" % nl] <= 1e-13, res
        if nl == 4:
            wd = lat.gauge(4)
            lat.stagDeriv(wd, xv)
            gr.zero()
            qb.stagDeriv(sop, gr, dx)
            gotf = down(gr)
            res["stagDeriv"] = max(rel_l2(gotf[k], mine(wd[k])) for k in range(4))
            assert res["stagDeriv"] <= 1e-13, res
        for d in (sop, gr, dc, dx):
            d.free()
    qb.finalize()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--cpu-sweep", action="store_true", help="with --cpu: also run the emulated multi-rank sweep over gloo")
    ap.add_argument("--lat", default="8,8,8,16")
    ap.add_argument("--rankgeom", default="1,1,1,2")
    ap.add_argument("--inner", default="1,1,1,1")
    ap.add_argument("--nlinks", type=int, default=4)
    ap.add_argument("--transport", default="nccl", choices=["nccl", "p2p"])
    ap.add_argument("--recon", type=int, default=18, choices=[12, 14, 18])
    ap.add_argument("--links", action="store_true", help="link pipeline / force case instead of the Dslash + CG case")
    a = ap.parse_args()
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    L = tuple(int(t) for t in a.lat.split(","))
    G = tuple(int(t) for t in a.rankgeom.split(","))
    inner = tuple(int(t) for t in a.inner.split(","))
    assert int(np.prod(G)) == world
    if a.cpu:
        dist.init_process_group("gloo")
        ok = cpu_plan_check(L, G, a.nlinks, rank, world, dist)
        out = {"ok": ok}
        if a.cpu_sweep:
            out["sweep_rel_l2"] = cpu_sweep_check(L, G, a.nlinks, rank, world, dist)
    else:
        import torch
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        if a.links:
            out = links_case(L, G, inner, rank, world, local_rank, dist)
        else:
            out = gpu_case(L, G, a.nlinks, inner, rank, world, local_rank, dist, a.transport, a.recon)
    dist.barrier()
    if rank == 0:
        print("MULTI_OK " + json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
