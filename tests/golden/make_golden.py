"""Writes the committed fixtures under tests/golden/.

trandgauge.json : copied literally from the reference's known-answer test
                  /root/reference/tests/reprod/trandgauge.nim:17 (non-FUELCompat branch).
rng_milc6.json  : first draws of RngMilc6 (src/rng/milcrng.nim:92-134) for site index 0, from a
                  pure-Python restatement independent of the C++ oracle.
staghmc_sh.json : the first "Begin H" line (Sf per pseudofermion, T) of the reference's golden HMC logs
                  /root/reference/tests/extra/staghmc_sh/ref.{0,1,2}, with the run parameters of
                  tests/extra/staghmc_sh/run:21-33 and the per-solver iteration statistics of trajectory 1.
                  These are numbers QEX itself produced (2 ranks x 2 threads) and checks at rel 2e-11 (run:43).
config1.json    : BASELINE config 1 solved by the oracle (self-consistency record, not a reference output;
                  the reference cannot be built here -- SURVEY.md 8c).
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))


def trandgauge():
    src = open("/root/reference/tests/reprod/trandgauge.nim").read()
    m = re.search(r"else:\s*\n\s*const P = @\[([^\]]+)\]", src)
    P = [float(t) for t in m.group(1).split(",")]
    json.dump({"source": "tests/reprod/trandgauge.nim:17", "lattice": [8, 8, 8, 8], "seed": 17 ** 7, "P": P},
              open(os.path.join(HERE, "trandgauge.json"), "w"), indent=1)


def rng():
    M32 = 0xFFFFFFFF
    MASK = 0x00FFFFFF

    def seed(seed0, index):
        s = seed0 & M32
        r = []
        for _ in range(7):
            s = ((69607 + 8 * index) * s + 12345) & M32
            r.append((s >> 8) & MASK)
        s = ((69607 + 8 * index) * s + 12345) & M32
        return r, s, (100005 + 8 * index) & M32

    def nxt(st):
        r, ic, mult = st
        t = ((((r[5] >> 7) | (r[6] << 17)) ^ ((r[4] >> 1) | (r[5] << 23))) & MASK)
        r[:] = [t] + r[:6]
        ic = (ic * mult + 12345) & M32
        st[1] = ic
        return t ^ ((ic >> 8) & MASK)

    r, ic, mult = seed(17 ** 7, 0)
    st = [r, ic, mult]
    vals = [nxt(st) for _ in range(16)]
    json.dump({"source": "src/rng/milcrng.nim:92-134", "seed": 17 ** 7, "index": 0, "site0": vals},
              open(os.path.join(HERE, "rng_milc6.json"), "w"), indent=1)


def staghmc_sh():
    base = "/root/reference/tests/extra/staghmc_sh/"
    runs = {"0": dict(mass=[0.1], hmasses=[[0.2, 0.4]]),
            "1": dict(mass=[0.1, 0.05], hmasses=[[0.2, 0.4], [0.2, 0.4]]),
            "2": dict(mass=[0.1, 0.05], hmasses=[[0.2, 0.4], [0.2, 0.4]])}
    out = {"source": "tests/extra/staghmc_sh/ref.{0,1,2} first 'Begin H' line; parameters tests/extra/staghmc_sh/run:21-33",
           "lattice": [8, 8, 8, 8], "seed": 987654321, "arsq": 1e-24, "rel_tol": 2e-11, "runs": {}}
    for tid, par in runs.items():
        lines = open(base + "ref." + tid).read().splitlines()
        k = next(i for i, l in enumerate(lines) if "Begin H:" in l)
        m = re.search(r"Begin H: (\S+)\s+Sg: (\S+)\s+Sf: @\[(.*)\]\s+T: (\S+)", lines[k])
        sf = [[float(t) for t in grp.split(",")] for grp in re.findall(r"@\[([^\]@]+)\]", m.group(3))]
        # Solver[action] statistics of trajectory 1: "A m=0.1 2: 56:99 ..." = calls: (iterations div calls):max
        a0 = next(i for i in range(k, len(lines)) if lines[i].startswith("Solver[action]:"))
        stats = []
        for l in lines[a0 + 1:]:
            mm = re.match(r"\s+A m=(\S+) (\d+): (\d+):(\d+)", l)
            if not mm:
                break
            stats.append({"m": float(mm.group(1)), "calls": int(mm.group(2)), "its_div_calls": int(mm.group(3)),
                          "its_max": int(mm.group(4))})
        out["runs"][tid] = dict(par, line=k + 1, H=float(m.group(1)), Sg=float(m.group(2)), Sf=sf, T=float(m.group(4)),
                                action_stats=stats)
    json.dump(out, open(os.path.join(HERE, "staghmc_sh.json"), "w"), indent=1)


def config1():
    from util import make_inputs
    lat, g, (b, _) = make_inputs((8, 8, 8, 8), 4)
    x = lat.vector()
    st = lat.solve(g, x, b, 0.1, 1e-8)
    json.dump({"config": "8^4 naive staggered, random SU(3), m=0.1, r2req=1e-8, seed 987654321",
               "iterations": st.iterations, "r2": st.r2, "x_norm2": lat.norm2(x), "b_norm2": lat.norm2(b)},
              open(os.path.join(HERE, "config1.json"), "w"), indent=1)


if __name__ == "__main__":
    trandgauge()
    rng()
    staghmc_sh()
    config1()
