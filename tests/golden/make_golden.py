"""Writes the committed fixtures under tests/golden/.

trandgauge.json : copied literally from the reference's known-answer test
                  /root/reference/tests/reprod/trandgauge.nim:17 (non-FUELCompat branch).
rng_milc6.json  : first draws of RngMilc6 (src/rng/milcrng.nim:92-134) for site index 0, from a
                  pure-Python restatement independent of the C++ oracle.
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
    config1()
