"""Summarise ncu outputs into profiles/: launch-list shares (csv from --metrics gpu__time_duration.sum)
and the --set full capture (raw page csv).  Usage:
  python tools/ncu_summary.py launches.csv raw.csv > profiles/rNN_summary.md   (also writes profiles/ncu_traffic.json)"""
import csv
import json
import os
import re
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def launch_table(path):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    hdr = rows[0]
    kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= mv:
            continue
        t = float(r[mv].replace(",", ""))
        a = agg.setdefault(r[kn], [0, 0.0])
        a[0] += 1
        a[1] += t
    unit = rows[1][hdr.index("Metric Unit")]
    scale = 1e-3 if unit in ("ns", "nsecond") else 1.0
    tot = sum(v[1] for v in agg.values())
    out = ["| kernel | launches | total us | avg us | share |", "|---|---:|---:|---:|---:|"]
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("| `%s` | %d | %.1f | %.2f | %.1f%% |" % (k[:80], n, t * scale, t * scale / n, 100 * t / tot))
    return "\n".join(out)


def full_table(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"]
    kn = hdr.index("Kernel Name")
    per = OrderedDict()
    for r in rows[2:]:
        per.setdefault(r[kn], []).append(r)
    to_bytes = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    out, traffic = [], {}
    for k, rs in per.items():
        out.append("\n### `%s` (%d launches captured)\n" % (k, len(rs)))
        out.append("| metric | " + " | ".join("launch %d" % (i + 1) for i in range(len(rs))) + " | unit |")
        out.append("|---|" + "---:|" * len(rs) + "---|")
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                out.append("| %s | %s | %s |" % (w, " | ".join(r[i] for r in rs), units[i]))
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        tr = sum(float(r[ir]) * to_bytes[units[ir]] + float(r[iw]) * to_bytes[units[iw]] for r in rs) / len(rs)
        m = re.search(r"k_sweep<(\d), (\d), ", k)
        if m:
            nl, rc = int(m.group(1)), int(m.group(2))
            rec = [12 if (rc >> j) & 1 else 18 for j in range(nl)]
            traffic["nl%d_%s" % (nl, "_".join(map(str, rec)))] = int(tr)
            alg = 524288 * (sum(8 * (96 if t == 12 else 144) for t in rec) + 96)
            out.append("\nDRAM traffic per launch %.1f MB vs algorithmic %.1f MB (ratio %.3f)." % (tr / 1e6, alg / 1e6, tr / alg))
    json.dump(traffic, open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w"), indent=1)
    return "\n".join(out)


if __name__ == "__main__":
    print("## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`, cold-cache, serialised: compare shares)\n")
    print(launch_table(sys.argv[1]))
    print("\n## `ncu --set full --clock-control none --import-source on` of the parity-sweep kernels (32^4, 524 288 output sites)")
    print(full_table(sys.argv[2]))
