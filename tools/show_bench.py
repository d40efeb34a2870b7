"""Pretty-print bench.py JSON lines (files given on the command line)."""
import json
import sys

for f in sys.argv[1:]:
    txt = open(f).read()
    lines = [l for l in txt.splitlines() if l.startswith("{")]
    if not lines:
        print(f, "no JSON line")
        continue
    d = json.loads(lines[-1])
    r = d["roofline"]
    print("%s: N=%d %s links %s | %.0f GF, %.2f ms/solve, e2e %.0f GF | sweep %.4f ms %.0f GB/s frac %.3f, CG-iteration frac %.3f | launches %d | clocks %s"
          % (f, d["n_gpus"], d["config"].get("transport"), d["config"].get("link_storage"), d["value"], d["ms_per_step"],
             d["e2e"]["value"], r["ms_per_launch"], r["achieved"], r["frac"], r.get("cg_iteration_frac", 0), d["gpu_launches"], d["clocks"]))
    for k, v in (d.get("variants") or {}).items():
        r = v["roofline"]
        print("    %-28s %.0f GF, %.2f ms/solve | sweep %.4f ms %.0f GB/s frac %.3f, CG-iteration frac %.3f"
              % (k, v["cg_gflops"], v["time_to_solution_ms"], r["ms_per_launch"], r["achieved"], r["frac"], r.get("cg_iteration_frac", 0)))
    if d.get("cpu_baseline"):
        c = d["cpu_baseline"]
        print("    cpu_baseline %.1f GF on %d cores (%s)" % (c["value"], c["cores"], c["sample"]))
