#!/bin/bash
# round 3, call 4 (1 GPU): GPU suite on the PDL build, then the bench line with QEXB200_PDL = 0 / 1 / 2 (A/B of the
# programmatic dependent launch in the CG loop and of the link prefetch before the wait)
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r03_gpu_suite_pdl.log 2>&1; echo "suite rc=$?" >> $O/r03_gpu_suite_pdl.log; tail -n 3 $O/r03_gpu_suite_pdl.log
for lv in 0 1 2 1 0 2; do
  QEXB200_PDL=$lv python bench.py --steps 10 --warmup 3 --no-cpu --no-links --no-cold > $O/r03_pdl_$lv.json 2> $O/r03_pdl_$lv.err
  echo "PDL=$lv rc=$?"
  python - $lv <<'PY'
import json,sys
lv=sys.argv[1]
d=json.loads(open("gpurun_out/r03_pdl_%s.json"%lv).read().strip().splitlines()[-1])
v=d["variants"]
print("PDL=%s naive18 %.3f ms (%d its) sweep %.4f ms | naive12 %.3f | hisq18/18 %.3f | hisq18/12 %.3f | clocks %s" % (lv, d["ms_per_step"], d["solve"]["cg_iterations"], d["roofline"]["ms_per_launch"],
   v["naive_reunit_links12"]["time_to_solution_ms"], v["hisq_fat_naik_links18_18"]["time_to_solution_ms"], v["hisq_fat_naik_reunit_links18_12"]["time_to_solution_ms"], d["clocks"].get("sm_mhz")))
PY
done 2>&1 | tee $O/r03_pdl_ab.log
