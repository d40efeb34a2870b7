#!/bin/bash
# round 3, call 6 (1 GPU): the final tree -- GPU suite, the no-allocation test alone in a fresh process, smoke, the default
# bench line, and the ncu launch list of a short bench command (kernel SHARES of the step; absolute times are serialised)
HASH=$1
O=gpurun_out
echo "commit $HASH" > $O/r03_gpu_suite_final.log
python -m pytest tests -m gpu -x -q >> $O/r03_gpu_suite_final.log 2>&1; echo "suite rc=$?" >> $O/r03_gpu_suite_final.log; tail -n 3 $O/r03_gpu_suite_final.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r03_noalloc_alone_final.log 2>&1; echo "alone rc=$?" >> $O/r03_noalloc_alone_final.log; tail -n 2 $O/r03_noalloc_alone_final.log
python __graft_entry__.py smoke > $O/r03_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r03_smoke.log
python bench.py > $O/r03_bench_n1_final.json 2> $O/r03_bench_n1_final.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r03_bench_n1_final.json").read().strip().splitlines()[-1])
print("ms/step %.3f value %.1f e2e %.1f frac %.3f iter_frac %.3f launches %s cpu %.1f" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["cg_iteration_frac"], d["gpu_launches"], d["cpu_baseline"]["value"]))
print(d["roofline"]["parity"], d["roofline"].get("target"), d["roofline"].get("hisq_weak"))
PY
S="--steps 2 --warmup 3 --no-cpu --no-links --no-cold --naive-only --no-parity"
python bench.py $S > $O/r03_bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/launches_r03_bench.csv python bench.py $S > $O/r03_ncu.log 2>&1
echo "ncu rc=$?"; wc -l $O/launches_r03_bench.csv
