#!/bin/bash
# Round-2 GPU call 4 (eight B200): per-role timing of the multi-rank sweep on both 8-GPU geometries, the weak-scaling
# bench line at 8, and the 8-rank cases of the multi-GPU parity suite.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run_roles() {  # tag lat geom nlinks
  timeout 240 $TR --nproc-per-node 8 --master-port 29521 tools/role_times.py --lat $2 --rankgeom $3 --nlinks $4 \
    2> gpurun_out/r02_roles_n8_$1.err | grep '^{' > gpurun_out/r02_roles_n8_$1.json
  python - <<PY
import json
d=json.load(open("gpurun_out/r02_roles_n8_$1.json"))
print("$1: sweep %.4f ms, cg iter %.4f ms" % (d["sweep_ms_events"], d["cg_iter_ms_events"]))
for r in d["ranks"]:
    c=r["cg"]; b=r["back_to_back_sweeps"]
    print("  rank %d: b2b kernel %.1f pack_end %.1f int_end %.1f face %.1f-%.1f waits %.1f/launch max %.1f | cg kernel %.1f waitsum %.1f xrank %.1f max %.1f"
          % (r["rank"], b["kernel_us"], b["pack_end_us"], b["interior_end_us"], b["face_start_us"], b["face_end_us"], b["wait_sum_per_launch_us"], b["wait_max_us"],
             c["kernel_us"], c["wait_sum_per_launch_us"], c.get("xrank_sum_mean_us", 0), c.get("xrank_sum_max_us", 0)))
PY
}
run_roles naive_1124 32,32,64,128 1,1,2,4 4
run_roles naive_1222 32,64,64,64 1,2,2,2 4
run_roles hisq_1124 32,32,64,128 1,1,2,4 8
timeout 400 $TR --nproc-per-node 8 --master-port 29512 bench.py --gpus 8 --steps 3 --no-links --no-cold --recon 18 \
  2> gpurun_out/r02_bench_n8.err | grep '^{' > gpurun_out/r02_bench_n8.json
tail -c 300 gpurun_out/r02_bench_n8.json; tail -2 gpurun_out/r02_bench_n8.err
QEXB200_P2P_TIMEOUT_S=20 QEXB200_TEST_NPROC=8 timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -rA 2>&1 | grep -v "^$" | tail -30 > gpurun_out/r02_multi_tests_n8.log
tail -12 gpurun_out/r02_multi_tests_n8.log
