#!/bin/bash
# Round-2 GPU call 7 (one B200): what the driver runs at round end -- the whole GPU suite, smoke(), the default bench line.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r02_gpu_tests_final.log
tail -3 gpurun_out/r02_gpu_tests_final.log
python __graft_entry__.py smoke 2>&1 | tail -3 > gpurun_out/r02_smoke.log; cat gpurun_out/r02_smoke.log
python bench.py 2> gpurun_out/r02_bench_n1_final.err | grep '^{' > gpurun_out/r02_bench_n1_final.json
head -c 500 gpurun_out/r02_bench_n1_final.json; echo; tail -2 gpurun_out/r02_bench_n1_final.err
