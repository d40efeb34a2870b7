#!/bin/bash
# What the next GPU session should run first (one B200, ~2 GPU-minutes), results under gpurun_out/:
#   1. the whole GPU suite;
#   2. A/B of the staple-derivative ring kernel (experimental, QEXB200_DERIV_PIPE=1): bit-identity + time;
#   3. launch list + one `ncu --set full` capture of k_staple_pipe (none was taken in round 1: budget spent).
# Usage on the box:  bash tools/first_gpu_call.sh
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/gpu_tests.log
python tools/ab_deriv_pipe.py 2>&1 | tail -6 | tee gpurun_out/ab_deriv_pipe.log
WHICH=hisq,nhyp REPS=1 python tools/prof_links.py > gpurun_out/prof_links_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_links_pipe.csv \
  env WHICH=hisq,nhyp REPS=1 python tools/prof_links.py > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_staple_pipe" -c 12 -o gpurun_out/staple_pipe_full -f \
  env WHICH=nhyp REPS=1 python tools/prof_links.py > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -8
