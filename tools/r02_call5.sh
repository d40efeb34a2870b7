#!/bin/bash
# Round-2 GPU call 5 (one B200): the configs the metric is quoted on at N=1 (c3 48^3x96 HISQ, c5 64^3x128 HISQ, c4 64^4
# nHYP trajectory solve sequence) and the launch list of the default bench command under ncu.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for c in c4 c3 c5; do
  timeout 900 python bench.py --config $c --steps 1 2> gpurun_out/r02_bench_${c}_n1.err | grep '^{' > gpurun_out/r02_bench_${c}_n1.json
  head -c 400 gpurun_out/r02_bench_${c}_n1.json; echo; tail -3 gpurun_out/r02_bench_${c}_n1.err
done
timeout 600 python bench.py --steps 2 --warmup 3 --no-links --no-cpu --no-cold --naive-only --recon 18 --no-parity > gpurun_out/r02_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r02_bench.csv \
  python bench.py --steps 2 --warmup 3 --no-links --no-cpu --no-cold --naive-only --recon 18 --no-parity > gpurun_out/r02_ncu_bench.log 2>&1
tail -2 gpurun_out/r02_ncu_bench.log | cut -c1-300
ls -la gpurun_out | tail -6
