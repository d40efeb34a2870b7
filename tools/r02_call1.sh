#!/bin/bash
# Round-2 GPU call 1 (one B200): whole GPU suite, big-lattice sweeps, ncu DRAM traffic of one sweep per size.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r02_gpu_tests.log
tail -3 gpurun_out/r02_gpu_tests.log
python tools/big_sweep.py > gpurun_out/r02_big_sweep.jsonl 2> gpurun_out/r02_big_sweep.err
cat gpurun_out/r02_big_sweep.jsonl; tail -3 gpurun_out/r02_big_sweep.err
for lat in 32,32,32,32 48,48,48,96 64,64,64,128; do
  tag=$(echo $lat | tr ',' 'x')
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct \
    --clock-control none -k regex:k_sweep --csv --log-file gpurun_out/r02_ncu_sweep_$tag.csv \
    python tools/big_sweep.py --short --lat $lat > gpurun_out/r02_ncu_sweep_$tag.log 2>&1
  tail -2 gpurun_out/r02_ncu_sweep_$tag.log
done
ls -la gpurun_out | tail -12
