#!/bin/bash
# Round-2 GPU call 2 (two B200): new single-GPU tests, the multi-GPU parity suite (2-rank cases), per-role timing of
# the one-kernel multi-rank sweep, and an ncu capture of it on rank 0 of a 2-rank job.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "solveEO or cgstate or reunit" 2>&1 | tail -8 > gpurun_out/r02_new_tests.log
tail -3 gpurun_out/r02_new_tests.log
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -25 > gpurun_out/r02_multi_tests_n2.log
tail -4 gpurun_out/r02_multi_tests_n2.log
for nl in 4 8; do
  timeout 300 $TR --nproc-per-node 2 --master-port 29521 tools/role_times.py --lat 32,32,32,64 --rankgeom 1,1,1,2 --nlinks $nl \
    > gpurun_out/r02_roles_n2_nl$nl.json 2> gpurun_out/r02_roles_n2_nl$nl.err
  tail -c 1500 gpurun_out/r02_roles_n2_nl$nl.json; tail -2 gpurun_out/r02_roles_n2_nl$nl.err
done
# ncu on rank 0 only; rank 1 runs the same loop beside it
export MASTER_ADDR=127.0.0.1 MASTER_PORT=29533 WORLD_SIZE=2
RANK=1 LOCAL_RANK=1 timeout 280 python tools/role_times.py --lat 32,32,32,64 --rankgeom 1,1,1,2 --short > gpurun_out/r02_ncu_mr_rank1.log 2>&1 &
RANK=0 LOCAL_RANK=0 timeout 280 ncu --set full --import-source on --clock-control none -k regex:k_sweep_mr -c 4 \
  -o gpurun_out/r02_sweep_mr_n2 -f python tools/role_times.py --lat 32,32,32,64 --rankgeom 1,1,1,2 --short > gpurun_out/r02_ncu_mr_rank0.log 2>&1
wait
tail -3 gpurun_out/r02_ncu_mr_rank0.log
ls -la gpurun_out | tail -12
