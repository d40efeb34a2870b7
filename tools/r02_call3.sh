#!/bin/bash
# Round-2 GPU call 3 (two B200): the new bench.py at N=1 and N=2, the one failing split-lattice link case with its
# output, role times of the fat+Naik sweep with persistent pack blocks, c3 at N=2, ncu of k_sweep_mr on a 2-rank job.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
QEXB200_ARENA_FIELDS=128 timeout 300 $TR --nproc-per-node 2 --master-port 29511 tests/multi_worker.py --lat 8,8,8,8 --rankgeom 1,1,1,2 \
  --inner 1,2,2,1 --links > gpurun_out/r02_links_case.log 2>&1
grep -v "^$" gpurun_out/r02_links_case.log | grep -B2 -A12 "Error\|assert" | head -60
timeout 600 python bench.py --steps 2 --ref-solves 2 --no-links > gpurun_out/r02_bench_n1_quick.json 2> gpurun_out/r02_bench_n1_quick.err
tail -c 600 gpurun_out/r02_bench_n1_quick.json; tail -3 gpurun_out/r02_bench_n1_quick.err
timeout 600 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --steps 2 --no-links > gpurun_out/r02_bench_n2_quick.json 2> gpurun_out/r02_bench_n2_quick.err
tail -c 600 gpurun_out/r02_bench_n2_quick.json; tail -3 gpurun_out/r02_bench_n2_quick.err
for pb in 148 296 592; do
  QEXB200_PACK_BLOCKS=$pb timeout 300 $TR --nproc-per-node 2 --master-port 29521 tools/role_times.py --lat 32,32,32,64 --rankgeom 1,1,1,2 --nlinks 8 \
    > gpurun_out/r02_roles_n2_nl8_pb$pb.json 2> gpurun_out/r02_roles_n2_nl8_pb$pb.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02_roles_n2_nl8_pb$pb.json"))
r=d["ranks"][0]
print("pack_blocks $pb: sweep %.4f ms, cg iter %.4f ms, pack_end %.1f us, kernel %.1f us" % (d["sweep_ms_events"], d["cg_iter_ms_events"], r["back_to_back_sweeps"]["pack_end_us"], r["back_to_back_sweeps"]["kernel_us"]))
PY
done
timeout 300 $TR --nproc-per-node 2 --master-port 29522 tools/role_times.py --lat 32,32,32,64 --rankgeom 1,1,1,2 --nlinks 4 \
    > gpurun_out/r02_roles_n2_nl4.json 2> gpurun_out/r02_roles_n2_nl4.err
timeout 600 $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --config c3 --steps 1 > gpurun_out/r02_bench_c3_n2.json 2> gpurun_out/r02_bench_c3_n2.err
tail -c 400 gpurun_out/r02_bench_c3_n2.json; tail -3 gpurun_out/r02_bench_c3_n2.err
# ncu over both ranks of a 2-rank job (the guard runs the command once without ncu first)
QEXB200_P2P_TIMEOUT_S=20 timeout 200 ncu --target-processes all --set full --import-source on --clock-control none -k regex:k_sweep_mr -c 4 \
  -o gpurun_out/r02_sweep_mr_n2 -f $TR --nproc-per-node 2 --master-port 29533 tools/role_times.py --lat 32,32,32,64 --rankgeom 1,1,1,2 --short \
  > gpurun_out/r02_ncu_mr.log 2>&1
tail -4 gpurun_out/r02_ncu_mr.log
ls -la gpurun_out | tail -8
