#!/bin/bash
# round 3, call 2 (2 GPUs): every 2-rank case of tests/test_gpu_multi.py on HEAD, the cache and rng tests, the bench line at
# N=2 (must exit 0 with roofline.parity.ok) and the same command on the fault-injected build (must exit 3).
HASH=$1
O=gpurun_out
echo "commit $HASH, 2 x B200, QEXB200_TEST_NPROC=2" > $O/r03_multi_tests_n2.log
QEXB200_TEST_NPROC=2 timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -rs >> $O/r03_multi_tests_n2.log 2>&1
tail -n 3 $O/r03_multi_tests_n2.log
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rng.py -m gpu -q -s -k "cache or plaquettes" > $O/r03_cache_rng.log 2>&1
grep -E "sum diff|passed|failed" $O/r03_cache_rng.log
B="--gpus 2 --steps 3 --warmup 3 --no-cpu --no-links --no-cold"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py $B > $O/r03_bench_n2.json 2> $O/r03_bench_n2.err
echo "good build rc=$?" | tee $O/r03_exit_codes.log
QEXB200_LIB=$PWD/qex_b200/build_fault/libqex_b200_fault.so timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py $B --naive-only > $O/r03_bench_n2_fault.json 2> $O/r03_bench_n2_fault.err
echo "fault-injected build rc=$?" | tee -a $O/r03_exit_codes.log
grep "FAILED checks" $O/r03_bench_n2_fault.err | head -2 | tee -a $O/r03_exit_codes.log
python - <<'PY'
import json
for f in ("gpurun_out/r03_bench_n2.json","gpurun_out/r03_bench_n2_fault.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["roofline"].get("parity"), d["roofline"]["frac"])
    except Exception as e:
        print(f, "no line", e)
PY
