#!/bin/bash
# round 3, call 5 (8 GPUs): every 8-rank case of tests/test_gpu_multi.py on HEAD, then the default bench line at N=8
HASH=$1
O=gpurun_out
bash tools/r03_call3.sh $HASH 8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 --no-links > $O/r03_bench_n8.json 2> $O/r03_bench_n8.err
echo "bench n8 rc=$?" | tee $O/r03_bench_n8.rc
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r03_bench_n8.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["parity"], d["roofline"].get("hisq_weak"), d["roofline"]["frac"])
PY
