#!/bin/bash
# Round-2 GPU call 6 (N B200, N = $1): the strong-scaling configs at N GPUs: c3 (48^3x96 HISQ), c5 (64^3x128 HISQ,
# overlap on and -- at 8 -- off), and at 8 the c4 trajectory solve sequence (64^4 nHYP).
set -u
N=${1:-8}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
run() {  # tag, args...
  tag=$1; shift
  timeout 400 $TR --master-port 29540 bench.py --gpus $N --steps 1 "$@" 2> gpurun_out/r02_bench_${tag}_n$N.err | grep '^{' > gpurun_out/r02_bench_${tag}_n$N.json
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02_bench_${tag}_n$N.json"))
    print("${tag} N=$N: %.0f GF/s, %.1f ms/step, its %s, sweep %.4f ms frac %.3f, cg-iter frac %.3f, parity %s" % (d["value"], d["ms_per_step"], d["solve"]["cg_iterations"],
          d["roofline"]["ms_per_launch"], d["roofline"]["frac"], d["roofline"]["cg_iteration_frac"], (d.get("parity") or {}).get("ok")))
except Exception as e:
    print("${tag} N=$N FAILED", e); print(open("gpurun_out/r02_bench_${tag}_n$N.err").read()[-1500:])
PY
}
run c3 --config c3
run c5 --config c5 --no-parity
if [ "$N" = "8" ]; then
  run c5_overlap_off --config c5 --overlap off --no-parity
  run c4 --config c4 --no-parity
fi
