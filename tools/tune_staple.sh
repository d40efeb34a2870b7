#!/bin/bash
# Rebuilds smear.cu with different launch bounds on the GPU box and times the HISQ / nHYP smears (32^4).
set -e
cd "$(dirname "$0")/.."
python -m qex_b200.build >/dev/null 2>&1
for cfg in "128 2" "128 3" "128 4" "64 4" "64 6" "256 1" "256 2"; do
  set -- $cfg
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -ccbin /usr/bin/g++ \
    -DQB_SB=$1 -DQB_STAPLE_MINB=$2 -c qex_b200/csrc/smear.cu -o qex_b200/build/smear.o 2>/dev/null
  nvcc -shared -o qex_b200/libqex_b200.so qex_b200/build/*.o -ccbin /usr/bin/g++ -ldl 2>/dev/null
  echo "== block $1 minblocks $2"
  WHICH=hisq,nhyp REPS=3 python tools/prof_links.py | tail -3
done
