#!/bin/bash
# round 3, call 3 (N GPUs): every N-rank case of tests/test_gpu_multi.py on HEAD (p2p, NCCL, 12-/14-real, link pipeline)
HASH=$1; N=$2
O=gpurun_out
echo "commit $HASH, $N x B200, QEXB200_TEST_NPROC=$N" > $O/r03_multi_tests_n$N.log
QEXB200_TEST_NPROC=$N timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -q -rs >> $O/r03_multi_tests_n$N.log 2>&1
tail -n 3 $O/r03_multi_tests_n$N.log
