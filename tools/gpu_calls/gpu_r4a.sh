#!/bin/bash
# first GPU run of the persistent TMA-staged R-MAC kernel (rmac_staged.cu): parity, then the A/B against the cluster kernel
OUT=gpurun_out/r4a
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -15 $OUT/pytest_rmac.log
timeout 300 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; cat $OUT/kbench_rmac.jsonl; tail -5 $OUT/kbench_rmac.err
timeout 400 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac_staged_kernel_many" > $OUT/memcheck_rmac.log 2>&1; echo "memcheck exit $?"; tail -8 $OUT/memcheck_rmac.log
