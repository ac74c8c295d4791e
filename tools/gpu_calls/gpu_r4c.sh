#!/bin/bash
# staged R-MAC kernel, third version (chunks of whole runs, idle warps off the walk; NCHW: 16-byte strips, in-place column maxima)
OUT=gpurun_out/r4c
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -15 $OUT/pytest_rmac.log
timeout 300 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; cat $OUT/kbench_rmac.jsonl; tail -5 $OUT/kbench_rmac.err
