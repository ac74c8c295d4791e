#!/bin/bash
# rmac_rows_kernel (channels-last maps straight from global memory into registers, no ring): parity, A/B against the staged and cluster kernels
OUT=gpurun_out/r4o
mkdir -p $OUT
timeout 150 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -5 $OUT/pytest_rmac.log
timeout 200 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; grep "nhwc" $OUT/kbench_rmac.jsonl | grep -v "B=32\|4 stages"; tail -5 $OUT/kbench_rmac.err
