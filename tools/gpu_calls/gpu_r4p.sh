#!/bin/bash
# staged R-MAC kernel two-step region stage (column-segment maxima, then regions): parity, then the A/B
OUT=gpurun_out/r4p
mkdir -p $OUT
timeout 150 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -5 $OUT/pytest_rmac.log
timeout 200 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; grep "nhwc" $OUT/kbench_rmac.jsonl | grep -v "B=32"; tail -5 $OUT/kbench_rmac.err
