#!/bin/bash
# warp-specialised staged R-MAC kernel with the batched norm reductions in the region warps: parity, timeline, A/B
OUT=gpurun_out/r4u
mkdir -p $OUT
timeout 120 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -3 $OUT/pytest_rmac.log
CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so timeout 200 python tools/rmac_prof.py > $OUT/rmac_prof.txt 2>&1; echo "prof exit $?"; grep -A7 "ws_kernel" $OUT/rmac_prof.txt | grep -v "^--"
timeout 120 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; grep "nhwc" $OUT/kbench_rmac.jsonl | grep -v "B=32\|4 stages\|cluster"; tail -5 $OUT/kbench_rmac.err
