#!/bin/bash
# staged R-MAC kernel, NCHW: segmented-shuffle cell pass + two-step region pass: parity, phase counters, A/B
OUT=gpurun_out/r4m
mkdir -p $OUT
timeout 150 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py -q -m gpu -x -k "rmac or image_kernels" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -5 $OUT/pytest_rmac.log
CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so timeout 200 python tools/rmac_prof.py > $OUT/rmac_prof.txt 2>&1; echo "prof exit $?"; grep -A12 "B=128 nchw" $OUT/rmac_prof.txt
timeout 200 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; grep "nchw" $OUT/kbench_rmac.jsonl | grep -v "B=32"; grep "B=256 nhwc f32" $OUT/kbench_rmac.jsonl; tail -5 $OUT/kbench_rmac.err
