#!/bin/bash
# staged R-MAC kernel, version 5b (region stage with 8 loads in flight, batched norm reductions; NCHW: 4 rows per batch, incremental item decoding, phase timers)
OUT=gpurun_out/r4g
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac" > $OUT/pytest_rmac.log 2>&1; echo "rmac tests exit $?"; tail -15 $OUT/pytest_rmac.log
CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so timeout 300 python tools/rmac_prof.py > $OUT/rmac_prof.txt 2>&1; echo "prof exit $?"; grep -A12 "B=128" $OUT/rmac_prof.txt
timeout 300 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; grep -v "B=32" $OUT/kbench_rmac.jsonl; tail -5 $OUT/kbench_rmac.err
