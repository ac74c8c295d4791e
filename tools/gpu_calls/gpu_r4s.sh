#!/bin/bash
# warp-specialised staged R-MAC kernel (16 streaming warps + 8 region warps): parity under short timeouts, then the A/B
OUT=gpurun_out/r4s
mkdir -p $OUT
timeout 90 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac_staged_kernel_vgg" > $OUT/pytest_rmac_a.log 2>&1; echo "vgg tests exit $?"; tail -5 $OUT/pytest_rmac_a.log
timeout 90 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "rmac_staged_kernel_many" > $OUT/pytest_rmac_b.log 2>&1; echo "many tests exit $?"; tail -5 $OUT/pytest_rmac_b.log
timeout 120 python tools/kbench.py rmac --iters 7 > $OUT/kbench_rmac.jsonl 2> $OUT/kbench_rmac.err; echo "kbench exit $?"; grep "nhwc" $OUT/kbench_rmac.jsonl | grep -v "B=32\|4 stages\|cluster"; tail -5 $OUT/kbench_rmac.err
