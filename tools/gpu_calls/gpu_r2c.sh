#!/bin/bash
# ncu --set full of the column crop kernel (NHWC4 f32, the bench's layout) + the crop band-path test again
mkdir -p gpurun_out/r2c
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "crop" > gpurun_out/r2c/pytest_crop.log 2>&1; echo pytest $?
tail -3 gpurun_out/r2c/pytest_crop.log
CMD="python tools/kbench.py crop --iters 1 --sizes 256"
timeout 200 $CMD > gpurun_out/r2c/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:crop_resize_col_kernel -s 8 -c 4 -o gpurun_out/r2c/crop_col $CMD > gpurun_out/r2c/ncu.log 2>&1
echo ncu $?
ls -la gpurun_out/r2c
