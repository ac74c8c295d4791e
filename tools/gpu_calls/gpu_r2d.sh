#!/bin/bash
mkdir -p gpurun_out/r2d
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_full_size.py -x -q -m gpu -k "crop" > gpurun_out/r2d/pytest_crop.log 2>&1; echo pytest $?
tail -3 gpurun_out/r2d/pytest_crop.log
timeout 200 python tools/kbench.py crop --iters 5 > gpurun_out/r2d/kbench_col.jsonl 2>gpurun_out/r2d/kbench_col.err
cat gpurun_out/r2d/kbench_col.jsonl | cut -c1-200
