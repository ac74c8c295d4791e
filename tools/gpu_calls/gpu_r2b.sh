#!/bin/bash
# crop kernel A/B (round 2): parity tests touching the crop kernels, then kbench with the column kernel (default) and the band kernel
mkdir -p gpurun_out/r2b
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_full_size.py tests/test_gpu_runtime.py -x -q -m gpu -k "crop or hot_path or four_channel or float16" > gpurun_out/r2b/pytest_crop.log 2>&1; echo pytest $?
tail -5 gpurun_out/r2b/pytest_crop.log
timeout 200 python tools/kbench.py crop --iters 5 > gpurun_out/r2b/kbench_col.jsonl 2>gpurun_out/r2b/kbench_col.err
timeout 200 python tools/kbench.py crop --iters 5 --tune crop_kernel=1 > gpurun_out/r2b/kbench_band.jsonl 2>gpurun_out/r2b/kbench_band.err
for r in 32 128 256; do timeout 100 python tools/kbench.py crop --iters 5 --sizes 256 --tune crop_band_rows=$r >> gpurun_out/r2b/kbench_col_rows.jsonl 2>&1; done
cat gpurun_out/r2b/kbench_col.jsonl gpurun_out/r2b/kbench_band.jsonl gpurun_out/r2b/kbench_col_rows.jsonl | cut -c1-230
