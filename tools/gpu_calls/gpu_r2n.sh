#!/bin/bash
mkdir -p gpurun_out/r2n
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_guards.py tests/test_gpu_runtime.py tests/test_gpu_sharded.py -x -q -m gpu -k "sim or nearest or topk or classifier or shard or search or scatter or hot_path" > gpurun_out/r2n/pytest_sim.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r2n/pytest_sim.log
timeout 240 python tools/sim_pair_check.py > gpurun_out/r2n/pair_check.log 2>&1; echo "pair check exit $?"; tail -3 gpurun_out/r2n/pair_check.log
timeout 150 python tools/kbench.py sim --iters 5 --tune sim_plan=1 > gpurun_out/r2n/kbench_plan1.jsonl 2> gpurun_out/r2n/kb.err; cat gpurun_out/r2n/kbench_plan1.jsonl | cut -c1-190
