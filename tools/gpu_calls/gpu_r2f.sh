#!/bin/bash
# persistent multicast tcgen05 kernel: parity tests, then kbench sim for every plan
mkdir -p gpurun_out/r2f
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_guards.py tests/test_gpu_runtime.py tests/test_gpu_sharded.py -x -q -m gpu -k "sim or nearest or topk or classifier or shard or search or scatter or hot_path" > gpurun_out/r2f/pytest_sim.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/r2f/pytest_sim.log
for plan in 1 2 3 4; do
  timeout 150 python tools/kbench.py sim --iters 5 --tune sim_plan=$plan > gpurun_out/r2f/kbench_plan$plan.jsonl 2> gpurun_out/r2f/kbench_plan$plan.err; echo "plan $plan exit $?"
  cat gpurun_out/r2f/kbench_plan$plan.jsonl | cut -c1-200
done
