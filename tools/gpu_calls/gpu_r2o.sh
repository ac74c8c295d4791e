#!/bin/bash
mkdir -p gpurun_out/r2o
timeout 240 python tools/sim_pair_check.py > gpurun_out/r2o/pair_check.log 2>&1; echo "pair check exit $?"; tail -12 gpurun_out/r2o/pair_check.log
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_guards.py tests/test_gpu_runtime.py tests/test_gpu_sharded.py -x -q -m gpu -k "sim or nearest or topk or classifier or shard or search or scatter or hot_path" > gpurun_out/r2o/pytest_sim.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r2o/pytest_sim.log
for cfg in "sim_plan=1" "sim_plan=2 sim_epi_groups=1" "sim_plan=2 sim_epi_groups=2"; do
  timeout 150 python tools/kbench.py sim --iters 5 --tune $cfg > gpurun_out/r2o/kb.jsonl 2> gpurun_out/r2o/kb.err; echo "$cfg exit $?"
  cat gpurun_out/r2o/kb.jsonl >> gpurun_out/r2o/kbench_all.jsonl
  grep -v "Q=200" gpurun_out/r2o/kb.jsonl | cut -c1-190
done
