#!/bin/bash
mkdir -p gpurun_out/r2g
timeout 60 python tools/sim_pair_check.py 1 > gpurun_out/r2g/pair_check_first.log 2>&1; echo "first shape exit $?"; tail -3 gpurun_out/r2g/pair_check_first.log
timeout 180 python tools/sim_pair_check.py > gpurun_out/r2g/pair_check.log 2>&1; echo "pair check exit $?"; tail -14 gpurun_out/r2g/pair_check.log
for plan in 1 2; do
  timeout 150 python tools/kbench.py sim --iters 5 --tune sim_plan=$plan > gpurun_out/r2g/kbench_plan$plan.jsonl 2> gpurun_out/r2g/kbench_plan$plan.err; echo "plan $plan exit $?"
  cat gpurun_out/r2g/kbench_plan$plan.jsonl | cut -c1-200
done
nvidia-smi --query-gpu=name,clocks.sm,temperature.gpu,power.draw --format=csv
