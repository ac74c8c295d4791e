#!/bin/bash
mkdir -p gpurun_out/r2j
timeout 240 python tools/sim_pair_check.py > gpurun_out/r2j/pair_check.log 2>&1; echo "pair check exit $?"; tail -12 gpurun_out/r2j/pair_check.log
for cfg in "sim_plan=1" "sim_plan=2 sim_epi_groups=1" "sim_plan=2 sim_epi_groups=2" "sim_plan=2 sim_epi_groups=2 sim_debug=1" "sim_plan=2 sim_epi_groups=2 sim_debug=2"; do
  timeout 150 python tools/kbench.py sim --iters 5 --tune $cfg > gpurun_out/r2j/kb.jsonl 2> gpurun_out/r2j/kb.err; echo "$cfg exit $?"
  cat gpurun_out/r2j/kb.jsonl >> gpurun_out/r2j/kbench_all.jsonl
  grep -v "Q=200" gpurun_out/r2j/kb.jsonl | cut -c1-190
done
