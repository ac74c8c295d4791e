#!/bin/bash
mkdir -p gpurun_out/r2p
i=0
for cfg in "sim_plan=1" "sim_plan=2 sim_epi_groups=1" "sim_plan=2 sim_epi_groups=1 sim_debug=1" "sim_plan=2 sim_epi_groups=2"; do
  i=$((i+1))
  CMD="python tools/kbench.py sim --quick --iters 1 --tune $cfg"
  timeout 200 $CMD > gpurun_out/r2p/plain$i.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__cycles_active.avg --clock-control none -c 60 --csv --log-file gpurun_out/r2p/launches$i.csv $CMD > gpurun_out/r2p/ncu$i.log 2>&1; echo "$cfg ncu $?"
done
