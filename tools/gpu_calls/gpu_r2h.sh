#!/bin/bash
mkdir -p gpurun_out/r2h
CMD="python tools/kbench.py sim --quick --iters 2 --tune sim_plan=2"
timeout 200 $CMD > gpurun_out/r2h/plain.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_active.avg,smsp__cycles_active.avg --clock-control none -c 200 --csv --log-file gpurun_out/r2h/sim_pair_launches.csv $CMD > gpurun_out/r2h/ncu.log 2>&1; echo ncu $?
