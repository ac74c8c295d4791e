#!/bin/bash
# sim kernel: epilogue groups + linear ranges: parity first (with a timeout: a pipeline bug would hang), then same-box A/B
OUT=gpurun_out/r2u
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_sharded.py -q -m gpu -x -k "sim or search or topk or sharded" > $OUT/pytest_sim.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -5 $OUT/pytest_sim.log
if [ $rc -ne 0 ]; then exit 0; fi
OLD=$PWD/cvpce_b200/variants/libold.so
CVPCE_B200_SO=$OLD timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
for t in "sim_plan=1 sim_epi_groups=1" "sim_plan=1 sim_epi_groups=2" "sim_plan=1 sim_epi_groups=4" "sim_plan=2 sim_epi_groups=1" "sim_plan=2 sim_epi_groups=2" "sim_plan=2 sim_epi_groups=4"; do
  timeout 300 python tools/sim_ab.py --tune $t >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
done
CVPCE_B200_SO=$OLD timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
cat $OUT/sim_ab.jsonl
