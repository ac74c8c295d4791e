#!/bin/bash
# key lists + linear ranges: parity, then same-box A/B (shipped kernel, grid+keys, linear+keys)
OUT=gpurun_out/r3d
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_sharded.py tests/test_gpu_full_size.py -q -m gpu -k "sim or search or topk or sharded or hot_path or classifier" > $OUT/pytest_sim.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -4 $OUT/pytest_sim.log
OLD=$PWD/cvpce_b200/variants/libold.so
for rep in 1 2; do
  CVPCE_B200_SO=$OLD timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py --tune sim_plan=1 >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py --tune sim_plan=2 >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
done
cat $OUT/sim_ab.jsonl
