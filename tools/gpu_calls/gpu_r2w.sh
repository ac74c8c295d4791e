#!/bin/bash
# sim kernel: software-pipelined TMEM loads in the epilogue: parity, then same-box A/B against the previous build (old, new, old, new)
OUT=gpurun_out/r2w
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_sharded.py -q -m gpu -x -k "sim or search or topk or sharded" > $OUT/pytest_sim.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -5 $OUT/pytest_sim.log
if [ $rc -ne 0 ]; then exit 0; fi
OLD=$PWD/cvpce_b200/variants/libold.so
for rep in 1 2; do
  CVPCE_B200_SO=$OLD timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  for v in $EXTRA_VARIANTS; do CVPCE_B200_SO=$PWD/cvpce_b200/variants/$v timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err; done
done
cat $OUT/sim_ab.jsonl
