#!/bin/bash
# L2 prefetch of the re-score operands during the search: parity, then same-box A/B (off / on with several start delays)
OUT=gpurun_out/r3l
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_guards.py tests/test_gpu_sharded.py tests/test_gpu_full_size.py -q -m gpu -k "sim or search or topk or sharded or hot_path or classifier" > $OUT/pytest_sim.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -4 $OUT/pytest_sim.log
for rep in 1 2; do
  for t in sim_prefetch=1 sim_prefetch=0 sim_prefetch=2 sim_prefetch=1000 sim_prefetch=10000 sim_prefetch=30000; do
    timeout 300 python tools/sim_ab.py --tune $t >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  done
done
cat $OUT/sim_ab.jsonl
