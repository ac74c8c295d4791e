#!/bin/bash
OUT=gpurun_out/r3e
mkdir -p $OUT
rm -f $OUT/sim_ab.jsonl
for rep in 1 2; do
  timeout 300 python tools/sim_ab.py >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py --tune sim_debug=3 >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py --tune sim_plan=1 sim_keys=1 >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
  timeout 300 python tools/sim_ab.py --tune sim_plan=1 sim_keys=1 sim_debug=3 >> $OUT/sim_ab.jsonl 2>> $OUT/sim_ab.err
done
cat $OUT/sim_ab.jsonl
