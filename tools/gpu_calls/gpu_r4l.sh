#!/bin/bash
# MAC pooling inside the step: streaming loads with an L2 evict-first hint (variant build) against the shipped loads, same box, alternating
OUT=gpurun_out/r4l
mkdir -p $OUT
ARGS="--steps 3 --warmup 2 --no-other-configs --no-extension-modes --no-cpu-baseline"
for i in 1 2; do
  timeout 300 python bench.py $ARGS > $OUT/bench_base_$i.json 2> $OUT/err.log; echo "base $i exit $?"
  CVPCE_B200_SO=cvpce_b200/variants/libpoolef.so timeout 300 python bench.py $ARGS > $OUT/bench_evict_first_$i.json 2>> $OUT/err.log; echo "evict-first $i exit $?"
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/r4l/bench_*.json")):
    d = json.load(open(f))
    k = d["kernels"]["region_pool_embed"]
    print(f.split("/")[-1], "value", d["value"], "ms/step", d["ms_per_step"], "MAC avg_ms", k["avg_ms"], "frac", k["frac"])
P
