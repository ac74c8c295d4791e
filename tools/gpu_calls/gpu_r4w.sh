#!/bin/bash
# trunk batch 256 (default) against 512 and 384: headline-only bench runs, same box, alternating
OUT=gpurun_out/r4w
mkdir -p $OUT
ARGS="--steps 3 --warmup 2 --no-other-configs --no-extension-modes --no-cpu-baseline"
for tb in 256 512 384 256 512; do
  timeout 300 python bench.py $ARGS --trunk-batch $tb > $OUT/bench_tb${tb}_$RANDOM.json 2>> $OUT/err.log; echo "tb $tb exit $?"
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/r4w/bench_tb*.json")):
    d = json.load(open(f))
    k = d["kernels"]
    print(f.split("/")[-1], "value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"], "MAC", k["region_pool_embed"]["avg_ms"], k["region_pool_embed"]["frac"],
          "maxpool", k["maxpool2x2"]["avg_ms"], k["maxpool2x2"]["frac"], "clk", d["clocks"]["sm_mhz"])
P
