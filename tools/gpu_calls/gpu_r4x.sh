#!/bin/bash
# trunk batch 512 against 640, 800, 1280 and 1600: headline-only bench runs, same box, alternating
OUT=gpurun_out/r4x
mkdir -p $OUT
ARGS="--steps 3 --warmup 2 --no-other-configs --no-extension-modes --no-cpu-baseline"
for tb in 512 640 800 1280 1600 800 512; do
  timeout 300 python bench.py $ARGS --trunk-batch $tb > $OUT/bench_tb${tb}_$RANDOM.json 2>> $OUT/err.log; echo "tb $tb exit $?"
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/r4x/bench_tb*.json")):
    d = json.load(open(f))
    k = d["kernels"]
    print(f.split("/")[-1], "value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"], "MAC", k["region_pool_embed"]["avg_ms"], k["region_pool_embed"]["frac"],
          "maxpool", k["maxpool2x2"]["avg_ms"], k["maxpool2x2"]["frac"], "clk", d["clocks"]["sm_mhz"])
P
