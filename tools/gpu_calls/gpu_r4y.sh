#!/bin/bash
# final tree of the round on one GPU: whole -m gpu suite, smoke(), default bench line
OUT=gpurun_out/r4y
mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
timeout 900 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench exit $?"; python - <<'P'
import json
d = json.load(open("gpurun_out/r4y/bench_default.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"], "roofline", d["roofline"]["kernel"], d["roofline"]["frac"], d["clocks"])
for k, v in d["kernels"].items(): print(" ", k, v["avg_ms"], v["frac"])
print("rmac_l3", d["extension_modes"]["rmac_l3"]["value"], d["extension_modes"]["rmac_l3"]["region_pool_embed"])
print("c5", d["other_configs"]["c5"]["value"], d["other_configs"]["c5"]["c5_parity"])
P
