#!/bin/bash
# final tree (trunk batch 800) under torchrun at N = 2: the default bench line, as the driver launches it
OUT=gpurun_out/r4z
mkdir -p $OUT
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 > $OUT/bench_2gpu.json 2> $OUT/bench_2gpu.err; echo "bench2 exit $?"
python - <<'P'
import json
txt = open("gpurun_out/r4z/bench_2gpu.json").read()
d = json.loads([l for l in txt.splitlines() if l.strip().startswith("{")][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"], d["ms_per_step_per_rank"], "MAC", d["kernels"]["region_pool_embed"]["frac"])
c5 = d["other_configs"]["c5"]
print("c5", c5["value"], c5["ms_per_step"], "e2e", c5["e2e"]["value"], c5["c5_parity"])
P
