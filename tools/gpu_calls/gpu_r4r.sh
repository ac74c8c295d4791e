#!/bin/bash
# final tree on two GPUs: whole -m gpu suite (with the two-GPU exchange test), default bench line at N = 2
OUT=gpurun_out/r4r
mkdir -p $OUT
timeout 900 python -m pytest tests -x -q -m gpu > $OUT/pytest_gpu_2gpus.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest_gpu_2gpus.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 > $OUT/bench_2gpu.json 2> $OUT/bench_2gpu.err; echo "bench2 exit $?"
python - <<'P'
import json
d = json.load(open("gpurun_out/r4r/bench_2gpu.json"))
print("value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"], d["ms_per_step_per_rank"])
c5 = d["other_configs"]["c5"]
print("c5", c5["value"], c5["ms_per_step"], "e2e", c5["e2e"]["value"], c5["c5_parity"], c5.get("exchange"))
P
