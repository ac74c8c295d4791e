#!/bin/bash
# final code at 8 GPUs: the default bench line exactly as the driver launches it
OUT=gpurun_out/r3n
mkdir -p $OUT
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 > $OUT/bench_8gpu.json 2> $OUT/bench_8gpu.err; echo "bench8 exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3n/bench_8gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('ms_per_step_per_rank'), d['e2e']['value'])
c=d['other_configs']['c5']; print(c['value'], c['ms_per_step'], c['e2e']['value'], c['e2e']['ms_per_step'], c['c5_parity'], c.get('exchange'))
PY
