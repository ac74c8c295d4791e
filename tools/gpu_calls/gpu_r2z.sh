#!/bin/bash
# final code on 2 GPUs: the whole -m gpu suite (incl. the two-GPU exchange-form test), smoke, and the default bench line at N=2
OUT=gpurun_out/r3m
mkdir -p $OUT
timeout 1200 python -m pytest tests -q -m gpu > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -1 $OUT/smoke.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > $OUT/bench_2gpu.json 2> $OUT/bench_2gpu.err; echo "bench2 exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3m/bench_2gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['kernels']['region_pool_embed']['frac'], d['kernels']['crop_resize']['frac'])
c=d['other_configs']['c5']; print(c['value'], c['ms_per_step'], c['e2e']['value'], c['c5_parity'], c.get('c5_parity_detail'))
PY
