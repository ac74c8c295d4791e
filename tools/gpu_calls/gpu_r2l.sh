#!/bin/bash
# 8 GPUs: default bench under torchrun at N = 8 and 4 (c3 headline + other_configs.c5), c5 alone with both exchanges / chunk counts
mkdir -p gpurun_out/r2l
run() { # name nproc args...
  local name=$1; local n=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" > gpurun_out/r2l/$name.json 2> gpurun_out/r2l/$name.err; echo "$name exit $?"
}
timeout 300 python -m pytest tests/test_gpu_sharded.py -x -q -m gpu > gpurun_out/r2l/pytest_sharded.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r2l/pytest_sharded.log
run bench_8gpu 8 --steps 5 --warmup 3
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2l/bench_8gpu.json').read().strip().splitlines()[-1])
print('c3 x8', l['value'], l['ms_per_step'], l['ms_per_step_per_rank'], 'e2e', l['e2e']['value'], l['e2e']['ms_per_step_per_rank'])
c=l['other_configs']['c5']; print('c5 x8', c['value'], c['ms_per_step'], c['ms_per_step_per_rank'], c['tflops_per_gpu'], 'e2e', c['e2e']['value'], c['e2e']['ms_per_step'], c['c5_parity'], c['c5_parity_detail'])
PY
run bench_4gpu 4 --steps 5 --warmup 3 --no-extension-modes
run c5_8gpu_p2p_chunks1 8 --workload c5 --steps 10 --warmup 3 --c5-chunks 1
run c5_8gpu_p2p_chunks2 8 --workload c5 --steps 10 --warmup 3 --c5-chunks 2
run c5_8gpu_nccl 8 --workload c5 --exchange nccl --steps 10 --warmup 3
for f in bench_4gpu c5_8gpu_p2p_chunks1 c5_8gpu_p2p_chunks2 c5_8gpu_nccl; do python - $f <<'PY'
import json,sys
l=json.loads(open('gpurun_out/r2l/%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
c=l['other_configs']['c5'] if 'other_configs' in l else l
print(sys.argv[1], l['value'], l['ms_per_step'], '| c5:', c['value'], c['ms_per_step'], 'e2e', c['e2e']['value'], c['e2e']['ms_per_step'], c['c5_parity'])
PY
done
