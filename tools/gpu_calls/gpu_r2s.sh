#!/bin/bash
# balanced streaming MAC kernel: parity + same-box A/B against the cluster kernel, CTAs-per-SM sweep, in-step number
OUT=gpurun_out/r2s
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "mac or rmac" > $OUT/pytest_mac.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_mac.log
for t in "mac_kernel=1" "mac_ctas_per_sm=2" "mac_ctas_per_sm=3" "mac_ctas_per_sm=4" "mac_ctas_per_sm=5" "mac_ctas_per_sm=6" "mac_ctas_per_sm=8"; do
  echo "== $t" >> $OUT/kbench_pool.jsonl
  timeout 300 python tools/kbench.py pool --iters 20 --tune $t 2>> $OUT/kbench_pool.err | grep "MAC" | grep "nhwc" >> $OUT/kbench_pool.jsonl
done
cat $OUT/kbench_pool.jsonl
timeout 900 python bench.py --no-other-configs --no-extension-modes --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2s/bench.json').read().strip().splitlines()[-1])
print(d['value'], d['kernels']['region_pool_embed'])
PY
