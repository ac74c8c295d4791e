#!/bin/bash
OUT=gpurun_out/r2x
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "hot_path or step_host" > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest.log
timeout 900 python bench.py --no-other-configs --no-extension-modes --no-cpu-baseline --e2e-prefetch 0 > $OUT/bench_noprefetch.json 2> $OUT/bench_noprefetch.err; echo "bench0 exit $?"
timeout 900 python bench.py --no-other-configs --no-extension-modes --no-cpu-baseline > $OUT/bench_prefetch.json 2> $OUT/bench_prefetch.err; echo "bench1 exit $?"
python - <<'PY'
import json
for n in ("noprefetch","prefetch"):
    d=json.loads(open('gpurun_out/r2x/bench_%s.json'%n).read().strip().splitlines()[-1])
    print(n, d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['kernels']['region_pool_embed']['frac'], d['kernels']['sim_topk']['frac'])
PY
