#!/bin/bash
mkdir -p gpurun_out/r2i
for mode in 0 1 2 3 4 5 6; do
  timeout 150 python tools/kbench.py sim --quick --iters 5 --tune sim_plan=2 sim_debug=$mode > gpurun_out/r2i/kbench_mode$mode.jsonl 2> gpurun_out/r2i/kbench_mode$mode.err; echo "mode $mode exit $?"
  grep "bf16-only" gpurun_out/r2i/kbench_mode$mode.jsonl | cut -c1-200
done
timeout 120 python - <<'PY' > gpurun_out/r2i/mode3_check.log 2>&1
import sys, torch
sys.path.insert(0, '.')
from cvpce_b200 import _lib, ops
gen = torch.Generator(device='cuda').manual_seed(1)
ok = True
for Q, G, D, k in ((1000, 8192, 1024, 5), (5000, 70000, 512, 5), (300, 5000, 512, 5), (129, 777, 64, 3)):
    q = torch.randn((Q, D), device='cuda', generator=gen); g = torch.randn((G, D), device='cuda', generator=gen)
    qb, qi = ops.rows_prepare(q); gb, gi = ops.rows_prepare(g)
    with _lib.tuning(sim_plan=1):
        a = ops.sim_topk(qb, gb, k, 0, q, g, qi, gi)
    with _lib.tuning(sim_plan=2, sim_debug=3):
        b = ops.sim_topk(qb, gb, k, 0, q, g, qi, gi)
    same = torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]); ok = ok and same
    print(Q, G, D, k, same)
print('MODE3 OK' if ok else 'MODE3 MISMATCH')
PY
tail -5 gpurun_out/r2i/mode3_check.log
