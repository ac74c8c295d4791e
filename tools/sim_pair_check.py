#!/usr/bin/env python
"""Development check of the CTA-pair (cta_group::2) search kernel against the round-1 kernel and the exact fp32 kernel:
bit-identical results on shapes that exercise whole-group rounds, the flattened remainder (several lists per row), odd
query-block counts, gallery tails, both re-score modes, D = 64 .. 1024 (resident and streaming query block).
Run under `timeout`: a protocol mistake in a 2-CTA kernel shows up as a hang."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import _lib, ops  # noqa: E402

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(11)
shapes = [(256, 256, 64, 1), (300, 5000, 512, 5), (129, 777, 64, 3), (1000, 8192, 1024, 5), (6400, 8192, 1024, 5), (64, 180, 1024, 1),
          (2500, 30000, 512, 5), (40000, 3000, 128, 5), (257, 100000, 256, 20), (5000, 70000, 512, 5)]
if len(sys.argv) > 1:
    shapes = shapes[: int(sys.argv[1])]
ok = True
for Q, G, D, k in shapes:
    q = torch.randn((Q, D), device=dev, generator=gen)
    g = torch.randn((G, D), device=dev, generator=gen)
    qb, qi = ops.rows_prepare(q)
    gb, gi = ops.rows_prepare(g)
    res = {}
    for plan, ew in ((1, 0), (2, 1), (2, 2)):
        with _lib.tuning(sim_plan=plan, sim_epi_groups=ew):
            res[plan + ew] = (ops.sim_topk(qb, gb, k, 5, q, g, qi, gi), ops.sim_topk(qb, gb, k, 5))
        torch.cuda.synchronize()
    same_exact = all(torch.equal(res[1][0][0], res[x][0][0]) and torch.equal(res[1][0][1], res[x][0][1]) for x in (3, 4))
    same_bf16 = all(torch.equal(res[1][1][0], res[x][1][0]) and torch.equal(res[1][1][1], res[x][1][1]) for x in (3, 4))
    line = "Q=%d G=%d D=%d k=%d: exact %s, bf16-only %s" % (Q, G, D, k, same_exact, same_bf16)
    if Q * G <= 3e8:
        sf, i_f = ops.sim_topk_f32(q, g, k, 5, normalize=True)
        agree = float((i_f == res[4][0][1]).float().mean())
        line += ", index agreement with the fp32 kernel %.5f" % agree
        ok = ok and agree > 0.999
    print(line, flush=True)
    ok = ok and same_exact and same_bf16
print("PAIR KERNEL OK" if ok else "PAIR KERNEL MISMATCH", flush=True)
sys.exit(0 if ok else 1)
