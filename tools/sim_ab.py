"""A/B of the search at the in-step C3 shape with in-step-like data (near-identical non-negative query rows, |N(0,1)| gallery):
CVPCE_B200_SO selects the library build."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import ops, _lib
tune = sys.argv[sys.argv.index("--tune") + 1:] if "--tune" in sys.argv else []
for t in tune:
    _lib.tuning_set(t.split("=")[0], int(t.split("=")[1]))
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(3)
g = torch.randn((8192, 1024), generator=gen, device=dev).abs_(); g /= g.norm(dim=1, keepdim=True)
base = torch.rand((1, 1024), generator=gen, device=dev)
res = {}
for name, q in (("dense", base + 0.05 * torch.rand((6400, 1024), generator=gen, device=dev)), ("random", torch.randn((6400, 1024), generator=gen, device=dev))):
    q = q / q.norm(dim=1, keepdim=True)
    gb, gi = ops.rows_prepare(g); flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for it in range(12):
        flush.zero_()
        qb, qi = ops.rows_prepare(q)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); s, i = ops.sim_topk(qb, gb, 5, 0, q, g, qi, gi); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res[name] = round(sorted(ts[2:])[len(ts[2:]) // 2], 4)
    res[name + "_checksum"] = int(i.sum())
g5 = torch.randn((125000, 512), generator=gen, device=dev); g5 /= g5.norm(dim=1, keepdim=True)
q5 = torch.randn((25000, 512), generator=gen, device=dev); q5 /= q5.norm(dim=1, keepdim=True)
gb, gi = ops.rows_prepare(g5); qb, qi = ops.rows_prepare(q5)
for nm, ex in (("c5q_exact", True), ("c5q_bf16", False)):
    ts = []
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s, i = ops.sim_topk(qb, gb, 5, 0, q5 if ex else None, g5 if ex else None, qi if ex else None, gi if ex else None)
        e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    res[nm] = round(sorted(ts[1:])[len(ts[1:]) // 2], 4)
print(json.dumps({"so": os.path.basename(os.environ.get("CVPCE_B200_SO") or "current"), "tune": tune, **res}))
