#!/usr/bin/env python
"""torchrun --nproc-per-node N tools/c5_exchange_check.py : the gallery-sharded search with the kernel-side peer scatter
(exchange="p2p") must return exactly what the NCCL all-gather formulation returns, step after step (double buffering)."""
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import dist as cdist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
G, D, Q, k = 200_000, 512, 20_000, 5
a = cdist.ShardedGallery.synthetic(G, D, world, rank, dev, seed=5, exchange="nccl")
b = cdist.ShardedGallery.synthetic(G, D, world, rank, dev, seed=5, exchange="p2p")
ok = True
for step in range(5):
    q = cdist.synthetic_queries(Q, D, dev, seed=6 + step)
    sa, ia = a.search(q, k)
    sb, ib = b.search(q, k)
    torch.cuda.synchronize()
    same = bool(torch.equal(sa, sb) and torch.equal(ia, ib))
    ok = ok and same
    if rank == 0:
        print("step %d: p2p == nccl: %s (exchange in use: %s)" % (step, same, b.exchange), flush=True)
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("ALL RANKS AGREE" if int(flag) == 1 else "MISMATCH", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag) == 1 else 1)
