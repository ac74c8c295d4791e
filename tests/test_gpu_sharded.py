"""Gallery-sharded search (SURVEY §8 e, BASELINE config 5) on the GPU: the fused prepare+replicate kernel, the partitioned /
host-fed forms on one GPU, and -- when the box has at least two GPUs -- the multi-rank bit-identity of the three exchange
forms (peer scatter == NCCL all-gather == one rank searching the whole gallery), step after step (buffer flips)."""
import os

import numpy as np
import pytest
import torch

from cvpce_b200 import dist as cdist

pytestmark = pytest.mark.gpu


def test_rows_prepare_scatter_writes_every_target(cuda):
    from cvpce_b200 import ops
    torch.manual_seed(3)
    R, D, Q, lo = 77, 192, 200, 50
    x = torch.randn((R, D), device=cuda) * torch.rand((R, 1), device=cuda) * 5
    x[5] = 0                                                   # zero row: inverse norm clamps at 1/eps, bf16 row is zero
    want_b, want_inv = ops.rows_prepare(x, normalize=True)
    bufs = [(torch.full((Q, D), 7.0, device=cuda), torch.full((Q, D), 7.0, device=cuda, dtype=torch.bfloat16),
             torch.full((Q,), 7.0, device=cuda)) for _ in range(3)]
    ops.rows_prepare_scatter(x, [b[0][lo:].data_ptr() for b in bufs], [b[1][lo:].data_ptr() for b in bufs],
                             [b[2][lo:].data_ptr() for b in bufs])
    torch.cuda.synchronize()
    for f32, b16, inv in bufs:
        assert torch.equal(f32[lo:lo + R], x) and torch.equal(b16[lo:lo + R], want_b) and torch.equal(inv[lo:lo + R], want_inv)
        for t in (f32, b16, inv):                               # rows outside the block are untouched
            assert bool((t[:lo] == 7).all()) and bool((t[lo + R:] == 7).all())
    with pytest.raises(RuntimeError):
        ops.rows_prepare_scatter(torch.randn((4, 6), device=cuda), [bufs[0][0].data_ptr()], [bufs[0][1].data_ptr()], [bufs[0][2].data_ptr()])


def test_search_host_and_partitioned_forms_on_one_gpu(cuda):
    G, D, Q, k = 20000, 512, 1500, 5
    sg = cdist.ShardedGallery.synthetic(G, D, 1, 0, cuda, seed=5)
    q = cdist.synthetic_queries(Q, D, cuda, 6, sg.g32)
    s, i = sg.search(q, k)
    s2, i2 = sg.search_partitioned(q, 0, Q, k)
    assert torch.equal(s, s2) and torch.equal(i, i2)
    assert bool((i[: Q // 10, 0] == torch.arange(Q // 10, device=cuda)).all())           # planted neighbours at rank 1
    for chunks in (1, 3):
        parts = cdist.partition_query_chunks(Q, 1, 0, chunks)
        blocks = [(q[g0:g0 + n].cpu().pin_memory(), lo, qc) for g0, n, lo, qc in parts]
        res = sg.search_host(blocks, k)
        assert torch.equal(torch.cat([r[0] for r in res]), s.cpu()) and torch.equal(torch.cat([r[1] for r in res]), i.cpu())
    # k clamps to the gallery size; unsupported D / k take the exact fp32 kernel per shard
    small = cdist.ShardedGallery(sg.g32[:3], 0)
    s3, i3 = small.search(q[:10], 5)
    assert tuple(i3.shape) == (10, 3) and bool((i3 >= 0).all())
    odd = cdist.ShardedGallery(sg.g32[:1000, :500].contiguous(), 0)
    s4, i4 = odd.search(q[:64, :500].contiguous(), 30)
    from cvpce_b200 import ops
    s5, i5 = ops.sim_topk_f32(q[:64, :500].contiguous(), sg.g32[:1000, :500].contiguous(), 30, normalize=True)
    assert torch.equal(i4, i5) and torch.equal(s4, s5)


def _rank_main(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    G, D, Q, k = 50000, 512, 3001, 5
    a = cdist.ShardedGallery.synthetic(G, D, world, rank, dev, seed=5, exchange="nccl")
    b = cdist.ShardedGallery.synthetic(G, D, world, rank, dev, seed=5, exchange="p2p")
    whole = cdist.ShardedGallery.synthetic(G, D, 1, 0, dev, seed=5)
    ok, notes = True, []
    for step in range(4):                                       # several steps: the gather buffers flip, the query buffers are reused
        q = cdist.synthetic_queries(Q, D, dev, 6 + step, whole.g32)
        sw, iw = whole.search(q, k)
        sa, ia = a.search(q, k)
        sb, ib = b.search(q, k)
        lo, hi = cdist.shard_range(Q, world, rank)
        sc, ic = b.search_partitioned(q[lo:hi], lo, Q, k)
        sd, id_ = a.search_partitioned(q[lo:hi], lo, Q, k)
        parts = cdist.partition_query_chunks(Q, world, rank, 2)
        res = b.search_host([(q[g0:g0 + n].cpu().pin_memory(), l, qc) for g0, n, l, qc in parts], k)
        torch.cuda.synchronize()
        for name, s_, i_ in (("nccl", sa, ia), ("p2p", sb, ib), ("p2p partitioned", sc, ic), ("nccl partitioned", sd, id_)):
            same = bool(torch.equal(s_, sw) and torch.equal(i_, iw))
            ok = ok and same
            if not same:
                notes.append("step %d: %s differs from the single-rank search" % (step, name))
        for (sh, ih), (g0, n, _, _) in zip(res, parts):
            same = bool(torch.equal(sh, sw[g0:g0 + n].cpu()) and torch.equal(ih, iw[g0:g0 + n].cpu()))
            ok = ok and same
            if not same:
                notes.append("step %d: search_host rows %d.. differ" % (step, g0))
    notes.append("exchange in use: %s" % b.exchange)
    with open(os.path.join(tmp, "rank%d" % rank), "w") as f:
        f.write(("OK " if ok else "FAIL ") + "; ".join(notes))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_search_bit_identical_across_exchange_forms_two_gpus(cuda, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on the box (bench.py --gpus N repeats this check in-run: other_configs.c5.c5_parity)")
    import torch.multiprocessing as mp
    port = 33500 + os.getpid() % 2000
    mp.spawn(_rank_main, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = [open(tmp_path / ("rank%d" % r)).read() for r in range(2)]
    assert all(g.startswith("OK") for g in got), got
