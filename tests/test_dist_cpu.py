"""CPU, world_size 2, gloo: the host-side logic of the N>1 paths (partitioning, all-gather plumbing, merge order).
The CUDA arithmetic is replaced by the oracle here (tests only); the product path itself has no CPU mode."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cvpce_b200 import dist as cdist


def test_shard_range_partitions_everything():
    for n in (0, 1, 7, 512, 1_000_000):
        for world in (1, 2, 3, 8):
            got = [cdist.shard_range(n, world, r) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
            sizes = [b - a for a, b in got]
            assert max(sizes) - min(sizes) <= 1
    assert cdist.partition_images(5, 2, 0) == [0, 1, 2] and cdist.partition_images(5, 2, 1) == [3, 4]


def test_synthetic_rows_do_not_depend_on_sharding():
    dev = torch.device("cpu")
    full = cdist.synthetic_rows(0, 3 * cdist.GEN_CHUNK // 2, 16, dev, 5)
    lo, hi = cdist.shard_range(len(full), 3, 1)
    assert torch.equal(cdist.synthetic_rows(lo, hi, 16, dev, 5), full[lo:hi])
    assert torch.allclose(full.norm(dim=1), torch.ones(len(full)), atol=1e-5)


def _merge_numpy(s_all, i_all):
    W, Q, k = s_all.shape
    s = s_all.permute(1, 0, 2).reshape(Q, W * k).numpy()
    i = i_all.permute(1, 0, 2).reshape(Q, W * k).numpy()
    out_s, out_i = np.empty((Q, k), np.float32), np.empty((Q, k), np.int64)
    for r in range(Q):
        order = sorted(range(W * k), key=lambda c: (i[r, c] < 0, -s[r, c], i[r, c]))[:k]
        out_s[r], out_i[r] = s[r, order], i[r, order]
    return torch.from_numpy(out_s), torch.from_numpy(out_i)


def _worker(rank, world, port, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    G, D, Q, k = 1000, 32, 40, 5
    full = cdist.synthetic_rows(0, G, D, torch.device("cpu"), 5)
    q = cdist.synthetic_queries(Q, D, torch.device("cpu"), 6, full)
    lo, hi = cdist.shard_range(G, world, rank)
    sim, idx = oracle.sim_topk(q.numpy(), full[lo:hi].numpy(), k)          # stands in for the CUDA shard search
    s_all, i_all = cdist.all_gather_topk(torch.from_numpy(sim.astype(np.float32)), torch.from_numpy(idx + lo))
    ms, mi = _merge_numpy(s_all, i_all)
    wsim, widx = oracle.sim_topk(q.numpy(), full.numpy(), k)
    ok = np.array_equal(mi.numpy(), widx) and np.allclose(ms.numpy(), wsim, atol=1e-6)
    # image data-parallel: every rank classifies its block; results gathered once at the end
    mine = torch.tensor(cdist.partition_images(9, world, rank), dtype=torch.int64)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine.tolist())
    ok = ok and sorted(sum(gathered, [])) == list(range(9))
    with open(os.path.join(tmp, "ok%d" % rank), "w") as f:
        f.write("1" if ok else "0")
    dist.destroy_process_group()


def test_gallery_sharded_search_world2_gloo(tmp_path):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / ("ok%d" % r)).read() for r in range(2)] == ["1", "1"]


def test_index_file_rows_per_rank(tmp_path):
    """f2: every rank reads only its row range of an index file in the reference's {'embedding','annotations'} format
    (production.py:52-56); the ranges tile the gallery and the annotation list stays whole."""
    emb = cdist.synthetic_rows(0, 101, 16, torch.device("cpu"), 9)
    ann = ["sku%03d" % i for i in range(101)]
    pth = str(tmp_path / "index.pt")
    torch.save({"embedding": emb, "annotations": ann}, pth)
    for world in (1, 2, 3, 8):
        parts, nxt = [], 0
        for rank in range(world):
            rows, lo, a = cdist.ShardedGallery.load_index_rows(pth, world, rank)
            assert lo == nxt and a == ann and rows.dtype == torch.float32 and rows.is_contiguous()
            nxt += rows.shape[0]
            parts.append(rows)
        assert torch.equal(torch.cat(parts), emb)
