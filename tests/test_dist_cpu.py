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


class _OracleBackend:
    """Stands in for cvpce_b200.ops inside ShardedGallery (tests only): the oracle's exact top-k on the host."""

    @staticmethod
    def rows_prepare(x, normalize=True):
        n = x.norm(dim=1).clamp_min(1e-8)
        return (x / n[:, None]).to(torch.bfloat16), 1.0 / n

    @staticmethod
    def sim_topk_workspace_bytes(Q, G, D, k):
        return 16

    @staticmethod
    def _topk(q32, g32, k, base):
        import oracle
        kk = min(k, len(g32))
        sim, idx = oracle.sim_topk(q32.numpy(), g32.numpy(), kk)
        s = np.full((len(q32), k), -np.inf, np.float32)
        i = np.full((len(q32), k), -1, np.int64)
        s[:, :kk], i[:, :kk] = sim, idx + base
        return torch.from_numpy(s), torch.from_numpy(i)

    def sim_topk(self, qb, gb, k, base, q32, g32, qinv, ginv, workspace=None):
        return self._topk(q32, g32, k, base)

    def sim_topk_f32(self, q, g, k, base, normalize=True):
        return self._topk(q, g, k, base)

    @staticmethod
    def topk_merge(s_all, i_all):
        return _merge_numpy(s_all, i_all)


def _worker_sharded(rank, world, port, tmp):
    """The real ShardedGallery host logic (k clamp, unsupported-D fallback, padded all-gather of partitioned queries, ranks
    without rows) on gloo, with the oracle standing in for the kernels."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    cpu = torch.device("cpu")
    ok = True
    for G, D, Q, k in ((1000, 64, 41, 5), (1000, 48, 40, 5), (3, 64, 7, 5), (1, 64, 5, 2)):   # D=48: fp32 fallback; G < k; G < world
        full = cdist.synthetic_rows(0, G, D, cpu, 5)
        q = cdist.synthetic_queries(Q, D, cpu, 6, full)
        lo, hi = cdist.shard_range(G, world, rank)
        sg = cdist.ShardedGallery(full[lo:hi], lo, world, rank, backend=_OracleBackend())
        ok = ok and sg.global_rows == G
        wsim, widx = oracle.sim_topk(q.numpy(), full.numpy(), min(k, G))
        s, i = sg.search(q, k)
        ok = ok and tuple(i.shape) == (Q, min(k, G)) and np.array_equal(i.numpy(), widx) and np.allclose(s.numpy(), wsim, atol=1e-6)
        qlo, qhi = cdist.shard_range(Q, world, rank)
        s2, i2 = sg.search_partitioned(q[qlo:qhi], qlo, Q, k)
        ok = ok and torch.equal(i2, i) and torch.equal(s2, s)
    parts = cdist.partition_query_chunks(10, world, rank, 3)
    got = [None] * world
    dist.all_gather_object(got, [(g0, n) for g0, n, _, _ in parts])
    rows = sorted(r for per in got for g0, n in per for r in range(g0, g0 + n))
    ok = ok and rows == list(range(10))
    with open(os.path.join(tmp, "sharded%d" % rank), "w") as f:
        f.write("1" if ok else "0")
    dist.destroy_process_group()


def test_sharded_gallery_host_logic_world2_gloo(tmp_path):
    port = 31500 + os.getpid() % 2000
    mp.spawn(_worker_sharded, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / ("sharded%d" % r)).read() for r in range(2)] == ["1", "1"]


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
            rows, lo, a, total = cdist.ShardedGallery.load_index_rows(pth, world, rank)
            assert lo == nxt and a == ann and total == 101 and rows.dtype == torch.float32 and rows.is_contiguous()
            nxt += rows.shape[0]
            parts.append(rows)
        assert torch.equal(torch.cat(parts), emb)
