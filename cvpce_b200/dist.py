"""One process per GPU (torchrun / mp.spawn, NCCL over NVLink): the two ways the hot path shards.

* image data-parallel (BASELINE configs 3-4): shelf images are independent units; rank r takes a contiguous block, the
  gallery is replicated, there is NO collective in the data path; results are gathered once at the end.
* gallery-sharded search (BASELINE config 5): rank r holds rows [lo_r, hi_r) of the gallery, every rank scores ALL queries
  against its shard with the fused tcgen05 similarity+top-k kernel (idx_base = lo_r), the per-rank (score, index)[Q,k]
  blocks are exchanged with one all-gather (W*Q*k*12 bytes: latency-, not bandwidth-bound) and merged by the
  topk_merge kernel with the same (score desc, index asc) order.

torch.distributed is plumbing only (process group + the NCCL all-gather); the arithmetic is in libcvpce_b200.so.
"""
import torch
import torch.distributed as dist

from . import ops

GEN_CHUNK = 65536  # synthetic galleries are generated in fixed chunks so that their content does not depend on world size


def shard_range(n, world, rank):
    """Contiguous block partition of n units: the first n % world ranks get one extra."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def partition_images(n_images, world, rank):
    lo, hi = shard_range(n_images, world, rank)
    return list(range(lo, hi))


def synthetic_rows(lo, hi, D, device, seed):
    """Rows [lo, hi) of the seeded synthetic unit-norm matrix N(0,1) (chunk c uses seed*1000003 + c)."""
    parts = []
    c = lo // GEN_CHUNK
    while c * GEN_CHUNK < hi:
        gen = torch.Generator(device=device).manual_seed(seed * 1000003 + c)
        chunk = torch.randn((GEN_CHUNK, D), generator=gen, device=device)
        a, b = max(lo, c * GEN_CHUNK) - c * GEN_CHUNK, min(hi, (c + 1) * GEN_CHUNK) - c * GEN_CHUNK
        parts.append(chunk[a:b])
        c += 1
    x = torch.cat(parts) if parts else torch.empty((0, D), device=device)
    return x / x.norm(dim=1, keepdim=True).clamp_min(1e-12)


def synthetic_queries(Q, D, device, seed, gallery_rows=None, planted_frac=0.1, noise=0.05):
    q = synthetic_rows(0, Q, D, device, seed)
    if gallery_rows is not None and len(gallery_rows):
        n = min(int(Q * planted_frac), len(gallery_rows))
        gen = torch.Generator(device=device).manual_seed(seed + 17)
        q[:n] = gallery_rows[:n] + noise * torch.randn((n, D), generator=gen, device=device)
        q = q / q.norm(dim=1, keepdim=True)
    return q.contiguous()


def all_gather_topk(scores, idx, group=None):
    """[Q,k] per rank -> ([W,Q,k], [W,Q,k]) on every rank."""
    world = dist.get_world_size(group)
    s_all = torch.empty((world,) + tuple(scores.shape), dtype=scores.dtype, device=scores.device)
    i_all = torch.empty((world,) + tuple(idx.shape), dtype=idx.dtype, device=idx.device)
    # views of one contiguous [W,...] buffer: NCCL gathers in place, gloo (CPU tests) copies
    dist.all_gather(list(s_all.unbind(0)), scores.contiguous(), group=group)
    dist.all_gather(list(i_all.unbind(0)), idx.contiguous(), group=group)
    return s_all, i_all


class ShardedGallery:
    def __init__(self, shard_f32, idx_base, world=1, rank=0, group=None, merge=None):
        self.world, self.rank, self.group = world, rank, group
        self.idx_base = int(idx_base)
        self.g32 = shard_f32.contiguous().float()
        self.gb, self.ginv = ops.rows_prepare(self.g32, normalize=True)
        self.merge = merge or ops.topk_merge
        self.launches = 1
        self._ws = None

    @classmethod
    def synthetic(cls, G, D, world, rank, device, seed=5, group=None):
        lo, hi = shard_range(G, world, rank)
        return cls(synthetic_rows(lo, hi, D, device, seed), lo, world, rank, group)

    def search(self, queries_f32, k, exact=True):
        """queries f32[Q,D] (present on every rank) -> (similarity f32[Q,k], GLOBAL gallery index i64[Q,k]) on every rank."""
        q = queries_f32.contiguous()
        Q, D = q.shape
        need = ops.sim_topk_workspace_bytes(Q, self.g32.shape[0], D, k)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 16), dtype=torch.uint8, device=q.device)
        qb, qinv = ops.rows_prepare(q, normalize=True)
        s, i = ops.sim_topk(qb, self.gb, k, self.idx_base, q if exact else None, self.g32 if exact else None,
                            qinv if exact else None, self.ginv if exact else None, workspace=self._ws)
        self.launches += 3
        if self.world == 1:
            return s, i
        s_all, i_all = all_gather_topk(s, i, self.group)
        self.launches += 1
        return self.merge(s_all, i_all)
