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


class PeerGather:
    """Double-buffered [W,Q,k] score / index gather buffers in symmetric memory (torch.distributed._symmetric_memory: every
    rank's buffer is mapped into every other rank's address space over NVLink).  Plumbing only: it hands raw peer
    addresses to cvpce_sim_topk_scatter and provides the cross-rank barrier; the bytes are moved by the kernel's stores."""

    def __init__(self, world, rank, Q, k, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.world, self.rank, self.Q, self.k = world, rank, Q, k
        self.s_bytes = world * Q * k * 4
        self.i_bytes = world * Q * k * 8
        self.half = (self.s_bytes + self.i_bytes + 255) // 256 * 256
        pg = group if group is not None else dist.group.WORLD
        try:  # needed by older torch builds, deprecated no-op in newer ones
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                if not symm_mem.is_symm_mem_enabled_for_group(pg.group_name):
                    symm_mem.enable_symm_mem_for_group(pg.group_name)
        except Exception:
            pass
        self.buf = symm_mem.empty((2 * self.half,), dtype=torch.uint8, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, pg)
        self.flip = 0
        self.hdl.barrier()

    def targets(self):
        """(score slot addresses, index slot addresses) of THIS rank's [Q,k] block in every rank's current half."""
        base = self.flip * self.half
        slot_s = base + self.rank * self.Q * self.k * 4
        slot_i = base + self.s_bytes + self.rank * self.Q * self.k * 8
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        return [p + slot_s for p in ptrs], [p + slot_i for p in ptrs]

    def gathered(self):
        base = self.flip * self.half
        s = self.buf[base: base + self.s_bytes].view(torch.float32).view(self.world, self.Q, self.k)
        i = self.buf[base + self.s_bytes: base + self.s_bytes + self.i_bytes].view(torch.int64).view(self.world, self.Q, self.k)
        return s, i

    def barrier_and_flip(self):
        # every rank's stores of this step are complete (kernel boundary) before its barrier signal; after the barrier all
        # W blocks of the current half are in place.  The other half is only rewritten after the NEXT barrier, by which time
        # every rank's merge of this half has been issued before that barrier on its stream.
        self.hdl.barrier()
        self.flip ^= 1


class ShardedGallery:
    def __init__(self, shard_f32, idx_base, world=1, rank=0, group=None, merge=None, exchange="nccl"):
        """exchange: "nccl" = all-gather of the per-rank [Q,k] blocks after the search (the reference formulation, SURVEY §8 e);
        "p2p" = the selection kernel stores its block straight into every rank's gather buffer over NVLink peer memory."""
        self.exchange = exchange
        self._peer = None
        self.world, self.rank, self.group = world, rank, group
        self.idx_base = int(idx_base)
        self.g32 = shard_f32.contiguous().float()
        self.gb, self.ginv = ops.rows_prepare(self.g32, normalize=True)
        self.merge = merge or ops.topk_merge
        self.launches = 1
        self._ws = None

    @classmethod
    def synthetic(cls, G, D, world, rank, device, seed=5, group=None, exchange="nccl"):
        lo, hi = shard_range(G, world, rank)
        return cls(synthetic_rows(lo, hi, D, device, seed), lo, world, rank, group, exchange=exchange)

    @staticmethod
    def load_index_rows(pth, world, rank):
        """This rank's row range of an index file in the reference's format ({'embedding', 'annotations'}, written by
        Classifier.save_index, cvpce/production.py:52-56): the file is memory-mapped, so a rank pages in only its own rows.
        Returns (rows f32 [hi-lo, D] on the host, first global row index, annotations list (whole: it maps global indices))."""
        try:
            idx = torch.load(pth, map_location="cpu", mmap=True, weights_only=False)
        except (RuntimeError, ValueError, TypeError):  # legacy (non-zip) serialisation cannot be mapped
            idx = torch.load(pth, map_location="cpu", weights_only=False)
        emb = idx["embedding"]
        lo, hi = shard_range(int(emb.shape[0]), world, rank)
        return emb[lo:hi].contiguous().float(), lo, idx["annotations"]

    @classmethod
    def from_index(cls, pth, world, rank, device, group=None, exchange="nccl"):
        """(ShardedGallery over this rank's rows of the index file, annotations)."""
        rows, lo, annotations = cls.load_index_rows(pth, world, rank)
        return cls(rows.to(device), lo, world, rank, group, exchange=exchange), annotations

    def search(self, queries_f32, k, exact=True):
        """queries f32[Q,D] (present on every rank) -> (similarity f32[Q,k], GLOBAL gallery index i64[Q,k]) on every rank."""
        q = queries_f32.contiguous()
        Q, D = q.shape
        need = ops.sim_topk_workspace_bytes(Q, self.g32.shape[0], D, k)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 16), dtype=torch.uint8, device=q.device)
        qb, qinv = ops.rows_prepare(q, normalize=True)
        if self.world > 1 and self.exchange == "p2p":
            if self._peer is None or self._peer.Q != Q or self._peer.k != k:
                try:
                    self._peer = PeerGather(self.world, self.rank, Q, k, q.device, self.group)
                except Exception as e:  # no peer mapping on this box (plumbing, not compute): use the NCCL exchange instead
                    import sys
                    sys.stderr.write("cvpce_b200.dist: symmetric memory unavailable (%r); using the NCCL all-gather\n" % (e,))
                    self.exchange = "nccl"
                    return self.search(queries_f32, k, exact)
            ps, pi = self._peer.targets()
            ops.sim_topk_scatter(qb, self.gb, k, self.idx_base, ps, pi, q if exact else None, self.g32 if exact else None,
                                 qinv if exact else None, self.ginv if exact else None, workspace=self._ws)
            s_all, i_all = self._peer.gathered()
            self._peer.barrier_and_flip()
            self.launches += 4
            return self.merge(s_all, i_all)
        s, i = ops.sim_topk(qb, self.gb, k, self.idx_base, q if exact else None, self.g32 if exact else None,
                            qinv if exact else None, self.ginv if exact else None, workspace=self._ws)
        self.launches += 3
        if self.world == 1:
            return s, i
        s_all, i_all = all_gather_topk(s, i, self.group)
        self.launches += 1
        return self.merge(s_all, i_all)
