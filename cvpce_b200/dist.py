"""One process per GPU (torchrun / mp.spawn, NCCL over NVLink): the two ways the hot path shards.

* image data-parallel (BASELINE configs 3-4): shelf images are independent units; rank r takes a contiguous block, the
  gallery is replicated, there is NO collective in the data path; results are gathered once at the end.
* gallery-sharded search (BASELINE config 5): rank r holds rows [lo_r, hi_r) of the gallery, every rank scores ALL queries
  against its shard with the fused tcgen05 similarity+top-k kernel (idx_base = lo_r), the per-rank (score, index)[Q,k]
  blocks are exchanged with one all-gather (W*Q*k*12 bytes: latency-, not bandwidth-bound) and merged by the
  topk_merge kernel with the same (score desc, index asc) order.

  The exchange has two forms with bit-identical results: exchange="nccl" (the all-gather of the reference formulation) and
  exchange="p2p": the selection kernel stores each finished row into every rank's gather buffer over NVLink peer memory
  (torch symmetric memory), so no collective launch follows the kernel.
* queries partitioned as well (the end-to-end form of config 5): each rank receives only 1/W of the queries from its host;
  cvpce_rows_prepare_scatter prepares the block and stores fp32 rows, bf16 rows and inverse norms into every rank's query
  buffers (again peer stores instead of a collective); each rank downloads only its block of the merged result.

torch.distributed is plumbing only (process group, symmetric-memory rendezvous / barrier, the NCCL all-gather of the
fallback form); the arithmetic and the peer stores are in libcvpce_b200.so.
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


class PeerExchange:
    """Symmetric-memory buffers of the gallery-sharded search (torch.distributed._symmetric_memory: every rank's buffer is
    mapped into every other rank's address space over NVLink): double-buffered [W,Q,k] score / index gather buffers and,
    when the queries arrive partitioned, the [Q,D] fp32 / [Q,D] bf16 / [Q] inverse-norm query buffers.  Plumbing only: it
    hands raw peer addresses to cvpce_sim_topk_scatter / cvpce_rows_prepare_scatter and provides the cross-rank barrier;
    the bytes are moved by the kernels' stores."""

    def __init__(self, world, rank, Q, k, device, group=None, D=0):
        import torch.distributed._symmetric_memory as symm_mem
        self.world, self.rank, self.Q, self.k, self.D = world, rank, Q, k, D
        self.s_bytes = world * Q * k * 4
        self.i_bytes = world * Q * k * 8
        self.half = (self.s_bytes + self.i_bytes + 255) // 256 * 256
        al = lambda n: (n + 255) // 256 * 256
        self.q32_off = 2 * self.half
        self.qb_off = self.q32_off + al(Q * D * 4)
        self.qi_off = self.qb_off + al(Q * D * 2)
        total = self.qi_off + al(Q * 4 if D else 0)
        pg = group if group is not None else dist.group.WORLD
        try:  # needed by older torch builds, deprecated no-op in newer ones
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                if not symm_mem.is_symm_mem_enabled_for_group(pg.group_name):
                    symm_mem.enable_symm_mem_for_group(pg.group_name)
        except Exception:
            pass
        self.buf = symm_mem.empty((total,), dtype=torch.uint8, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, pg)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.flip = 0
        self.hdl.barrier()

    def fits(self, Q, k, D):
        return Q <= self.Q and k == self.k and (D == 0 or D == self.D)

    # ---- per-rank result blocks
    def targets(self, Q):
        """(score slot addresses, index slot addresses) of THIS rank's [Q,k] block in every rank's current half."""
        base = self.flip * self.half
        slot_s = base + self.rank * Q * self.k * 4
        slot_i = base + self.world * Q * self.k * 4 + self.rank * Q * self.k * 8
        return [p + slot_s for p in self.ptrs], [p + slot_i for p in self.ptrs]

    def gathered(self, Q):
        base = self.flip * self.half
        nb_s, nb_i = self.world * Q * self.k * 4, self.world * Q * self.k * 8
        s = self.buf[base: base + nb_s].view(torch.float32).view(self.world, Q, self.k)
        i = self.buf[base + nb_s: base + nb_s + nb_i].view(torch.int64).view(self.world, Q, self.k)
        return s, i

    # ---- replicated query block
    def query_targets(self, lo):
        """Addresses of row `lo` of the fp32 / bf16 / inverse-norm query buffers on every rank."""
        D = self.D
        return ([p + self.q32_off + lo * D * 4 for p in self.ptrs], [p + self.qb_off + lo * D * 2 for p in self.ptrs],
                [p + self.qi_off + lo * 4 for p in self.ptrs])

    def queries(self, Q):
        D = self.D
        q32 = self.buf[self.q32_off: self.q32_off + Q * D * 4].view(torch.float32).view(Q, D)
        qb = self.buf[self.qb_off: self.qb_off + Q * D * 2].view(torch.bfloat16).view(Q, D)
        qi = self.buf[self.qi_off: self.qi_off + Q * 4].view(torch.float32)
        return q32, qb, qi

    def barrier(self):
        self.hdl.barrier()

    def barrier_and_flip(self):
        # every rank's stores of this step are complete (kernel boundary) before its barrier signal; after the barrier all
        # W blocks of the current half are in place.  The other half is only rewritten after the NEXT barrier, by which time
        # every rank's merge of this half has been issued before that barrier on its stream.
        self.hdl.barrier()
        self.flip ^= 1


PeerGather = PeerExchange  # round-1 name

class ShardedGallery:
    """This rank's shard of the gallery + the search over all shards.

    Contract of the multi-rank calls: every rank of the group calls the same method with the same Q, k (and, for search(),
    the SAME queries); the image data-parallel mode, where ranks hold different queries, does not use this class."""

    def __init__(self, shard_f32, idx_base, world=1, rank=0, group=None, merge=None, exchange="nccl", global_rows=None, backend=None):
        """exchange: "nccl" = all-gather of the per-rank [Q,k] blocks after the search (the reference formulation, SURVEY §8 e);
        "p2p" = the selection kernel stores its block straight into every rank's gather buffer over NVLink peer memory.
        global_rows: rows of the whole gallery (default: summed over the group once); backend: object providing rows_prepare /
        sim_topk / sim_topk_f32 / topk_merge (default: cvpce_b200.ops; the gloo CPU tests substitute the oracle)."""
        self.ops = backend or ops
        self.exchange = exchange
        self._peer = None
        self.world, self.rank, self.group = world, rank, group
        self.idx_base = int(idx_base)
        self.g32 = shard_f32.contiguous().float()
        self.D = int(self.g32.shape[1])
        self.tc_ok = self.D % 64 == 0 and 64 <= self.D <= 4096  # else: exact fp32 CUDA-core search per shard
        if self.tc_ok and self.g32.shape[0]:
            self.gb, self.ginv = self.ops.rows_prepare(self.g32, normalize=True)
        else:  # fp32 path, or a rank without rows (only when the gallery has fewer rows than ranks)
            self.gb = torch.empty((0, self.D), dtype=torch.bfloat16, device=self.g32.device)
            self.ginv = torch.empty((0,), dtype=torch.float32, device=self.g32.device)
        self.merge = merge or self.ops.topk_merge
        if global_rows is None:
            global_rows = int(self.g32.shape[0])
            if world > 1:
                t = torch.tensor([global_rows], dtype=torch.int64, device=self.g32.device)
                dist.all_reduce(t, group=group)
                global_rows = int(t.item())
        self.global_rows = int(global_rows)
        self.launches = 1
        self._ws = None

    @classmethod
    def synthetic(cls, G, D, world, rank, device, seed=5, group=None, exchange="nccl"):
        lo, hi = shard_range(G, world, rank)
        return cls(synthetic_rows(lo, hi, D, device, seed), lo, world, rank, group, exchange=exchange, global_rows=G)

    @staticmethod
    def load_index_rows(pth, world, rank):
        """This rank's row range of an index file in the reference's format ({'embedding', 'annotations'}, written by
        Classifier.save_index, cvpce/production.py:52-56): the file is memory-mapped, so a rank pages in only its own rows.
        Returns (rows f32 [hi-lo, D] on the host, first global row index, annotations list (whole: it maps global indices),
        rows of the whole gallery)."""
        try:
            idx = torch.load(pth, map_location="cpu", mmap=True, weights_only=False)
        except (RuntimeError, ValueError, TypeError):  # legacy (non-zip) serialisation cannot be mapped
            idx = torch.load(pth, map_location="cpu", weights_only=False)
        emb = idx["embedding"]
        lo, hi = shard_range(int(emb.shape[0]), world, rank)
        return emb[lo:hi].contiguous().float(), lo, idx["annotations"], int(emb.shape[0])

    @classmethod
    def from_index(cls, pth, world, rank, device, group=None, exchange="nccl"):
        """(ShardedGallery over this rank's rows of the index file, annotations)."""
        rows, lo, annotations, total = cls.load_index_rows(pth, world, rank)
        return cls(rows.to(device), lo, world, rank, group, exchange=exchange, global_rows=total), annotations

    # ------------------------------------------------------------------------------------------------------------ helpers
    def _k_eff(self, k):
        return max(0, min(int(k), self.global_rows))

    def _use_tc(self, k, exact):
        # the tcgen05 kernel keeps k' >= k + 8 candidates per row for the exact re-score (k <= 24), k <= 32 without it
        return self.tc_ok and k <= (24 if exact else 32)

    def _workspace(self, Q, k, device):
        need = self.ops.sim_topk_workspace_bytes(Q, max(int(self.g32.shape[0]), 1), self.D, k)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(max(need, 16), dtype=torch.uint8, device=device)
        return self._ws

    def _peer_for(self, Q, k, device, D=0):
        if self._peer is None or not self._peer.fits(Q, k, D):
            grow = max(Q, self._peer.Q) if self._peer is not None and self._peer.k == k else Q
            self._peer = PeerExchange(self.world, self.rank, grow, k, device, self.group, D=D or (self._peer.D if self._peer else 0))
        return self._peer

    def _p2p_failed(self, e):
        import sys
        sys.stderr.write("cvpce_b200.dist: symmetric memory unavailable (%r); using the NCCL all-gather\n" % (e,))
        self.exchange = "nccl"

    def _local_topk(self, q32, qb, qinv, k, exact):
        """This shard's (scores f32 [Q,k], GLOBAL indices i64 [Q,k]) for prepared queries (qb / qinv None: fp32 path)."""
        Q = q32.shape[0]
        if self.g32.shape[0] == 0:  # a rank without gallery rows contributes padding
            return (torch.full((Q, k), float("-inf"), dtype=torch.float32, device=q32.device),
                    torch.full((Q, k), -1, dtype=torch.int64, device=q32.device))
        if qb is None:
            self.launches += 1
            kk = min(k, int(self.g32.shape[0]))
            res = self.ops.sim_topk_f32(q32, self.g32, kk, self.idx_base, normalize=True)
            return res if kk == k else self._pad(res, k)
        self.launches += 2
        return self.ops.sim_topk(qb, self.gb, k, self.idx_base, q32 if exact else None, self.g32 if exact else None,
                                 qinv if exact else None, self.ginv if exact else None, workspace=self._workspace(Q, k, q32.device))

    @staticmethod
    def _pad(si, k):
        s, i = si
        n = k - s.shape[1]
        return (torch.cat([s, torch.full((s.shape[0], n), float("-inf"), dtype=s.dtype, device=s.device)], 1),
                torch.cat([i, torch.full((i.shape[0], n), -1, dtype=i.dtype, device=i.device)], 1))

    # ------------------------------------------------------------------------------------------------------------- search
    def search(self, queries_f32, k, exact=True):
        """queries f32[Q,D], the SAME on every rank -> (similarity f32[Q,k'], GLOBAL gallery index i64[Q,k']) on every rank,
        k' = min(k, rows of the whole gallery)."""
        q = queries_f32.contiguous().float()
        Q = q.shape[0]
        k = self._k_eff(k)
        if Q == 0 or k == 0:
            return (torch.empty((Q, k), dtype=torch.float32, device=q.device), torch.empty((Q, k), dtype=torch.int64, device=q.device))
        tc = self._use_tc(k, exact)
        qb = qinv = None
        if tc:
            qb, qinv = self.ops.rows_prepare(q, normalize=True)
            self.launches += 1
        if self.world > 1 and self.exchange == "p2p" and tc:
            exact = exact and self.g32.shape[0] > 0  # a rank without rows has no candidate to re-score
            try:
                peer = self._peer_for(Q, k, q.device)
            except Exception as e:  # no peer mapping on this box (plumbing, not compute): use the NCCL exchange instead
                self._p2p_failed(e)
                return self.search(queries_f32, k, exact)
            ps, pi = peer.targets(Q)
            self.ops.sim_topk_scatter(qb, self.gb, k, self.idx_base, ps, pi, q if exact else None, self.g32 if exact else None,
                                      qinv if exact else None, self.ginv if exact else None, workspace=self._workspace(Q, k, q.device))
            s_all, i_all = peer.gathered(Q)
            peer.barrier_and_flip()
            self.launches += 3
            return self.merge(s_all, i_all)
        s, i = self._local_topk(q, qb, qinv, k, exact)
        if self.world == 1:
            return s, i
        s_all, i_all = all_gather_topk(s, i, self.group)
        self.launches += 1
        return self.merge(s_all, i_all)

    def search_partitioned(self, q_block, lo, Q, k, exact=True):
        """The queries arrive PARTITIONED: this rank holds rows [lo, lo + len(q_block)) of the Q queries (blocks in rank order,
        shard_range(Q, world, rank)).  The block is prepared and replicated to every rank (p2p: cvpce_rows_prepare_scatter
        stores into every rank's query buffers; nccl: all-gather of the blocks), then searched as in search().
        Returns the merged (similarity [Q,k'], index [Q,k']) of ALL queries on every rank."""
        q_block = q_block.contiguous().float()
        dev = q_block.device
        k = self._k_eff(k)
        if Q == 0 or k == 0:
            return (torch.empty((Q, k), dtype=torch.float32, device=dev), torch.empty((Q, k), dtype=torch.int64, device=dev))
        if self.world == 1:
            return self.search(q_block, k, exact)
        tc = self._use_tc(k, exact)
        if self.exchange == "p2p" and tc:
            exact = exact and self.g32.shape[0] > 0
            try:
                peer = self._peer_for(Q, k, dev, D=self.D)
            except Exception as e:
                self._p2p_failed(e)
                return self.search_partitioned(q_block, lo, Q, k, exact)
            p32, pb, pinv = peer.query_targets(lo)
            self.ops.rows_prepare_scatter(q_block, p32, pb, pinv, normalize=True)
            peer.barrier()  # every rank's block is in place in every rank's query buffers
            q32, qb, qinv = peer.queries(Q)
            ps, pi = peer.targets(Q)
            self.ops.sim_topk_scatter(qb, self.gb, k, self.idx_base, ps, pi, q32 if exact else None, self.g32 if exact else None,
                                      qinv if exact else None, self.ginv if exact else None, workspace=self._workspace(Q, k, dev))
            s_all, i_all = peer.gathered(Q)
            peer.barrier_and_flip()
            self.launches += 4  # prepare+replicate, tcgen05 search, select+scatter, merge (the barriers are torch's)
            return self.merge(s_all, i_all)
        # NCCL form: all-gather of the (padded) blocks, then the resident-queries search
        rows = (Q + self.world - 1) // self.world
        mine = torch.zeros((rows, self.D), dtype=torch.float32, device=dev)
        mine[: q_block.shape[0]] = q_block
        parts = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(parts, mine, group=self.group)
        self.launches += 1
        q = torch.cat([parts[r][: shard_range(Q, self.world, r)[1] - shard_range(Q, self.world, r)[0]] for r in range(self.world)])
        return self.search(q, k, exact)

    def search_host(self, blocks_h, k, exact=True, out=None):
        """End-to-end form: `blocks_h` = list of (q_block PINNED host f32 [n_c, D], lo_c, Q_c), one entry per query chunk c: this
        rank's block of chunk c (rows [lo_c, lo_c + n_c) of the chunk's Q_c queries).  Chunk c+1 is uploaded on a copy stream
        while chunk c is searched.  Returns a list of (similarity, index) PINNED host tensors [n_c, k'] for the same rows (each
        rank downloads only its own block of the merged result); synchronises once at the end."""
        dev = self.g32.device
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream(dev)
        staged = []
        with torch.cuda.stream(self._copy_stream):
            for qh, lo, Qc in blocks_h:
                qd = qh.to(dev, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
                staged.append((qd, ev))
        res = []
        kk = self._k_eff(k)
        for c, ((qh, lo, Qc), (qd, ev)) in enumerate(zip(blocks_h, staged)):
            main.wait_event(ev)
            qd.record_stream(main)
            s, i = self.search_partitioned(qd, lo, Qc, k, exact)
            n = qh.shape[0]
            if out is not None:
                sh, ih = out[c]
            else:
                sh = torch.empty((n, kk), dtype=torch.float32).pin_memory()
                ih = torch.empty((n, kk), dtype=torch.int64).pin_memory()
            rows = slice(lo, lo + n) if self.world > 1 else slice(0, n)
            sh.copy_(s[rows], non_blocking=True)
            ih.copy_(i[rows], non_blocking=True)
            res.append((sh, ih))
        main.synchronize()
        return res


def partition_query_chunks(Q, world, rank, chunks):
    """Row ranges for search_host: chunk c covers queries [c0, c1); inside it this rank owns rows shard_range(c1 - c0, world,
    rank).  Returns a list of (global first row, rows, row offset inside the chunk, chunk length)."""
    out = []
    for c in range(chunks):
        c0, c1 = shard_range(Q, chunks, c)
        lo, hi = shard_range(c1 - c0, world, rank)
        out.append((c0 + lo, hi - lo, lo, c1 - c0))
    return out
