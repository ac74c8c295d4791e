"""GPU: the kernels at BASELINE.json's FULL sizes, checked through size-independent properties (the oracle would take too long
there): NMS idempotence / maximality / order on 32 x 1000 and 512 x ~750 candidates, identity crops and constant images on
2048 x 1536 shelves, exact maxima on 256-crop trunk batches, planted neighbours / shard-merge invariance / scale invariance on
a C5 shard (125k x 512 gallery rows)."""
import numpy as np
import pytest
import torch

from cvpce_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(cuda):
    from cvpce_b200 import _lib, ops as _ops
    assert _lib.load().cvpce_device_check() == 0, _lib.last_error()
    return _ops


def _shelves(n, products, seed0):
    boxes, scores, seg = [], [], [0]
    for i in range(n):
        b, s = synth.shelf_candidates(seed0 + i, n_products=products)
        boxes.append(b)
        scores.append(s)
        seg.append(seg[-1] + len(s))
    return np.concatenate(boxes), np.concatenate(scores), np.array(seg, dtype=np.int32)


@pytest.mark.parametrize("n_images,products", [(32, 200), (512, 150)])   # configs[2] (C3) and configs[3] (C4)
def test_nms_properties_at_full_size(ops, n_images, products):
    boxes, scores, seg = _shelves(n_images, products, 3000)
    bd, sd, segd = (torch.from_numpy(x).cuda() for x in (boxes, scores, seg))
    max_n = int(np.diff(seg).max())
    keep, counts, kb, ks = ops.nms_batched(bd, sd, segd, max_n, 0.5, want_boxes=True, want_scores=True)
    counts_h = counts.cpu().numpy()
    assert (counts_h > 0).all() and (counts_h <= np.diff(seg)).all()
    thr = np.float32(0.5)
    kept_rows, kept_seg = [], [0]
    for b in range(n_images):
        lo, c = int(seg[b]), int(counts_h[b])
        loc = keep[lo:lo + c]
        assert int(loc.min()) >= 0 and int(loc.max()) < seg[b + 1] - seg[b] and loc.unique().numel() == c
        assert torch.equal(kb[lo:lo + c], bd[lo + loc]) and torch.equal(ks[lo:lo + c], sd[lo + loc])   # gathered outputs
        assert bool((ks[lo:lo + c][:-1] >= ks[lo:lo + c][1:]).all())                                    # score-descending
        kept_rows.append(kb[lo:lo + c])
        kept_seg.append(kept_seg[-1] + c)
        if b < 8:
            iou_kk = ops.box_iou(kb[lo:lo + c], kb[lo:lo + c])
            iou_kk.fill_diagonal_(0)
            assert float(iou_kk.max()) <= thr                                           # no kept pair overlaps beyond thr
            removed = torch.ones(int(seg[b + 1] - seg[b]), dtype=torch.bool, device="cuda")
            removed[loc] = False
            if removed.any():                                                           # maximality: every removed box is
                iou_rk = ops.box_iou(bd[lo:int(seg[b + 1])][removed], kb[lo:lo + c])    # suppressed by a kept box that
                higher = ks[lo:lo + c][None, :] >= sd[lo:int(seg[b + 1])][removed][:, None]   # does not score lower
                assert bool(((iou_rk > thr) & higher).any(dim=1).all())
    # idempotence: NMS of the kept boxes keeps all of them, in the same order
    kept_b = torch.cat(kept_rows)
    kept_s = torch.cat([ks[int(seg[b]):int(seg[b]) + int(counts_h[b])] for b in range(n_images)])
    kseg = torch.tensor(kept_seg, dtype=torch.int32, device="cuda")
    keep2, counts2, _, _ = ops.nms_batched(kept_b, kept_s, kseg, int(counts_h.max()), 0.5)
    assert torch.equal(counts2, counts)
    for b in range(0, n_images, max(1, n_images // 16)):
        c = int(counts_h[b])
        assert torch.equal(keep2[kept_seg[b]:kept_seg[b] + c], torch.arange(c, device="cuda"))


def test_crop_identity_and_constant_at_shelf_size(ops):
    H, W = synth.SHELF_HW
    S = 256
    g = torch.Generator(device="cuda").manual_seed(1)
    img = torch.rand((3, H, W), device="cuda", generator=g)
    const = torch.full((3, H, W), 0.25, device="cuda")
    table = ops.make_image_table([img, const])
    rs = np.random.RandomState(2)
    # S x S boxes at integer positions (incl. the image corners): scale 1 => every filter has one tap of weight 1 => exact copy
    xs = np.concatenate([[0, W - S], rs.randint(0, W - S, size=62)])
    ys = np.concatenate([[0, H - S], rs.randint(0, H - S, size=62)])
    boxes = np.stack([xs, ys, xs + S, ys + S], 1).astype(np.float32)
    boxes[5:10] += 0.7                                   # fractional coordinates truncate toward zero (production.py:20)
    crops = ops.crop_resize(table, 2, torch.from_numpy(boxes).cuda(), torch.zeros(64, dtype=torch.int32, device="cuda"), S)
    for i in range(64):
        x, y = int(boxes[i, 0]), int(boxes[i, 1])
        assert torch.equal(crops[i], img[:, y:y + S, x:x + S]), i
    # a constant image gives constant crops for square boxes of any size (weights sum to 1); a non-square box is padded with
    # 0.5 on its short side, the square part stays constant
    sq = np.array([[100, 100, 100 + s, 100 + s] for s in (17, 64, 255, 256, 257, 640, 1400)], dtype=np.float32)
    c2 = ops.crop_resize(table, 2, torch.from_numpy(sq).cuda(), torch.ones(len(sq), dtype=torch.int32, device="cuda"), S)
    assert float((c2 - 0.25).abs().max()) <= 1e-6
    wide = np.array([[0, 0, 2048, 512]], dtype=np.float32)                  # L = 2048: rows below 512/8 = 64 are padding
    c3 = ops.crop_resize(table, 2, torch.from_numpy(wide).cuda(), torch.ones(1, dtype=torch.int32, device="cuda"), S)
    assert float((c3[0, :, :60] - 0.25).abs().max()) <= 1e-6 and float((c3[0, :, 68:] - 0.5).abs().max()) <= 1e-6
    # all 6400 crops of a C3 step are finite and inside the value range of the source ([0,1) image, 0.5 padding)
    b, s, seg = _shelves(4, 200, 3100)
    idx = np.repeat(np.arange(4), np.diff(seg)).astype(np.int32)
    many = ops.crop_resize(ops.make_image_table([img] * 4), 4, torch.from_numpy(b).cuda(), torch.from_numpy(idx).cuda(), S)
    assert many.shape == (len(b), 3, S, S) and bool(torch.isfinite(many).all())
    assert float(many.min()) >= 0.0 and float(many.max()) <= 1.0


@pytest.mark.parametrize("B", [800, 256, 130])   # the trunk batch of the bench step, a smaller one (clusters of 2) and a ragged last batch (clusters of 4)
def test_mac_exact_maxima_at_trunk_batch_size(ops, B):
    g = torch.Generator(device="cuda").manual_seed(3)
    for cl in (False, True):
        f4 = torch.rand((B, 512, 16, 16), device="cuda", generator=g)       # pool4(relu4_3) and relu5_3 of a trunk batch
        f5 = torch.rand((B, 512, 16, 16), device="cuda", generator=g)
        if cl:
            f4, f5 = f4.contiguous(memory_format=torch.channels_last), f5.contiguous(memory_format=torch.channels_last)
        emb = ops.region_pool_embed([f4, f5])
        raw = torch.cat([f4.amax(dim=(2, 3)), f5.amax(dim=(2, 3))], 1)
        norm = raw.double().norm(dim=1, keepdim=True)
        assert torch.allclose(emb.double(), raw.double() / norm, rtol=1e-6, atol=0)
        unnorm = ops.region_pool_embed([f4, f5], final_l2=False)
        assert torch.equal(unnorm, raw)                                       # maxima are exact
        perm = torch.randperm(256, device="cuda", generator=g)
        f4p = f4.flatten(2)[:, :, perm].reshape(f4.shape)
        f5p = f5.flatten(2)[:, :, perm].reshape(f5.shape)
        if cl:
            f4p, f5p = f4p.contiguous(memory_format=torch.channels_last), f5p.contiguous(memory_format=torch.channels_last)
        assert torch.equal(ops.region_pool_embed([f4p, f5p]), emb)            # pixel order does not matter


def test_search_properties_on_a_c5_shard(ops):
    G, D, Q, k = 125_000, 512, 100_000, 5           # configs[4] on one of 8 GPUs: its gallery shard, all 100k queries
    g = torch.Generator(device="cuda").manual_seed(5)
    gal = torch.nn.functional.normalize(torch.randn((G, D), device="cuda", generator=g), dim=1)
    qry = torch.nn.functional.normalize(torch.randn((Q, D), device="cuda", generator=g), dim=1)
    planted = torch.randint(0, G, (Q // 8,), device="cuda", generator=g)
    qry[: Q // 8] = torch.nn.functional.normalize(gal[planted] + 0.05 * torch.randn((Q // 8, D), device="cuda", generator=g) / D ** 0.5, dim=1)
    gb, gi = ops.rows_prepare(gal)
    qb, qi = ops.rows_prepare(qry)
    sim, idx = ops.sim_topk(qb, gb, k, 0, qry, gal, qi, gi)
    assert torch.equal(idx[: Q // 8, 0], planted)                                   # planted neighbours at rank 1
    assert bool((sim[:, :-1] >= sim[:, 1:]).all())                                  # descending
    assert bool((idx.sort(dim=1).values[:, 1:] != idx.sort(dim=1).values[:, :-1]).all())   # distinct
    exact = (qry[:, None, :].double() * gal[idx].double()).sum(-1)                  # scores are the exact cosines
    assert float((sim.double() - exact).abs().max()) < 2e-6
    # no gallery row beats the k-th returned score (checked in fp32 on a sample of queries against the whole shard)
    rows = torch.arange(0, Q, 61, device="cuda")
    full = qry[rows] @ gal.T
    assert bool((full > (sim[rows, -1:] + 2e-6)).sum(dim=1).le(k - 1).all())
    # shard-merge invariance: top-k over two half shards + merge == top-k over the whole shard
    h = G // 2
    s0, i0 = ops.sim_topk(qb, gb[:h].contiguous(), k, 0, qry, gal[:h].contiguous(), qi, gi[:h].contiguous())
    s1, i1 = ops.sim_topk(qb, gb[h:].contiguous(), k, h, qry, gal[h:].contiguous(), qi, gi[h:].contiguous())
    ms, mi = ops.topk_merge(torch.stack([s0, s1]), torch.stack([i0, i1]))
    assert torch.equal(mi, idx) and torch.equal(ms, sim)
    # cosine is scale-invariant in the query: the indices do not change
    qs = qry * 3.0
    qb3, qi3 = ops.rows_prepare(qs)
    _, idx3 = ops.sim_topk(qb3, gb, k, 0, qs, gal, qi3, gi)
    assert float((idx3 == idx).float().mean()) > 0.9999
