"""GPU parity tests proper: the sm_100a kernels, called through the C ABI (cvpce_b200.ops -> libcvpce_b200.so), against
the CPU oracle (oracle/) and the golden vectors written from the real reference (tests/golden/).

Bars: NMS keep lists and top-k indices bit-exact (ties by the stated order); crops <= 1e-5 abs; pooled maxima exact;
embeddings / similarities <= 1e-3 relative (observed ~1e-6 in fp32).
"""
import numpy as np
import pytest
import torch

import oracle
from cvpce_b200 import synth
from conftest import topk_check

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(cuda):
    from cvpce_b200 import _lib, ops as _ops
    lib = _lib.load()  # raises if the .so is missing: no fallback
    assert lib.cvpce_device_check() == 0, _lib.last_error()
    return _ops


def dev(a, device="cuda"):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


# ---------------------------------------------------------------------------------------------------------------- NMS
def test_nms_golden_single(ops, golden):
    for name, (boxes, scores, thr, per_img) in synth.nms_cases().items():
        keep = ops.nms(dev(boxes), dev(scores), thr).cpu().numpy()
        assert np.array_equal(keep, golden["nms"][name + "/keep"]), name


def test_nms_batched_csr_matches_golden_and_oracle(ops, golden):
    cases = synth.nms_cases()
    names = list(cases)
    boxes = np.concatenate([cases[n][0] for n in names] + [np.zeros((0, 4), np.float32)])
    scores = np.concatenate([cases[n][1] for n in names])
    sizes = [len(cases[n][1]) for n in names]
    # an empty image in the middle of the batch
    sizes.insert(3, 0)
    names.insert(3, None)
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    keep, counts, kboxes, kscores = ops.nms_batched(dev(boxes), dev(scores), dev(seg), max(sizes), 0.5, want_boxes=True, want_scores=True)
    keep, counts, kboxes, kscores = keep.cpu().numpy(), counts.cpu().numpy(), kboxes.cpu().numpy(), kscores.cpu().numpy()
    for i, n in enumerate(names):
        if n is None:
            assert counts[i] == 0
            continue
        b, s, thr, _ = cases[n]
        want = oracle.nms(b, s, 0.5)
        got = keep[seg[i]: seg[i] + counts[i]]
        assert np.array_equal(got, want), n
        assert np.array_equal(kboxes[seg[i]: seg[i] + counts[i]], b[want]), n
        assert np.array_equal(kscores[seg[i]: seg[i] + counts[i]], s[want]), n
        if thr == 0.5:
            assert np.array_equal(got, golden["nms"][n + "/keep"]), n


def test_nms_postprocess_options(ops, golden):
    b, s, thr, per_img = synth.nms_cases()["random5000"]
    seg = dev(np.array([0, len(s)], dtype=np.int32))
    keep, counts, _, _ = ops.nms_batched(dev(b), dev(s), seg, len(s), thr, max_keep=per_img)
    got = keep[: int(counts.item())].cpu().numpy()
    assert np.array_equal(got, golden["nms"]["random5000/keep_postprocess"])  # batched_nms(...)[:detections_per_img]
    # confidence filter of production.py:15 folded in, plus the box rescale of transform.postprocess
    scale = np.array([[1.5, 0.75]], dtype=np.float32)
    keep, counts, kb, _ = ops.nms_batched(dev(b), dev(s), seg, len(s), thr, min_score=0.5, max_keep=per_img, box_scale=dev(scale),
                                          want_boxes=True)
    want = oracle.postprocess_nms(b, s, thr, per_img, min_score=0.5)
    n = int(counts.item())
    assert np.array_equal(keep[:n].cpu().numpy(), want)
    assert np.array_equal(kb[:n].cpu().numpy(), b[want] * np.array([1.5, 0.75, 1.5, 0.75], dtype=np.float32))


def test_nms_random_batch_vs_oracle(ops):
    rs = np.random.RandomState(5)
    sizes = [int(v) for v in rs.randint(0, 1500, 48)] + [8192, 1, 64, 65, 63]
    all_b, all_s = [], []
    for i, n in enumerate(sizes):
        b, s = synth.random_boxes(1000 + i, n, extent=float(rs.choice([200, 600, 1500])), tie_every=int(rs.choice([0, 0, 5])))
        all_b.append(b)
        all_s.append(s)
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    thr = 0.5
    keep, counts, _, _ = ops.nms_batched(dev(np.concatenate(all_b)), dev(np.concatenate(all_s)), dev(seg), max(sizes), thr)
    keep, counts = keep.cpu().numpy(), counts.cpu().numpy()
    for i, n in enumerate(sizes):
        want = oracle.nms(all_b[i], all_s[i], thr)
        assert counts[i] == len(want), (i, n)
        assert np.array_equal(keep[seg[i]: seg[i] + counts[i]], want), (i, n)


def test_nms_threshold_double_semantics(ops):
    b = dev(np.array([[0, 0, 2, 1], [1, 0, 3, 1]], dtype=np.float32))
    s = dev(np.array([0.9, 0.8], dtype=np.float32))
    assert ops.nms(b, s, 1 / 3).tolist() == [0]
    assert ops.nms(b, s, float(np.float32(1 / 3))).tolist() == [0, 1]


def test_nms_limits_and_errors(ops):
    b, s = synth.random_boxes(3, 100)
    seg = dev(np.array([0, 100], dtype=np.int32))
    keep, counts, _, _ = ops.nms_batched(dev(b), dev(s), seg, 50, 0.5)  # max_n too small: flagged, not silently wrong
    assert int(counts.item()) == -1
    with pytest.raises(RuntimeError):
        ops.nms_batched(dev(b), dev(s), seg, 10000, 0.5)
    with pytest.raises(RuntimeError):
        ops.nms(torch.from_numpy(b), torch.from_numpy(s), 0.5)  # CPU tensors: no fallback
    assert ops.nms(dev(np.zeros((0, 4), np.float32)), dev(np.zeros((0,), np.float32)), 0.5).numel() == 0


# --------------------------------------------------------------------------------------------------------------- crop
def test_crops_golden_and_oracle(ops, golden):
    img = synth.shelf_image(synth.CROP_IMAGE_SEED, *synth.CROP_IMAGE_HW)
    boxes = synth.crop_boxes()
    gimg = dev(img)
    table = ops.make_image_table([gimg])
    for key, S, aa in (("aa_S256", 256, True), ("aa_S224", 224, True), ("noaa_S256", 256, False)):
        out, valid = ops.crop_resize(table, 1, dev(boxes), None, S, aa, return_valid=True)
        out = out.cpu().numpy()
        assert valid.cpu().numpy().all()
        got = out.reshape(len(boxes), -1)[:, synth.sample_pixels(S)]
        np.testing.assert_allclose(got, golden["crops"][key], rtol=0, atol=1e-5)
        want, _ = oracle.crop_resize(img, boxes, S, aa)
        np.testing.assert_allclose(out, want, rtol=0, atol=1e-5)


def test_crops_multi_image_layouts_affine(ops):
    from cvpce_b200 import classification
    imgs = [synth.shelf_image(60 + i, h, w) for i, (h, w) in enumerate([(300, 400), (128, 96), (513, 257)])]
    gimgs = [dev(im) for im in imgs]
    rs = np.random.RandomState(9)
    boxes, idx = [], []
    for i, im in enumerate(imgs):
        H, W = im.shape[1:]
        for _ in range(9):
            x1, y1 = rs.uniform(0, W - 8), rs.uniform(0, H - 8)
            boxes.append([x1, y1, x1 + rs.uniform(2, W - x1), y1 + rs.uniform(2, H - y1)])
            idx.append(i)
    boxes = np.asarray(boxes, dtype=np.float32)
    idx = np.asarray(idx, dtype=np.int32)
    perm = rs.permutation(len(idx))
    boxes, idx = boxes[perm], idx[perm]
    S = 96
    want = np.stack([oracle.crop_resize(imgs[i], boxes[j:j + 1], S, True)[0][0] for j, i in enumerate(idx)])
    table = ops.make_image_table(gimgs)
    out = ops.crop_resize(table, 3, dev(boxes), dev(idx), S, True).cpu().numpy()
    np.testing.assert_allclose(out, want, rtol=0, atol=1e-5)
    # fused scale_to_tanh + MACVGG normalisation (utils.py:280-281, classification.py:41-44), channels_last, bf16
    aff = classification.tanh_normalize_affine()
    mean = np.array(classification.NORM_MEAN, dtype=np.float32).reshape(1, 3, 1, 1)
    std = np.array(classification.NORM_STD, dtype=np.float32).reshape(1, 3, 1, 1)
    want_aff = ((want * 2 - 1) - mean) / std
    out = ops.crop_resize(table, 3, dev(boxes), dev(idx), S, True, affine=aff, channels_last=True)
    assert out.shape == (len(idx), 3, S, S) and out.is_contiguous(memory_format=torch.channels_last)
    np.testing.assert_allclose(out.cpu().numpy(), want_aff, rtol=0, atol=2e-5)
    out = ops.crop_resize(table, 3, dev(boxes), dev(idx), S, True, affine=aff, out_dtype=torch.bfloat16)
    np.testing.assert_allclose(out.float().cpu().numpy(), want_aff, rtol=1e-2, atol=1e-2)


def test_crops_band_kernel_paths(ops):
    """The column-marching kernel against the oracle and against the direct kernel: band heights 8..256 (few / many crops),
    up-sampling, 2-4 and 5-6 tap filters, the direct fallback inside the kernel for wider filters (the 900 x 700 box at
    S = 32 / 64), crops narrower than the filter, crops on the image border, all output layouts."""
    from cvpce_b200 import _lib
    img = synth.shelf_image(77, 700, 900)
    gimg = dev(img)
    table = ops.make_image_table([gimg])
    rs = np.random.RandomState(4)
    few = np.array([[0, 0, 900, 700], [10, 20, 600, 500], [100, 100, 140, 130], [5, 5, 300, 690],
                    # crops narrower than the filter (1 and 3 pixels wide; up-sampling, 2-4 and 5-6 taps), a 3 x 3 crop, and
                    # crops touching the right / bottom image border (the shifted tap window must stay inside the row)
                    [100, 100, 101, 300], [10, 10, 13, 600], [50, 50, 53, 53], [700, 10, 900, 400], [880, 650, 900, 700],
                    [899, 0, 900, 700], [0, 699, 900, 700]], dtype=np.float32)
    x1, y1 = rs.uniform(0, 700, 700), rs.uniform(0, 500, 700)
    many = np.stack([x1, y1, x1 + rs.uniform(3, 200, 700), y1 + rs.uniform(3, 200, 700)], 1).astype(np.float32)
    for boxes, S in ((few, 256), (few, 64), (few, 32), (many, 64), (few, 224)):
        want, _ = oracle.crop_resize(img, boxes, S, True)
        out = ops.crop_resize(table, 1, dev(boxes), None, S, True)
        np.testing.assert_allclose(out.cpu().numpy(), want, rtol=0, atol=1e-5)
        nhwc = ops.crop_resize(table, 1, dev(boxes), None, S, True, channels_last=True)
        assert torch.equal(nhwc.contiguous(), out)
        b16 = ops.crop_resize(table, 1, dev(boxes), None, S, True, out_dtype=torch.bfloat16)
        b16n = ops.crop_resize(table, 1, dev(boxes), None, S, True, out_dtype=torch.bfloat16, channels_last=True)
        assert torch.equal(b16, out.to(torch.bfloat16)) and torch.equal(b16n.contiguous(), b16)
        n4 = ops.crop_resize(table, 1, dev(boxes), None, S, True, channels_last=True, pad_channel=True)
        assert torch.equal(n4[:, :3].contiguous(), out) and float(n4[:, 3].abs().max()) == 0.0
        with _lib.tuning(crop_kernel=3):
            direct = ops.crop_resize(table, 1, dev(boxes), None, S, True)
        np.testing.assert_allclose(direct.cpu().numpy(), out.cpu().numpy(), rtol=0, atol=2e-6)
        for rows in (8, 32, 256):
            with _lib.tuning(crop_band_rows=rows):
                assert torch.equal(ops.crop_resize(table, 1, dev(boxes), None, S, True), out)
    # odd sizes (no vector width to respect in the column kernel) and a size above the band kernel's 256
    for S in (33, 100, 320):
        want, _ = oracle.crop_resize(img, few, S, True)
        out = ops.crop_resize(table, 1, dev(few), None, S, True)
        np.testing.assert_allclose(out.cpu().numpy(), want, rtol=0, atol=1e-5)


def test_crops_strided_view_empty_and_zero(ops):
    img = synth.shelf_image(2, 64, 80)
    big = dev(np.pad(img, ((0, 0), (3, 5), (7, 9))))
    view = big[:, 3:67, 7:87]  # non-contiguous view with unit pixel stride
    table = ops.make_image_table([view])
    boxes = np.array([[3, 4, 50, 60], [10, 10, 10, 10], [20, 5, 20, 40]], dtype=np.float32)
    out, valid = ops.crop_resize(table, 1, dev(boxes), None, 32, True, return_valid=True)
    want, wvalid = oracle.crop_resize(img, boxes, 32, True)
    np.testing.assert_allclose(out.cpu().numpy(), want, rtol=0, atol=1e-5)
    assert valid.cpu().numpy().astype(bool).tolist() == wvalid.tolist() == [True, False, True]
    empty = ops.crop_resize(table, 1, dev(np.zeros((0, 4), np.float32)), None, 256, True)
    assert empty.shape == (0, 3, 256, 256)  # production.py:18-19


# --------------------------------------------------------------------------------------------------------------- pool
def test_mac_golden_and_exact_maxima(ops, golden):
    f4, f5 = synth.feature_maps()
    out = ops.region_pool_embed([dev(f4), dev(f5)]).cpu().numpy()
    np.testing.assert_allclose(out, golden["mac"]["mac_given_maps"], rtol=1e-6, atol=1e-7)
    raw = ops.region_pool_embed([dev(f4), dev(f5)], final_l2=False).cpu().numpy()
    assert np.array_equal(raw, oracle.region_pool_embed([f4, f5], final_l2=False))  # max is exact


@pytest.mark.parametrize("B", [1, 3, 37])
def test_mac_vgg_shapes_layouts_dtypes(ops, B):
    rs = np.random.RandomState(B)
    f4 = np.maximum(rs.randn(B, 512, 32, 32), 0).astype(np.float32)
    f5 = np.maximum(rs.randn(B, 512, 16, 16), 0).astype(np.float32)
    want = oracle.region_pool_embed([f4, f5])
    g4, g5 = dev(f4), dev(f5)
    out = ops.region_pool_embed([g4, g5]).cpu().numpy()
    np.testing.assert_allclose(out, want, rtol=2e-6, atol=1e-7)
    out_cl = ops.region_pool_embed([g4.contiguous(memory_format=torch.channels_last), g5.contiguous(memory_format=torch.channels_last)])
    np.testing.assert_allclose(out_cl.cpu().numpy(), want, rtol=2e-6, atol=1e-7)
    b4, b5 = g4.bfloat16(), g5.bfloat16()
    want_bf = oracle.region_pool_embed([b4.float().cpu().numpy(), b5.float().cpu().numpy()])
    np.testing.assert_allclose(ops.region_pool_embed([b4, b5]).cpu().numpy(), want_bf, rtol=2e-6, atol=1e-7)
    np.testing.assert_allclose(ops.region_pool_embed([b4.contiguous(memory_format=torch.channels_last),
                                                      b5.contiguous(memory_format=torch.channels_last)]).cpu().numpy(),
                               want_bf, rtol=2e-6, atol=1e-7)


def test_rmac_regions_mode(ops):
    assert ops.rmac_regions(16, 16, 3) == [tuple(r) for r in oracle.rmac_regions(16, 16, 3).tolist()]
    assert ops.rmac_regions(14, 20, 3) == [tuple(r) for r in oracle.rmac_regions(14, 20, 3).tolist()]
    rs = np.random.RandomState(4)
    feats = [np.maximum(rs.randn(6, 512, 16, 16), 0).astype(np.float32), np.maximum(rs.randn(6, 96, 40, 24), 0).astype(np.float32),
             np.maximum(rs.randn(6, 30, 7, 7), 0).astype(np.float32)]
    regions = [oracle.rmac_regions(f.shape[2], f.shape[3], 3).tolist() for f in feats]
    want = oracle.region_pool_embed(feats, regions, per_region_l2=True)
    out = ops.region_pool_embed([dev(f) for f in feats], regions=regions, per_region_l2=True).cpu().numpy()
    np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-7)
    want = oracle.region_pool_embed(feats, regions, per_region_l2=False, final_l2=False)
    out = ops.region_pool_embed([dev(f) for f in feats], regions=regions, per_region_l2=False, final_l2=False).cpu().numpy()
    np.testing.assert_allclose(out, want, rtol=1e-6, atol=0)


@pytest.mark.parametrize("hw", [(32, 16), (28, 14)])
def test_rmac_vgg_shapes_layouts_dtypes(ops, hw):
    """R-MAC (north-star extension; oracle = Tolias et al. restatement, parity unpinned by the reference) on VGG-sized
    maps through every kernel path: NCHW separable, channels-last separable, bf16, and the generic path (a region
    list with more than 8 distinct ranges)."""
    rs = np.random.RandomState(hw[0])
    B = 5
    f4 = np.maximum(rs.randn(B, 512, hw[0], hw[0]), 0).astype(np.float32)
    f5 = np.maximum(rs.randn(B, 512, hw[1], hw[1]), 0).astype(np.float32)
    regions = [oracle.rmac_regions(hw[0], hw[0], 3).tolist(), oracle.rmac_regions(hw[1], hw[1], 3).tolist()]
    want = oracle.region_pool_embed([f4, f5], regions, per_region_l2=True)
    want_raw = oracle.region_pool_embed([f4, f5], regions, per_region_l2=False, final_l2=False)
    g4, g5 = dev(f4), dev(f5)
    c4, c5 = g4.contiguous(memory_format=torch.channels_last), g5.contiguous(memory_format=torch.channels_last)
    for a, b in ((g4, g5), (c4, c5), (g4, c5)):
        out = ops.region_pool_embed([a, b], regions=regions, per_region_l2=True).cpu().numpy()
        np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-7)
        raw = ops.region_pool_embed([a, b], regions=regions, per_region_l2=False, final_l2=False).cpu().numpy()
        np.testing.assert_allclose(raw, want_raw, rtol=1e-6, atol=0)  # sums of exact maxima
    b4, b5 = g4.bfloat16(), g5.bfloat16()
    want_bf = oracle.region_pool_embed([b4.float().cpu().numpy(), b5.float().cpu().numpy()], regions, per_region_l2=True)
    for a, b in ((b4, b5), (b4.contiguous(memory_format=torch.channels_last), b5.contiguous(memory_format=torch.channels_last))):
        np.testing.assert_allclose(ops.region_pool_embed([a, b], regions=regions, per_region_l2=True).cpu().numpy(), want_bf,
                                   rtol=1e-5, atol=1e-7)
    # irregular list: 12 regions with 12 distinct row and column ranges -> generic (non-separable) path
    irregular = [[(i, i // 2, hw[0] - (i % 3), hw[0] - i // 3) for i in range(12)], [(0, 0, hw[1], hw[1])]]
    want_irr = oracle.region_pool_embed([f4, f5], irregular, per_region_l2=True)
    for a, b in ((g4, g5), (c4, c5)):
        np.testing.assert_allclose(ops.region_pool_embed([a, b], regions=irregular, per_region_l2=True).cpu().numpy(), want_irr,
                                   rtol=1e-5, atol=1e-7)


def _rmac_both_kernels(ops, feats, **kw):
    """The same call through the persistent TMA-staged kernel (forced), its warp-specialised form (forced) and the cluster kernel."""
    from cvpce_b200 import _lib
    with _lib.tuning(rmac_kernel=2):
        staged = ops.region_pool_embed(feats, **kw).cpu().numpy()
    with _lib.tuning(rmac_kernel=4):  # channels-last only calls: the warp-specialised kernel (16 streaming + 8 region warps); else as 2
        ws = ops.region_pool_embed(feats, **kw).cpu().numpy()
    # the same maxima; the norms sum the squares over 8 instead of 16 warps
    np.testing.assert_allclose(ws, staged, rtol=2e-6, atol=1e-7)
    if not kw.get("per_region_l2", False) and not kw.get("final_l2", True):
        assert np.array_equal(ws, staged)
    with _lib.tuning(rmac_kernel=1):
        cluster = ops.region_pool_embed(feats, **kw).cpu().numpy()
    return staged, cluster


@pytest.mark.parametrize("hw", [(32, 16), (28, 14)])
def test_rmac_staged_kernel_vgg_shapes_layouts_dtypes(ops, hw):
    """rmac_staged.cu (ring of cp.async.bulk stages, one persistent CTA per image) against the oracle and against the cluster
    kernel on the VGG maps: NCHW, channels-last and mixed layouts, f32 and bf16, R-MAC levels 3 and 4, raw sums."""
    rs = np.random.RandomState(100 + hw[0])
    B = 7
    f4 = rs.randn(B, 512, hw[0], hw[0]).astype(np.float32)  # signed values: nothing relies on ReLU outputs
    f5 = np.maximum(rs.randn(B, 512, hw[1], hw[1]), 0).astype(np.float32)
    g4, g5 = dev(f4), dev(f5)
    c4, c5 = g4.contiguous(memory_format=torch.channels_last), g5.contiguous(memory_format=torch.channels_last)
    for levels in (3, 4):
        regions = [oracle.rmac_regions(hw[0], hw[0], levels).tolist(), oracle.rmac_regions(hw[1], hw[1], levels).tolist()]
        want = oracle.region_pool_embed([f4, f5], regions, per_region_l2=True)
        want_raw = oracle.region_pool_embed([f4, f5], regions, per_region_l2=False, final_l2=False)
        for a, b in ((g4, g5), (c4, c5), (g4, c5), (c4, g5)):
            staged, cluster = _rmac_both_kernels(ops, [a, b], regions=regions, per_region_l2=True)
            np.testing.assert_allclose(staged, want, rtol=1e-5, atol=1e-7)
            np.testing.assert_allclose(staged, cluster, rtol=1e-5, atol=1e-7)
            raw, raw_cluster = _rmac_both_kernels(ops, [a, b], regions=regions, per_region_l2=False, final_l2=False)
            np.testing.assert_allclose(raw, want_raw, rtol=1e-6, atol=0)
            assert np.array_equal(raw, raw_cluster)  # sums of exact maxima in the same region order
    regions = [oracle.rmac_regions(hw[0], hw[0], 3).tolist(), oracle.rmac_regions(hw[1], hw[1], 3).tolist()]
    b4, b5 = g4.bfloat16(), g5.bfloat16()
    want_bf = oracle.region_pool_embed([b4.float().cpu().numpy(), b5.float().cpu().numpy()], regions, per_region_l2=True)
    if (hw[1] * 2) % 16 == 0:  # NCHW rows must be whole 16-byte units for the staged kernel (else the call takes the cluster kernel)
        layouts = ((b4, b5), (b4.contiguous(memory_format=torch.channels_last), b5.contiguous(memory_format=torch.channels_last)))
    else:
        layouts = ((b4.contiguous(memory_format=torch.channels_last), b5.contiguous(memory_format=torch.channels_last)),)
    for a, b in layouts:
        staged, cluster = _rmac_both_kernels(ops, [a, b], regions=regions, per_region_l2=True)
        np.testing.assert_allclose(staged, want_bf, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(staged, cluster, rtol=1e-5, atol=1e-7)
    # a region list that leaves rows and columns uncovered, plus a whole-map second map with per-region L2
    partial = [[(2, 3, 9, 11), (5, 1, hw[0] - 1, 8), (4, 4, 12, hw[0] - 2)], [(0, 0, hw[1], hw[1])]]
    want_p = oracle.region_pool_embed([f4, f5], partial, per_region_l2=True)
    for a, b in ((g4, g5), (c4, c5)):
        staged, _ = _rmac_both_kernels(ops, [a, b], regions=partial, per_region_l2=True)
        np.testing.assert_allclose(staged, want_p, rtol=1e-5, atol=1e-7)


def test_rmac_staged_kernel_many_images_per_cta_and_odd_maps(ops):
    """More images than SMs (every CTA walks several images and the ring runs across image boundaries), non-square maps,
    narrow planes (two and four NCHW planes per warp pass), pixels smaller than a ring stage (chunks spanning rows),
    three maps, an output row stride, and the automatic choice (staged kernel from B >= SMs/2)."""
    from cvpce_b200 import _lib
    rs = np.random.RandomState(5)
    B = 333
    feats = [rs.randn(B, 128, 14, 20).astype(np.float32), rs.randn(B, 256, 7, 12).astype(np.float32), rs.randn(B, 128, 5, 4).astype(np.float32)]
    regions = [oracle.rmac_regions(f.shape[2], f.shape[3], 3).tolist() for f in feats]
    want = oracle.region_pool_embed(feats, regions, per_region_l2=True)
    nchw = [dev(f) for f in feats]
    nhwc = [t.contiguous(memory_format=torch.channels_last) for t in nchw]
    for maps in (nchw, nhwc, [nchw[0], nhwc[1], nchw[2]]):
        staged, cluster = _rmac_both_kernels(ops, maps, regions=regions, per_region_l2=True)
        np.testing.assert_allclose(staged, want, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(staged, cluster, rtol=1e-5, atol=1e-7)
        auto = ops.region_pool_embed(maps, regions=regions, per_region_l2=True).cpu().numpy()
        again = ops.region_pool_embed(maps, regions=regions, per_region_l2=True).cpu().numpy()
        assert np.array_equal(auto, again)  # B >= SMs/2: the automatic choice is one of the staged kernels, both deterministic
        np.testing.assert_allclose(auto, staged, rtol=2e-6, atol=1e-7)
        wide = torch.full((B, 600), -7.0, device="cuda")
        with _lib.tuning(rmac_kernel=2):
            ops.region_pool_embed(maps, regions=regions, per_region_l2=True, out=wide)
        assert np.array_equal(wide[:, :512].cpu().numpy(), staged) and bool((wide[:, 512:] == -7.0).all())
    # fewer ring stages than chunks per image
    base, _ = _rmac_both_kernels(ops, nhwc, regions=regions, per_region_l2=True)
    with _lib.tuning(rmac_kernel=4):
        base_ws = ops.region_pool_embed(nhwc, regions=regions, per_region_l2=True).cpu().numpy()
    for stages in (3, 4, 5):
        with _lib.tuning(rmac_kernel=2, rmac_stages=stages):
            out = ops.region_pool_embed(nhwc, regions=regions, per_region_l2=True).cpu().numpy()
        assert np.array_equal(out, base)
        with _lib.tuning(rmac_kernel=4, rmac_stages=stages):
            out = ops.region_pool_embed(nhwc, regions=regions, per_region_l2=True).cpu().numpy()
        assert np.array_equal(out, base_ws)


def test_l2norm_rows(ops):
    x = np.random.RandomState(1).randn(77, 1000).astype(np.float32)
    x[5] = 0
    out = ops.l2norm_rows_(dev(x)).cpu().numpy()
    want = x / np.maximum(np.linalg.norm(x.astype(np.float64), axis=1, keepdims=True), 1e-8)
    np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-7)


# ---------------------------------------------------------------------------------------------------------------- sim
def test_nearest_neighbors_reference_kat(ops):
    from cvpce_b200 import classification
    a, q, expected = synth.nn_kat()
    got = classification.nearest_neighbors(dev(a), dev(q))[:, 0].cpu().numpy()  # reference test/models/classification_test.py
    assert np.array_equal(got, expected)


def test_sim_topk_f32_vs_oracle_and_golden(ops, golden):
    for name, (Q, G, D, k, seed) in synth.NN_CASES.items():
        g, q = synth.nn_case(Q, G, D, seed)
        sim, idx = ops.sim_topk_f32(dev(q), dev(g), k)
        osim, oidx = oracle.sim_topk(q, g, k)
        topk_check(q, g, idx.cpu().numpy(), osim, oidx)
        assert np.array_equal(idx.cpu().numpy(), golden["nn"][name]), name
        np.testing.assert_allclose(sim.cpu().numpy(), osim, rtol=1e-3, atol=1e-6)


def test_sim_topk_f32_unnormalised_ties_padding(ops):
    rs = np.random.RandomState(2)
    g = (rs.randn(500, 96) * rs.uniform(0.1, 5, (500, 1))).astype(np.float32)
    q = (rs.randn(70, 96) * 3).astype(np.float32)
    g[7] = 0
    g[100] = g[50]
    g[200] = g[50] * 2  # same direction: exact tie after normalisation up to rounding
    sim, idx = ops.sim_topk_f32(dev(q), dev(g), 8, idx_base=1000)
    osim, oidx = oracle.sim_topk(q, g, 8)
    topk_check(q, g, idx.cpu().numpy() - 1000, osim, oidx, tol=1e-5)
    # fewer gallery rows than k: padded with (-inf, -1)
    sim, idx = ops.sim_topk_f32(dev(q[:3]), dev(g[:2]), 4)
    assert (idx[:, 2:] == -1).all() and torch.isinf(sim[:, 2:]).all()
    # exact duplicates: ascending index
    gg = np.array([[1, 0], [1, 0], [0, 1], [1, 0]], dtype=np.float32)
    _, idx = ops.sim_topk_f32(dev(np.array([[1, 0]], dtype=np.float32)), dev(gg), 3)
    assert idx.tolist() == [[0, 1, 3]]


def test_topk_merge(ops):
    rs = np.random.RandomState(3)
    W, Q, k = 8, 300, 5
    scores = rs.randn(W, Q, k).astype(np.float32)
    scores[:, ::3, :] = np.round(scores[:, ::3, :], 1)  # force ties
    scores = -np.sort(-scores, axis=2)
    idx = np.stack([rs.permutation(1000)[:k] + w * 1000 for w in range(W) for _ in range(Q)]).reshape(W, Q, k).astype(np.int64)
    scores[3, 10, 2:] = -np.inf
    idx[3, 10, 2:] = -1
    s, i = ops.topk_merge(dev(scores), dev(idx))
    for r in range(Q):
        cand = [(-(scores[w, r, j]), idx[w, r, j]) for w in range(W) for j in range(k) if idx[w, r, j] >= 0]
        cand.sort()
        assert i[r].tolist() == [c[1] for c in cand[:k]], r
        assert s[r].tolist() == [-c[0] for c in cand[:k]], r


@pytest.mark.parametrize("Q,G,D,k", [(64, 180, 1024, 1), (300, 5000, 512, 5), (1000, 8192, 1024, 5), (130, 777, 64, 3),
                                     (257, 40000, 512, 5), (5, 100000, 512, 5)])
def test_sim_topk_tensor_core_vs_oracle(ops, Q, G, D, k):
    rs = np.random.RandomState(Q + G)
    g = synth.unit_rows(rs, G, D, nonneg=(D == 1024))
    q, _ = synth.planted_queries(rs, g, Q, frac=0.3)
    gq, gg = dev(q), dev(g)
    qb, qinv = ops.rows_prepare(gq)
    gb, ginv = ops.rows_prepare(gg)
    sim, idx = ops.sim_topk(qb, gb, k, 0, gq, gg, qinv, ginv)
    osim, oidx = oracle.sim_topk(q, g, k)
    rate = topk_check(q, g, idx.cpu().numpy(), osim, oidx)
    assert rate > 0.999, rate
    np.testing.assert_allclose(sim.cpu().numpy(), osim, rtol=1e-3, atol=1e-6)
    # cross-check with the exact fp32 CUDA-core kernel on the device
    sim32, idx32 = ops.sim_topk_f32(gq, gg, k)
    assert (idx32 == idx).float().mean().item() > 0.999
    # bf16-operand scores without the re-score: within bf16 rounding of the exact ones, top-1 agreement reported
    simb, idxb = ops.sim_topk(qb, gb, k)
    np.testing.assert_allclose(simb[:, 0].cpu().numpy(), osim[:, 0], rtol=0, atol=2e-3)


@pytest.mark.parametrize("Q,G,D,k", [(1300, 5000, 1024, 5), (2000, 3000, 512, 5), (700, 20000, 64, 1), (1500, 6000, 256, 20)])
def test_sim_topk_gallery_split_counts_agree(ops, Q, G, D, k):
    """The tcgen05 search gives the same bits in exact mode whatever distributes the work: the number of gallery splits
    (candidate lists per row) the plan or a tuning override selects, the (query block, split) grid or linear item ranges over
    all SMs (a CTA's range crosses query blocks: several list lifetimes per CTA, a varying number of lists per row), key lists
    or compare / select lists; exact mode also equals the fp32 CUDA-core kernel."""
    from cvpce_b200 import _lib
    rs = np.random.RandomState(Q + G + D)
    g = synth.unit_rows(rs, G, D, nonneg=(D == 1024))
    q, _ = synth.planted_queries(rs, g, Q, frac=0.3)
    gq, gg = dev(q), dev(g)
    qb, qinv = ops.rows_prepare(gq)
    gb, ginv = ops.rows_prepare(gg)
    ref = None
    for split in ({}, {"sim_split": 1}, {"sim_split": 2}, {"sim_split": 3}, {"sim_split": 7}, {"sim_plan": 1}, {"sim_plan": 2}):
        with _lib.tuning(**split):
            exact = ops.sim_topk(qb, gb, k, 7, gq, gg, qinv, ginv)
            approx = ops.sim_topk(qb, gb, k, 7)
        torch.cuda.synchronize()
        if ref is None:
            ref = (exact, approx)
            s32, i32 = ops.sim_topk_f32(gq, gg, k, idx_base=7)
            assert (i32 == exact[1]).float().mean().item() > 0.999
        else:
            assert torch.equal(exact[0], ref[0][0]) and torch.equal(exact[1], ref[0][1]), split
            # bf16-only scores: lists that live for <= 16 tiles are kept as keys (scores truncated to 11 mantissa bits, ties
            # by column), longer ones in full precision: the same rows up to near-ties below that resolution
            assert (approx[1] == ref[1][1]).float().mean().item() > 0.98, split
            np.testing.assert_allclose(approx[0].cpu().numpy(), ref[1][0].cpu().numpy(), rtol=6e-4, atol=1e-6)
    with _lib.tuning(sim_keys=1):                                     # key lists off: the full-precision lists everywhere
        exact = ops.sim_topk(qb, gb, k, 7, gq, gg, qinv, ginv)
    assert torch.equal(exact[0], ref[0][0]) and torch.equal(exact[1], ref[0][1])


def test_sim_topk_shard_offsets_and_merge(ops):
    rs = np.random.RandomState(8)
    G, D, Q, k, W = 4096, 512, 200, 5, 4
    g = synth.unit_rows(rs, G, D, nonneg=False)
    q, _ = synth.planted_queries(rs, g, Q)
    gq = dev(q)
    qb, qinv = ops.rows_prepare(gq)
    parts_s, parts_i = [], []
    for w in range(W):
        gs = dev(g[w * G // W:(w + 1) * G // W])
        gb, ginv = ops.rows_prepare(gs)
        s, i = ops.sim_topk(qb, gb, k, w * G // W, gq, gs, qinv, ginv)
        parts_s.append(s)
        parts_i.append(i)
    s, i = ops.topk_merge(torch.stack(parts_s), torch.stack(parts_i))
    osim, oidx = oracle.sim_topk(q, g, k)
    assert topk_check(q, g, i.cpu().numpy(), osim, oidx) > 0.999


# --------------------------------------------------------------------------------------------------- encoder / pipeline
def test_macvgg_mirror_matches_reference_embeddings(ops, golden):
    from cvpce_b200 import classification
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    probe = torch.cat([enc.block1[0].weight.reshape(-1)[:8], enc.block2[-2].weight.reshape(-1)[:8]]).numpy()
    if not np.allclose(probe, golden["mac"]["macvgg_weight_probe"], atol=1e-7):
        pytest.skip("seeded VGG16 init differs from the authoring container's torch build")
    enc = enc.cuda()
    with torch.no_grad():
        emb = enc(dev(synth.encoder_inputs())).cpu().numpy()
    np.testing.assert_allclose(emb, golden["mac"]["macvgg_embeddings"], rtol=1e-3, atol=1e-5)


def test_classifier_mirror_matches_reference_labels(ops, golden):
    from cvpce_b200 import classification, production
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).cuda()
    gallery, queries = synth.classify_case()
    clf = production.Classifier(enc, synth.GalleryDataset(gallery), batch_size=4, num_workers=0, k=3)
    res, emb = clf.classify(torch.from_numpy(queries), return_embedding=True)
    np.testing.assert_allclose(clf.embedding.cpu().numpy(), golden["classify"]["gallery_embedding"], rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(emb.cpu().numpy(), golden["classify"]["query_embedding"], rtol=1e-3, atol=1e-5)
    labels = np.array([[int(a[1:]) for a in row] for row in res])
    assert np.array_equal(labels, golden["classify"]["labels"])
    assert clf.classify(torch.empty((0, 3, 64, 64))) == []
    # f2: index file round trip in the reference's format, then the same labels through a sharded gallery
    # (rank 1 of 2 holds only the upper half of the rows; merged with rank 0's block the result is the unsharded one)
    import os
    import tempfile
    from cvpce_b200 import dist as cdist
    with tempfile.TemporaryDirectory() as tmp:
        pth = os.path.join(tmp, "index.pt")
        clf.save_index(pth)
        clf2 = production.Classifier(enc, None, batch_size=4, num_workers=0, k=3, load=pth)
        assert torch.equal(clf2.embedding, clf.embedding) and clf2.annotations == clf.annotations
        clf2.shard_gallery(1, 0)
        assert clf2.classify(torch.from_numpy(queries)) == res
        blocks = []
        for rank in range(2):
            sg, ann = cdist.ShardedGallery.from_index(pth, 2, rank, torch.device("cuda"))
            assert ann == clf.annotations and sg.g32.shape[0] in (len(ann) // 2, len(ann) - len(ann) // 2)
            sg.world = 1                                      # single process: search the shard, merge by hand below
            blocks.append(sg.search(emb, 3))
        _, merged = ops.topk_merge(torch.stack([b[0] for b in blocks]), torch.stack([b[1] for b in blocks]))
        assert [[clf.annotations[j] for j in row] for row in merged.cpu().tolist()] == res


# ------------------------------------------------------------------------------ f1: pre-NMS candidate selection (+ a1, a2)
def _head_to_device(name):
    logits, regression, anchors, sizes, C, topk = synth.head_case(name)
    # like torchvision: one concatenated head output split into per-level VIEWS (tv:retinanet.py:640-648)
    lg_all = dev(np.concatenate(logits, axis=1))
    rg_all = dev(np.concatenate(regression, axis=1))
    n_per = [a.shape[0] for a in anchors]
    return (list(lg_all.split(n_per, dim=1)), list(rg_all.split(n_per, dim=1)), [dev(a) for a in anchors], sizes, C, topk,
            (logits, regression, anchors))


@pytest.mark.parametrize("name", list(synth.HEAD_CASES))
def test_select_candidates_bit_exact_vs_oracle(ops, name):
    lg, rg, an, sizes, C, topk, host = _head_to_device(name)
    boxes, scores, labels, seg_counts, level_counts, stride = ops.select_candidates(lg, rg, an, sizes, 0.05, topk)
    want = oracle.select_candidates(*host, sizes, 0.05, topk)
    boxes, scores, labels = boxes.cpu().numpy(), scores.cpu().numpy(), labels.cpu().numpy()
    seg_counts, level_counts = seg_counts.cpu().numpy(), level_counts.cpu().numpy()
    assert stride == len(lg) * topk
    for b, (wb, ws, wl, wc) in enumerate(want):
        n = int(seg_counts[b])
        assert n == len(ws), (name, b)
        assert np.array_equal(level_counts[b], wc), (name, b)
        sl = slice(b * stride, b * stride + n)
        assert np.array_equal(scores[sl].view(np.uint32), ws.view(np.uint32)), (name, b)   # bit-exact scores
        assert np.array_equal(labels[sl], wl), (name, b)
        assert np.array_equal(boxes[sl].view(np.uint32), wb.view(np.uint32)), (name, b)     # bit-exact decoded boxes


@pytest.mark.parametrize("name", ["dense_2img", "sparse", "fullsize"])
def test_postprocess_detections_matches_reference_gln(ops, golden, name):
    from cvpce_b200 import detection
    from conftest import rows_check
    lg, rg, an, sizes, C, topk, host = _head_to_device(name)
    dets = detection.postprocess_detections({"cls_logits": lg, "bbox_regression": rg}, [an] * len(sizes), sizes,
                                            score_thresh=0.05, topk_candidates=topk, nms_thresh=0.5, detections_per_img=1000)
    want = oracle.postprocess_detections(*host, sizes, 0.05, topk, 0.5, 1000)
    for b, d in enumerate(dets):
        pre = "%s/%d/" % (name, b)
        assert np.array_equal(d["boxes"].cpu().numpy(), want[b][0])       # bit-exact vs the oracle
        assert np.array_equal(d["scores"].cpu().numpy(), want[b][1])
        rows_check(d["boxes"].cpu().numpy(), d["scores"].cpu().numpy(), golden["head"][pre + "det_boxes"],
                   golden["head"][pre + "det_scores"])                     # and the real GLN post-processing within ulps
        assert np.array_equal(d["labels"].cpu().numpy(), golden["head"][pre + "det_labels"])


def test_select_candidates_edge_cases(ops):
    lg, rg, an, sizes, C, topk, host = _head_to_device("sparse")
    # nothing passes / NaN logits / topk = 1 / constant logits (one giant tie group: the lowest flat indices win)
    _, _, _, cnt, _, _ = ops.select_candidates(lg, rg, an, sizes, 0.999999, topk)
    assert int(cnt.sum()) == 0
    nan_lg = [torch.full_like(l, float("nan")) for l in lg]
    _, _, _, cnt, _, _ = ops.select_candidates(nan_lg, rg, an, sizes, 0.05, topk)
    assert int(cnt.sum()) == 0
    _, s1, _, cnt, lc, st = ops.select_candidates(lg, rg, an, sizes, 0.05, 1)
    want = oracle.select_candidates(*host, sizes, 0.05, 1)[0]
    assert np.array_equal(lc.cpu().numpy()[0], want[3]) and np.array_equal(s1[: int(cnt[0])].cpu().numpy(), want[1])
    const_lg = [torch.full_like(l, 0.25) for l in lg]
    b, s, l, cnt, lc, st = ops.select_candidates(const_lg, rg, an, sizes, 0.05, topk)
    host_const = ([np.full_like(x, 0.25) for x in host[0]], host[1], host[2])
    want = oracle.select_candidates(*host_const, sizes, 0.05, topk)[0]
    n = int(cnt[0])
    assert n == len(want[1]) and np.array_equal(b[:n].cpu().numpy(), want[0])
    # negative / zero threshold: every finite logit passes
    _, s0, _, cnt0, _, _ = ops.select_candidates(lg, rg, an, sizes, 0.0, 50)
    want = oracle.select_candidates(*host, sizes, 0.0, 50)[0]
    assert np.array_equal(s0[: int(cnt0[0])].cpu().numpy(), want[1])
    with pytest.raises(RuntimeError):
        ops.select_candidates(lg, rg, an, sizes, 0.05, 4096)     # topk above CVPCE_SELECT_MAX_TOPK


def test_nms_segments_equals_csr_form(ops):
    b, s, thr, per_img = synth.nms_cases()["random5000"]
    n = len(s)
    stride = n + 37
    boxes = np.zeros((2 * stride, 4), np.float32)
    scores = np.zeros(2 * stride, np.float32)
    boxes[:n], scores[:n] = b, s
    m = n // 3
    boxes[stride:stride + m], scores[stride:stride + m] = b[:m], s[:m]
    keep, counts, kb, _ = ops.nms_segments(dev(boxes), dev(scores), stride, dev(np.array([n, m], np.int32)), thr, want_boxes=True)
    keep, counts = keep.cpu().numpy(), counts.cpu().numpy()
    assert np.array_equal(keep[: counts[0]], oracle.nms(b, s, thr))
    assert np.array_equal(keep[stride: stride + counts[1]], oracle.nms(b[:m], s[:m], thr))


@pytest.mark.parametrize("thr", [0.05, 0.3, 0.5, 0.7, 0.0])
def test_nms_grid_window_adversarial_geometry(ops, thr):
    """The kernel only tests a candidate against kept boxes whose centre lies in a window around it (a necessary condition
    for IoU > thr) and files kept boxes in a 32x32 grid with an overflow list: nested / concentric boxes (one cell
    overflows), clusters, extreme aspect ratios, tiny and huge boxes, degenerate and duplicated boxes -- keep lists must
    stay bit-exact with the all-pairs oracle at every threshold."""
    rs = np.random.RandomState(int(thr * 100) + 3)
    parts = []
    c = rs.uniform(400, 600, (1, 2))
    half = np.linspace(2, 380, 700)[:, None] * rs.uniform(0.9, 1.1, (700, 2))          # concentric: all centres in one cell
    parts.append(np.concatenate([c - half, c + half], 1))
    ctr = rs.uniform(0, 1000, (40, 2))[rs.randint(0, 40, 1500)] + rs.normal(0, 6, (1500, 2))   # 40 tight clusters
    wh = rs.uniform(20, 90, (1500, 2))
    parts.append(np.concatenate([ctr - wh / 2, ctr + wh / 2], 1))
    ctr = rs.uniform(0, 1000, (600, 2))                                                 # extreme aspect ratios
    wh = np.stack([rs.uniform(2, 900, 600), rs.uniform(2, 30, 600)], 1)[:, rs.permutation(2)]
    parts.append(np.concatenate([ctr - wh / 2, ctr + wh / 2], 1))
    ctr = rs.uniform(0, 1000, (300, 2))                                                 # tiny boxes
    wh = rs.uniform(1e-3, 0.5, (300, 2))
    parts.append(np.concatenate([ctr - wh / 2, ctr + wh / 2], 1))
    parts.append(np.array([[0, 0, 1000, 1000], [-5000, -5000, 6000, 6000], [10, 10, 10, 50], [30, 40, 20, 60]], np.float64))
    boxes = np.concatenate(parts).astype(np.float32)
    boxes = np.concatenate([boxes, boxes[rs.randint(0, len(boxes), 200)]])               # exact duplicates
    scores = rs.uniform(0, 1, len(boxes)).astype(np.float32)
    scores[rs.randint(0, len(scores), 300)] = 0.5                                       # score ties
    want = oracle.nms(boxes, scores, thr)
    got = ops.nms(dev(boxes), dev(scores), thr).cpu().numpy()
    assert np.array_equal(got, want), (thr, len(got), len(want))


def test_hot_path_from_head_matches_hot_path_from_boxes(ops):
    """HotPath.step_from_head (f1 -> a1 -> ...) on head outputs that encode a shelf's scored candidates gives the same kept
    boxes (to fp32 box-coder rounding) and the same matches as HotPath.step on the candidates themselves."""
    from cvpce_b200 import classification, pipeline
    torch.backends.cudnn.allow_tf32 = False
    H, W = 384, 512
    frame = (300, 400)
    img = dev(synth.shelf_image(3, H, W))
    boxes, scores = synth.shelf_candidates(3, H, W, n_products=12)
    boxes = np.clip(boxes, 0, [W, H, W, H]).astype(np.float32)
    lg, rg, anc, scale = synth.head_from_candidates(boxes, scores, (H, W), 5, frame=frame)
    n_per = [a.shape[0] for a in anc]
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).cuda()
    gallery = dev(synth.unit_rows(np.random.RandomState(2), 300, 1024))
    hp = pipeline.HotPath(enc, gallery, k=1, size=64, trunk_batch=8, detections_per_img=200, min_score=0.3)
    seg = dev(np.array([0, len(scores)], dtype=np.int32))
    a = hp.step([img], dev(boxes), dev(scores), seg, len(scores))
    b = hp.step_from_head([img], list(dev(lg)[None].split(n_per, 1)), list(dev(rg)[None].split(n_per, 1)), [dev(x) for x in anc],
                          [frame], dev(np.array([scale], dtype=np.float32)))
    assert int(a["counts"][0]) == int(b["counts"][0]) > 0
    np.testing.assert_allclose(b["boxes"].cpu().numpy(), a["boxes"].cpu().numpy(), atol=2e-2)
    # oracle for the head path itself: selection + NMS + rescale
    L = [lg[None, o:o + n] for o, n in zip(np.cumsum([0] + n_per[:-1]), n_per)]
    R = [rg[None, o:o + n] for o, n in zip(np.cumsum([0] + n_per[:-1]), n_per)]
    wb, ws, _ = oracle.postprocess_detections(L, R, anc, [frame], 0.05, 1000, 0.5, 200)[0]
    wb = wb[ws > np.float32(0.3)] * np.array([scale[0], scale[1], scale[0], scale[1]], dtype=np.float32)
    assert np.array_equal(b["boxes"].cpu().numpy(), wb)
    agree = (a["topk_idx"] == b["topk_idx"]).float().mean().item()
    assert agree >= 0.9, agree


def test_macresnet_mirror_matches_reference_embeddings(ops, golden):
    """f4: MACResNet (classification.py:53-85) over a seeded random-init resnet50, descriptor layers (2, 3) -> 1536-d."""
    from torchvision.models import resnet
    from cvpce_b200 import classification
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(synth.WEIGHT_SEED + 1)
    enc = classification.MACResNet(resnet.resnet50(weights=None), (2, 3)).eval().requires_grad_(False)
    probe = enc.blocks[0][0][0].weight.reshape(-1)[:8].numpy()
    if not np.allclose(probe, golden["mac"]["macresnet_weight_probe"], atol=1e-7):
        pytest.skip("seeded resnet50 init differs from the authoring container's torch build")
    enc = enc.cuda()
    with torch.no_grad():
        emb = enc(dev(synth.encoder_inputs())).cpu().numpy()
    assert emb.shape == (4, 1536)
    np.testing.assert_allclose(emb, golden["mac"]["macresnet_embeddings"], rtol=1e-3, atol=1e-5)


@pytest.mark.parametrize("shape", [(3, 64, 32, 32), (2, 128, 17, 23), (1, 512, 2, 2), (5, 8, 6, 4)])
def test_maxpool2x2_nhwc_equals_torch(ops, shape):
    """a7 driving: the trunk's nn.MaxPool2d(2, 2) on channels-last maps -- bit-exact, NaN propagating, f32 and bf16."""
    g = torch.Generator(device="cuda").manual_seed(shape[1])
    x = torch.randn(shape, device="cuda", generator=g)
    x[0, 1, 0, 1] = float("nan")
    x[-1, -1, -1, -1] = float("-inf")
    for dt in (torch.float32, torch.bfloat16):
        xc = x.to(dt).contiguous(memory_format=torch.channels_last)
        if not ops.maxpool2x2_supported(xc):
            assert dt == torch.float32 or shape[1] % 8
            continue
        got = ops.maxpool2x2(xc)
        want = torch.nn.functional.max_pool2d(xc, 2, 2)
        assert got.shape == want.shape and got.is_contiguous(memory_format=torch.channels_last)
        assert torch.equal(torch.nan_to_num(got.float(), nan=12345.0), torch.nan_to_num(want.float(), nan=12345.0))
    with pytest.raises(RuntimeError):
        ops.maxpool2x2(x)  # NCHW-contiguous input is rejected (the caller keeps torch's module for it)


def test_fast_trunk_features_equal_plain_trunk(ops):
    from cvpce_b200 import classification
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).cuda()
    x = dev(synth.encoder_inputs())
    with torch.no_grad():
        p4, p5 = enc.features(x)
        enc.enable_fast_trunk()
        f4, f5 = enc.features(x)
    np.testing.assert_allclose(f4.cpu().numpy(), p4.cpu().numpy(), rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(f5.cpu().numpy(), p5.cpu().numpy(), rtol=1e-3, atol=1e-5)


@pytest.mark.parametrize("shapes", [[(512, 32, 32), (512, 16, 16)], [(512, 7, 5), (1024, 4, 3)], [(64, 9, 9), (128, 3, 3), (256, 1, 2)],
                                    [(4, 2, 2)], [(1024, 16, 16)]])
def test_mac_channels_last_pixel_split_kernel(ops, shapes):
    """Dense channels-last MAC (the fast trunk's layout) goes through the pixel-split cluster kernel: odd pixel counts, 1..3
    maps, VGG / ResNet channel counts, f32 and bf16, with and without the final L2."""
    rs = np.random.RandomState(len(shapes) * 7 + shapes[0][0])
    B = 5
    feats = [rs.randn(B, c, h, w).astype(np.float32) for c, h, w in shapes]
    for final in (True, False):
        want = oracle.region_pool_embed(feats, final_l2=final)
        dv = [dev(f).contiguous(memory_format=torch.channels_last) for f in feats]
        got = ops.region_pool_embed(dv, final_l2=final).cpu().numpy()
        if final:
            np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-7)
        else:
            assert np.array_equal(got, want)
        bf = [d.bfloat16() for d in dv]
        want_bf = oracle.region_pool_embed([b.float().cpu().numpy() for b in bf], final_l2=final)
        np.testing.assert_allclose(ops.region_pool_embed(bf, final_l2=final).cpu().numpy(), want_bf, rtol=2e-6, atol=1e-7)


@pytest.mark.parametrize("B", [1, 3, 37, 256, 300])
def test_mac_balanced_streaming_kernel_equals_cluster_kernel(ops, B):
    """The balanced streaming MAC kernel (workspace form, default behind ops.region_pool_embed) against the per-image cluster
    kernel and the oracle: maxima bit-equal, descriptors within rounding of the norm; the workspace counters are left clean
    (a second and third launch on the same workspace give the same bits); ragged pixel counts; tiny grids (B = 1: fewer chunks
    than CTAs) and batches whose images straddle CTA ranges."""
    from cvpce_b200 import _lib
    rs = np.random.RandomState(B)
    shapes = [(512, 16, 16), (512, 16, 16)] if B >= 256 else [(512, 9, 7), (256, 5, 3), (1024, 2, 2)]
    feats = [rs.randn(B, c, h, w).astype(np.float32) for c, h, w in shapes]
    dv = [dev(f).contiguous(memory_format=torch.channels_last) for f in feats]
    for final in (False, True):
        with _lib.tuning(mac_kernel=1):
            want = ops.region_pool_embed(dv, final_l2=final)
        got = [ops.region_pool_embed(dv, final_l2=final) for _ in range(3)]
        assert torch.equal(got[0], got[1]) and torch.equal(got[0], got[2])
        if final:
            np.testing.assert_allclose(got[0].cpu().numpy(), want.cpu().numpy(), rtol=2e-6, atol=1e-7)
        else:
            assert torch.equal(got[0], want)
            if B <= 37:
                assert np.array_equal(got[0].cpu().numpy(), oracle.region_pool_embed(feats, final_l2=False))
    raw = ops.region_pool_embed(dv, final_l2=False)
    for per_sm in (1, 7):                                             # other grid sizes: other segmentations, same answer
        with _lib.tuning(mac_ctas_per_sm=per_sm):
            assert torch.equal(ops.region_pool_embed(dv, final_l2=False), raw)
    hf = [d.half() for d in dv]
    with _lib.tuning(mac_kernel=1):
        want_h = ops.region_pool_embed(hf)
    np.testing.assert_allclose(ops.region_pool_embed(hf).cpu().numpy(), want_h.cpu().numpy(), rtol=2e-6, atol=1e-7)
    from cvpce_b200 import ops as _ops
    assert all(int(w[:65536].view(torch.int32).abs().sum()) == 0 for w in _ops._POOL_WS.values())


def test_mac_of_pooled_relu4_3_is_bit_identical(ops):
    """MACVGG.features(pooled_f4=True): max over pool4(relu4_3) == max over relu4_3 when the map has even sides."""
    from cvpce_b200 import classification
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).cuda().enable_fast_trunk()
    x = dev(synth.encoder_inputs())
    with torch.no_grad():
        f4, f5 = enc.features(x)
        p4, q5 = enc.features(x, pooled_f4=True)
        assert p4.shape[2] * 2 == f4.shape[2] and torch.equal(f5, q5)
        assert torch.equal(ops.region_pool_embed([f4, f5]), ops.region_pool_embed([p4, q5]))
        odd = enc.features(x[:, :, :40, :40], pooled_f4=True)[0]      # relu4_3 is 5x5: not tiled by 2x2 windows -> unchanged
        assert odd.shape[2] == 5


def test_sim_topk_scatter_writes_every_target(ops):
    """The gallery-sharded search's kernel-side scatter: the [Q,k] block lands in every target buffer (here two local
    buffers standing in for two ranks' gather slots) and equals the plain sim_topk result."""
    rs = np.random.RandomState(5)
    Q, G, D, k = 300, 5000, 512, 5
    q, g = dev(synth.unit_rows(rs, Q, D, nonneg=False)), dev(synth.unit_rows(rs, G, D, nonneg=False))
    qb, qi = ops.rows_prepare(q, True)
    gb, gi = ops.rows_prepare(g, True)
    s, i = ops.sim_topk(qb, gb, k, 1000, q, g, qi, gi)
    W = 3
    gs = torch.full((2, W, Q, k), -7.0, device="cuda")
    gidx = torch.full((2, W, Q, k), -7, dtype=torch.int64, device="cuda")
    slot = 1
    ops.sim_topk_scatter(qb, gb, k, 1000, [gs[r, slot].data_ptr() for r in range(2)], [gidx[r, slot].data_ptr() for r in range(2)],
                         q, g, qi, gi)
    for r in range(2):
        assert torch.equal(gs[r, slot], s) and torch.equal(gidx[r, slot], i)
        assert bool((gs[r, 0] == -7).all()) and bool((gidx[r, 2] == -7).all())   # other slots untouched


def test_float16_hand_off_kernels(ops):
    """Extension mode (--trunk f16): the crop kernel writes IEEE half directly, the max-pool and the dense channels-last MAC
    read it.  Crops: the fp32 result of the oracle rounded to half; pools / MAC: exact on the half inputs."""
    img = synth.shelf_image(synth.CROP_IMAGE_SEED, *synth.CROP_IMAGE_HW)
    boxes = synth.crop_boxes()[:6]
    want, _ = oracle.crop_resize(img, boxes, 64, True)
    table = ops.make_image_table([dev(img)])
    for cl in (False, True):
        got = ops.crop_resize(table, 1, dev(boxes), None, 64, True, 0.5, None, torch.float16, cl)
        assert got.dtype == torch.float16 and got.is_contiguous(memory_format=torch.channels_last if cl else torch.contiguous_format)
        ref16 = torch.from_numpy(want).half()
        diff = (got.cpu().float() - ref16.float()).abs().max().item()
        assert diff <= 2 ** -11, diff     # at most one half ulp apart where the fp32 values differ in the last bits
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn((3, 64, 10, 12), device="cuda", generator=g).half().contiguous(memory_format=torch.channels_last)
    assert torch.equal(ops.maxpool2x2(x), torch.nn.functional.max_pool2d(x, 2, 2))
    feats = [torch.randn((4, 512, 6, 6), device="cuda", generator=g).half().contiguous(memory_format=torch.channels_last),
             torch.randn((4, 512, 3, 3), device="cuda", generator=g).half().contiguous(memory_format=torch.channels_last)]
    want = oracle.region_pool_embed([f.float().cpu().numpy() for f in feats])
    np.testing.assert_allclose(ops.region_pool_embed(feats).cpu().numpy(), want, rtol=2e-6, atol=1e-7)
    with pytest.raises(RuntimeError):
        ops.region_pool_embed([f.contiguous() for f in feats])   # float16 NCHW maps: unsupported, fails loudly


# ------------------------------------------------------------------------------ f3: box IoU + greedy IoU claim (matching)
def test_box_iou_bit_equal_to_torchvision_golden_and_oracle(ops, golden):
    g = golden["match"]
    for name, (targets, preds) in synth.match_cases().items():
        got = ops.box_iou(dev(preds), dev(targets)).cpu().numpy()
        assert np.array_equal(got.view(np.uint32), g[name + "/iou"].view(np.uint32)), name
    rs = np.random.RandomState(5)
    for na, nb in ((1, 1), (3, 5), (257, 1023), (1000, 1024), (64, 4099)):   # unaligned and vector-store row pitches
        a, _ = synth.random_boxes(rs.randint(1 << 30), na, extent=400.0)
        b, _ = synth.random_boxes(rs.randint(1 << 30), nb, extent=400.0)
        b[::7, 2:] = b[::7, :2]                                               # zero-area boxes
        a[0] = b[0]
        got = ops.box_iou(dev(a), dev(b)).cpu().numpy()
        want = oracle.box_iou(a, b)
        nan = np.isnan(want)                                                   # 0/0: NaN on both sides (payload bits differ by ISA)
        assert nan.any() and np.array_equal(np.isnan(got), nan), (na, nb)
        assert np.array_equal(got.view(np.uint32)[~nan], want.view(np.uint32)[~nan]), (na, nb)
    assert ops.box_iou(dev(np.zeros((0, 4), np.float32)), dev(b)).shape == (0, 4099)
    assert ops.box_iou(dev(a), dev(np.zeros((0, 4), np.float32))).shape == (64, 0)


def test_iou_match_equals_reference_golden(ops, golden):
    g = golden["match"]
    cases = synth.match_cases()
    for name, (targets, preds) in cases.items():
        for thr in synth.MATCH_THRESHOLDS:
            tp, used = ops.iou_match(dev(preds), dev(targets), thr)
            assert np.array_equal(tp.cpu().numpy().astype(np.uint8), g[name + "/tp_%g" % thr]), (name, thr)
            assert np.array_equal(used.cpu().numpy(), oracle.iou_match(preds, targets, thr)[1]), (name, thr)
    # the reference test's own known answer (test/metrics_test.py:61-68)
    tp, _ = ops.iou_match(dev(cases["kat0"][1]), dev(cases["kat0"][0]), 0.65)
    assert tp.int().tolist() == [1, 0, 0, 1, 0]
    # all images in ONE launch (CSR offsets), plus an image without predictions and one without targets
    names = list(cases)
    rows = [cases[n][1] for n in names] + [np.zeros((0, 4), np.float32), cases["kat1"][1]]
    cols = [cases[n][0] for n in names] + [cases["kat1"][0], np.zeros((0, 4), np.float32)]
    ro = np.cumsum([0] + [len(r) for r in rows]).astype(np.int32)
    co = np.cumsum([0] + [len(c) for c in cols]).astype(np.int32)
    for thr in synth.MATCH_THRESHOLDS:
        tp, used = ops.iou_match(dev(np.concatenate(rows)), dev(np.concatenate(cols)), thr, row_offsets=dev(ro), col_offsets=dev(co),
                                 max_cols=max(len(c) for c in cols))
        tp, used = tp.cpu().numpy(), used.cpu().numpy()
        for i, n in enumerate(names):
            assert np.array_equal(tp[ro[i]:ro[i + 1]].astype(np.uint8), g[n + "/tp_%g" % thr]), (n, thr)
        assert not tp[ro[-2]:].any() and not used[co[-3]:].any()


def test_iou_match_planogram_labels_equal_reference_golden(ops, golden):
    from cvpce_b200 import planograms
    g = golden["match"]
    for seed in (51, 52):
        b1, l1, b2, l2, _ = synth.planogram_case(seed)
        for thr in (0.5, 0.8):
            key = "planogram%d_%g" % (seed, thr)
            exp = torch.from_numpy(g[key + "/expected_positions"])
            me, ma, missing = planograms.match_expected(exp, torch.from_numpy(b2), l1, l2, thr)   # host tensors, like the reference
            assert not me.is_cuda
            assert np.array_equal(me.numpy().astype(np.uint8), g[key + "/matched_expected"]), key
            assert np.array_equal(ma.numpy().astype(np.uint8), g[key + "/matched_actual"]), key
            assert np.array_equal(missing.numpy(), g[key + "/missing_expected"]), key


def test_iou_match_edges_and_large_problem(ops):
    # fp32 threshold comparison, NaN claims, no stop after the first hit (see tests/test_match_golden.py for the reference facts)
    tp, _ = ops.iou_match(dev(np.array([[0, 0, 10, 10]], np.float32)), dev(np.array([[0, 0, 7, 10]], np.float32)), 0.7)
    assert tp.tolist() == [True]
    z = dev(np.array([[5, 5, 5, 5]], np.float32))
    tp, used = ops.iou_match(z, z, 0.5)
    assert tp.tolist() == [True] and used.tolist() == [True]
    rows = dev(np.array([[0, 0, 10, 10], [0, 0, 10, 10]], np.float32))
    cols = dev(np.array([[0, 0, 10, 10], [0, 0, 10, 9]], np.float32))
    tp, used = ops.iou_match(rows, cols, 0.5)
    assert tp.tolist() == [True, False] and used.tolist() == [True, True]
    tp, used = ops.iou_match(dev(np.zeros((0, 4), np.float32)), cols, 0.5)
    assert tp.numel() == 0 and used.tolist() == [False, False]
    tp, used = ops.iou_match(rows, dev(np.zeros((0, 4), np.float32)), 0.5)
    assert tp.tolist() == [False, False] and used.numel() == 0
    # several panels of rows, several words of columns, labels, dense overlaps
    rs = np.random.RandomState(9)
    for n_rows, n_cols, n_labels, thr in ((5000, 3000, 0, 0.3), (777, 5001, 6, 0.2), (33, 16384, 0, 0.1), (2000, 31, 3, 0.5)):
        r, _ = synth.random_boxes(rs.randint(1 << 30), n_rows, extent=600.0)
        c, _ = synth.random_boxes(rs.randint(1 << 30), n_cols, extent=600.0)
        rl = cl = None
        if n_labels:
            rl, cl = rs.randint(n_labels, size=n_rows).astype(np.int64), rs.randint(n_labels, size=n_cols).astype(np.int64)
        tp, used = ops.iou_match(dev(r), dev(c), thr, None if rl is None else dev(rl), None if cl is None else dev(cl))
        want_tp, want_used = oracle.iou_match(r, c, thr, rl, cl)
        assert np.array_equal(tp.cpu().numpy(), want_tp) and np.array_equal(used.cpu().numpy(), want_used), (n_rows, n_cols)
        assert 0 < want_tp.sum() and 0 < want_used.sum() < n_cols
    with pytest.raises(RuntimeError):
        ops.iou_match(rows, dev(np.zeros((16385, 4), np.float32)), 0.5)
    with pytest.raises(RuntimeError):
        ops.iou_match(rows.cpu(), cols, 0.5)


def test_metrics_mirror_process_batch(ops, golden):
    from cvpce_b200 import metrics
    g = golden["match"]
    cases = synth.match_cases()
    names = list(cases)
    rs = np.random.RandomState(3)
    targets, preds, confs, perms = [], [], [], []
    for n in names:                      # hand the predictions over UNSORTED with their confidences, like detection_eval.py does
        t, p = cases[n]
        conf = np.sort(rs.uniform(size=len(p)).astype(np.float32))[::-1].copy()
        perm = rs.permutation(len(p))
        targets.append(torch.from_numpy(t))
        preds.append(torch.from_numpy(p[perm]))
        confs.append(torch.from_numpy(conf[perm]))
    res = metrics.process_batch(targets, preds, confs, synth.MATCH_THRESHOLDS)
    for n, (m, conf, n_t), t in zip(names, res, targets):
        assert n_t == len(t) and bool((conf[:-1] >= conf[1:]).all())
        for thr in synth.MATCH_THRESHOLDS:
            tp = m[thr]["true_positives"]
            assert tp.dtype == torch.float32 and not tp.is_cuda
            assert np.array_equal(tp.numpy().astype(np.uint8), g[n + "/tp_%g" % thr]), (n, thr)
            assert torch.equal(tp + m[thr]["false_positives"], torch.ones_like(tp))
            assert float(m[thr]["recall_300"]) == pytest.approx(float(tp[:300].sum()) / len(t))
    s, i = metrics.iou_matrices(targets[0], torch.from_numpy(cases[names[0]][1]))
    assert np.array_equal(s.numpy(), g[names[0] + "/sorted_iou"])
    one = metrics.process_one(targets[3], preds[3], confs[3], (0.5,))
    assert torch.equal(one[0][0.5]["true_positives"], res[3][0][0.5]["true_positives"])


def test_hot_path_step_host_overlapped_upload_equals_step(ops):
    """HotPath.step_host (pinned host images uploaded group by group on a copy stream, crops cut per group between trunk
    batches) returns exactly what HotPath.step returns on device-resident inputs -- repeatedly, with the staging buffers
    reused, with an image without detections in the middle and at the end, for fp32 and the channels-last half hand-off."""
    from cvpce_b200 import classification, pipeline
    torch.backends.cudnn.allow_tf32 = False
    H, W = 384, 512
    n_img = 7
    images, boxes, scores, seg = [], [], [], [0]
    for i in range(n_img):
        images.append(torch.from_numpy(synth.shelf_image(20 + i, H, W)))
        if i in (3, 6):
            b, s = np.zeros((0, 4), np.float32), np.zeros((0,), np.float32)
        else:
            b, s = synth.shelf_candidates(20 + i, H, W, n_products=5 + 2 * i)
            b = np.clip(b, 0, [W, H, W, H]).astype(np.float32)
        boxes.append(b)
        scores.append(s)
        seg.append(seg[-1] + len(s))
    boxes, scores, seg = np.concatenate(boxes), np.concatenate(scores), np.array(seg, dtype=np.int32)
    max_n = int(np.diff(seg).max())
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).cuda()
    gallery = dev(synth.unit_rows(np.random.RandomState(2), 300, 1024))
    images_h = [im.pin_memory() for im in images]
    bh, sh, gh = torch.from_numpy(boxes).pin_memory(), torch.from_numpy(scores).pin_memory(), torch.from_numpy(seg).pin_memory()
    for kwargs in (dict(), dict(trunk_dtype=torch.float16, channels_last=True)):
        if kwargs:
            enc = enc.half().to(memory_format=torch.channels_last)
            enc.enable_fast_trunk()
        hp = pipeline.HotPath(enc, gallery, k=3, size=64, trunk_batch=8, detections_per_img=200, **kwargs)
        want = hp.step([im.cuda() for im in images], dev(boxes), dev(scores), dev(seg), max_n)
        assert int(want["counts"].sum()) > 20 and int(want["counts"][3]) == 0
        for group in (1, 2, 4, 16):
            for _ in range(2):
                got = hp.step_host(images_h, bh, sh, gh, max_n, upload_group=group)
                assert torch.equal(got["counts"], want["counts"])
                assert torch.equal(got["boxes"], want["boxes"])
                assert torch.equal(got["embedding"], want["embedding"]), (kwargs, group)
                assert torch.equal(got["topk_idx"], want["topk_idx"]) and torch.equal(got["topk_sim"], want["topk_sim"])
        # prefetch: the next step's images (a DIFFERENT set: the order reversed) go up into the second staging buffer before this
        # step runs; three steps in flight over two buffers
        rev_h = images_h[::-1]
        rev_seg = np.concatenate([[0], np.cumsum(np.diff(seg)[::-1])]).astype(np.int32)
        rev_boxes = np.concatenate([boxes[seg[i]:seg[i + 1]] for i in range(n_img - 1, -1, -1)])
        rev_scores = np.concatenate([scores[seg[i]:seg[i + 1]] for i in range(n_img - 1, -1, -1)])
        want_rev = hp.step([im.cuda() for im in images[::-1]], dev(rev_boxes), dev(rev_scores), dev(rev_seg), max_n)
        up_a = hp.upload_images(images_h, 2)
        up_b = hp.upload_images(rev_h, 2)
        got_a = hp.step(up_a[0], dev(boxes), dev(scores), dev(seg), max_n, uploads=up_a[1])
        up_c = hp.upload_images(images_h, 3)      # reuses the first buffer: waits for step A's crops
        got_b = hp.step(up_b[0], dev(rev_boxes), dev(rev_scores), dev(rev_seg), max_n, uploads=up_b[1])
        got_c = hp.step(up_c[0], dev(boxes), dev(scores), dev(seg), max_n, uploads=up_c[1])
        for got, w in ((got_a, want), (got_b, want_rev), (got_c, want)):
            assert torch.equal(got["boxes"], w["boxes"]) and torch.equal(got["embedding"], w["embedding"]) and torch.equal(got["topk_idx"], w["topk_idx"])
    # a step with no detections at all releases the staging buffers too
    empty = hp.step_host(images_h, bh[:0], sh[:0], torch.zeros(n_img + 1, dtype=torch.int32).pin_memory(), 1, upload_group=2)
    assert int(empty["counts"].sum()) == 0 and any(e is not None for e in hp._stage_free)


def test_patch_layer_on_a_reference_shaped_package(ops):
    """The reference itself cannot travel to the GPU box, so the patcher is exercised on a package with the reference's
    attribute layout (the real reference is patched and reverted in tests/test_reference_pin.py): after patch.apply the
    swapped call sites run on the GPU and agree with the torch / torchvision functions they replace."""
    import types
    import torchvision
    from torchvision.models.detection import retinanet
    from cvpce_b200 import patch

    def ref_nn(anchors, queries, k=1):
        d = 1 - torch.nn.functional.cosine_similarity(anchors[None, :, :], queries[:, None, :], dim=2)
        return d.argsort(dim=1)[:, :k]

    def ref_resize(img):
        _, h, w = img.shape
        L = max(h, w)
        canvas = torch.full((3, L, L), 0.5, dtype=img.dtype, device=img.device)
        canvas[:, :h, :w] = img
        return torch.nn.functional.interpolate(canvas[None], size=(256, 256), mode="bilinear", antialias=True, align_corners=False)[0]

    class _Enc(torch.nn.Module):
        def forward(self, x, eps=1e-8):
            return x

    class _PG:
        def generate_proposals_and_images(self, image):
            raise NotImplementedError

    class _Clf:
        def classify(self, images, return_embedding=False):
            raise NotImplementedError

    seen = {}

    def do_calculate(thresholds, matches, confidences, total_targets):
        seen.update(thresholds=thresholds, matches=matches, confidences=confidences, total=total_targets)
        return "bookkeeping result"

    pkg = types.SimpleNamespace(
        models=types.SimpleNamespace(classification=types.SimpleNamespace(nearest_neighbors=ref_nn, MACVGG=type("MACVGG", (_Enc,), {}),
                                                                          MACResNet=type("MACResNet", (_Enc,), {}))),
        production=types.SimpleNamespace(nearest_neighbors=ref_nn, ProposalGenerator=_PG, Classifier=_Clf),
        datautils=types.SimpleNamespace(resize_for_classification=ref_resize),
        utils=types.SimpleNamespace(nearest_neighbors=ref_nn),
        metrics=types.SimpleNamespace(iou_matrices=ref_nn, _process_one=ref_nn, calculate_metrics=ref_nn, _do_calculate=do_calculate),
        planograms=types.SimpleNamespace(tvops=torchvision.ops))
    tv_batched_nms = retinanet.box_ops.batched_nms
    patch.apply(pkg)
    try:
        # a11: the reference's own known answer (test/models/classification_test.py:8-25) through every alias
        anchors, queries, expected = synth.nn_kat()
        for mod in (pkg.models.classification, pkg.production, pkg.utils):
            assert mod.nearest_neighbors(dev(anchors), dev(queries)).flatten().tolist() == list(expected)
        # a1: torchvision's RetinaNet call site
        b, s = synth.random_boxes(77, 900, extent=300.0)
        labels = torch.zeros(900, dtype=torch.int64, device="cuda")
        assert retinanet.box_ops.batched_nms is not tv_batched_nms
        assert torch.equal(retinanet.box_ops.batched_nms(dev(b), dev(s), labels, 0.5), tv_batched_nms(dev(b), dev(s), labels, 0.5))
        # a4: same signature, CUDA image in
        img = dev(synth.shelf_image(8, 150, 90))
        assert float((pkg.datautils.resize_for_classification(img) - ref_resize(img)).abs().max()) <= 1e-5
        # ... and the two shared helpers hand HOST tensors (dataset workers, other detectors) to the original functions
        assert torch.equal(pkg.datautils.resize_for_classification(img.cpu()), ref_resize(img.cpu()))
        assert torch.equal(retinanet.box_ops.batched_nms(torch.from_numpy(b), torch.from_numpy(s), labels.cpu(), 0.5),
                           tv_batched_nms(torch.from_numpy(b), torch.from_numpy(s), labels.cpu(), 0.5))
        # f3: matching on the GPU, bookkeeping handed to the package's own _do_calculate; box_iou behind the tvops name
        cases = synth.match_cases()
        t = [torch.from_numpy(cases[n][0]) for n in ("kat0", "kat1", "kat2")]
        p = [torch.from_numpy(cases[n][1]) for n in ("kat0", "kat1", "kat2")]
        conf = [torch.linspace(1, 0.1, len(x)) for x in p]
        assert pkg.metrics.calculate_metrics(t, p, conf, (0.5, 0.65)) == "bookkeeping result"
        assert seen["total"] == 9 and seen["matches"][0.65]["true_positives"][0].tolist() == [1, 0, 0, 1, 0]
        assert [float(r) for r in seen["matches"][0.5]["recall_300"]] == pytest.approx([1, 3 / 4, 1 / 2])   # test/metrics_test.py:70-78
        got = pkg.planograms.tvops.box_iou(p[0], t[0])
        assert not got.is_cuda and torch.equal(got, torchvision.ops.box_iou(p[0], t[0]))
        assert pkg.planograms.tvops.nms is torchvision.ops.nms
    finally:
        patch.revert()
    assert retinanet.box_ops.batched_nms is tv_batched_nms and pkg.models.classification.nearest_neighbors is ref_nn
    assert pkg.planograms.tvops is torchvision.ops and pkg.metrics.calculate_metrics is ref_nn


def test_nms_and_box_iou_on_degenerate_integer_grids(ops):
    """Small integer coordinates: exact score ties, IoUs exactly at the threshold, zero-area and nested boxes -- the same
    generator on which tests/test_oracle_golden.py pins the oracle to the compiled torchvision ops."""
    rs = np.random.RandomState(2025)
    for thr in (0.25, 1 / 3, 0.5, 0.7):
        boxes, scores, seg = [], [], [0]
        for case in range(40):
            n = int(rs.choice([0, 1, 2, 7, 40, 150, 700]))
            xy = rs.randint(0, 12, size=(n, 2)).astype(np.float32)
            wh = rs.randint(0, 7, size=(n, 2)).astype(np.float32)
            boxes.append(np.concatenate([xy, xy + wh], axis=1))
            scores.append((rs.randint(0, 8, size=n) / 8.0).astype(np.float32))
            seg.append(seg[-1] + n)
        b, s = np.concatenate(boxes), np.concatenate(scores)
        keep, counts, _, _ = ops.nms_batched(dev(b), dev(s), dev(np.array(seg, np.int32)), 700, thr)
        keep, counts = keep.cpu().numpy(), counts.cpu().numpy()
        for i in range(40):
            want = oracle.nms(boxes[i], scores[i], thr)
            assert counts[i] == len(want) and np.array_equal(keep[seg[i]:seg[i] + counts[i]], want), (thr, i, len(scores[i]))
        got = ops.box_iou(dev(boxes[5]), dev(boxes[6])).cpu().numpy() if len(boxes[5]) and len(boxes[6]) else None
        if got is not None:
            want = oracle.box_iou(boxes[5], boxes[6])
            nan = np.isnan(want)
            assert np.array_equal(np.isnan(got), nan) and np.array_equal(got[~nan], want[~nan])


def test_crops_equal_the_reference_torch_cpu_ops_on_random_boxes(ops):
    """The crop kernel against the reference's own torch CPU ops (python slice + pad 0.5 + F.interpolate(antialias=True),
    oracle/torch_port.py, pinned to the live reference in tests/test_reference_pin.py) on random boxes: up- and down-sampling
    by up to ~5x, thin strips, boxes touching and crossing the image border, fractional coordinates."""
    from oracle import torch_port
    H, W = 700, 900
    img = synth.shelf_image(12, H, W)
    rs = np.random.RandomState(77)
    xy = rs.uniform(-20, [W - 10, H - 10], size=(48, 2))
    wh = np.exp(rs.uniform(np.log(6), np.log(680), size=(48, 2)))
    boxes = np.concatenate([np.maximum(xy, 0), xy + wh], axis=1).astype(np.float32)
    boxes[:, 2] = np.maximum(boxes[:, 2], boxes[:, 0] + 2)
    boxes[:, 3] = np.maximum(boxes[:, 3], boxes[:, 1] + 2)
    boxes[:, 0] = np.minimum(boxes[:, 0], W - 3)
    boxes[:, 1] = np.minimum(boxes[:, 1], H - 3)
    want = torch_port.crops_for_boxes(torch.from_numpy(img), torch.from_numpy(boxes)).numpy()
    table = ops.make_image_table([dev(img)])
    got = ops.crop_resize(table, 1, dev(boxes), torch.zeros(len(boxes), dtype=torch.int32, device="cuda"), 256).cpu().numpy()
    assert got.shape == want.shape
    err = np.abs(got - want).reshape(len(boxes), -1).max(axis=1)
    assert err.max() <= 1e-5, (int(err.argmax()), float(err.max()), boxes[int(err.argmax())])


def test_hot_path_end_to_end_equals_the_reference_torch_cpu_path(ops):
    """One shelf through HotPath.step on the GPU (NMS -> crops -> VGG16 -> MAC -> cosine top-k, S = 256, fp32 convolutions)
    against the same shelf through the reference's torch CPU ops in the reference's order (oracle/torch_port.hot_path:
    torchvision nms, python crop loop + F.interpolate, VGG16, amax / cat / norm, the gather-and-argsort nearest_neighbors)."""
    from oracle import torch_port
    from cvpce_b200 import classification, pipeline
    torch.backends.cudnn.allow_tf32 = False
    H, W = 420, 560
    img = synth.shelf_image(14, H, W)
    boxes, scores = synth.shelf_candidates(14, H, W, n_products=9)
    boxes = np.clip(boxes, 0, [W, H, W, H]).astype(np.float32)
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    port = torch_port.MACVGGPort().eval().requires_grad_(False)
    port.block1.load_state_dict(enc.block1.state_dict())
    port.block2.load_state_dict(enc.block2.state_dict())
    gallery = torch.from_numpy(synth.unit_rows(np.random.RandomState(4), 257, 1024))
    want_boxes, want_idx = torch_port.hot_path(port, gallery, torch.from_numpy(img), torch.from_numpy(boxes), torch.from_numpy(scores), k=3)
    hp = pipeline.HotPath(enc.cuda(), gallery.cuda(), k=3, size=256, trunk_batch=4, detections_per_img=200)
    out = hp.step([dev(img)], dev(boxes), dev(scores), dev(np.array([0, len(scores)], np.int32)), len(scores))
    assert torch.equal(out["boxes"].cpu(), want_boxes)                                  # NMS: bit-exact keep list, in order
    K = want_boxes.shape[0]
    assert K >= 8 and out["topk_idx"].shape == (K, 3)
    # embeddings of the CPU path for the tolerance check and the tie-aware index comparison
    crops = torch_port.crops_for_boxes(torch.from_numpy(img), want_boxes)
    with torch.no_grad():
        emb = port(crops * 2 - 1)                                                        # utils.scale_to_tanh, production.py:63
    np.testing.assert_allclose(out["embedding"].cpu().numpy(), emb.numpy(), rtol=1e-3, atol=2e-5)
    sims = (emb.double() @ gallery.double().T)
    top = sims.sort(dim=1, descending=True)
    clear = (top.values[:, :3] - top.values[:, 1:4]) > 1e-5                              # ranks decided by more than rounding
    got = out["topk_idx"].cpu()
    assert torch.equal(got[clear], want_idx[clear]) and torch.equal(want_idx[clear], top.indices[:, :3][clear])
    assert float(clear.float().mean()) > 0.9


def test_four_channel_crops_and_padded_first_convolution(ops):
    """NHWC4 crops are the NHWC crops plus a zero 4th channel (every dtype, both kernels), the fast trunk on them gives the
    same feature maps bit for bit (the extra products are exact zeros), and HotPath uses them without changing any result."""
    from cvpce_b200 import classification, pipeline
    torch.backends.cudnn.allow_tf32 = False
    img = dev(synth.shelf_image(15, 300, 410))
    table = ops.make_image_table([img])
    boxes = dev(np.array([[3, 4, 200, 280], [0, 0, 410, 300], [100, 100, 101, 250], [50.5, 60.5, 90.2, 75.9], [7, 7, 7, 7]], np.float32))
    idx = torch.zeros(5, dtype=torch.int32, device="cuda")
    aff = classification.tanh_normalize_affine()
    for S in (256, 64, 37):                                                      # band kernel and direct kernel
        for dtype in (torch.float32, torch.bfloat16, torch.float16):
            c3 = ops.crop_resize(table, 1, boxes, idx, S, True, 0.5, aff, dtype, True)
            c4 = ops.crop_resize(table, 1, boxes, idx, S, True, 0.5, aff, dtype, True, pad_channel=True)
            assert c4.shape == (5, 4, S, S) and c4.is_contiguous(memory_format=torch.channels_last)
            assert torch.equal(c4[:, :3], c3) and not bool(c4[:, 3].any()), (S, dtype)
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).cuda()
    enc.enable_fast_trunk()
    c3 = ops.crop_resize(table, 1, boxes, idx, 64, True, 0.5, aff, torch.float32, True)
    c4 = ops.crop_resize(table, 1, boxes, idx, 64, True, 0.5, aff, torch.float32, True, pad_channel=True)
    with torch.no_grad():
        a4, a5 = enc.features(c3, normalized=True)
        b4, b5 = enc.features(c4, normalized=True)
    assert torch.equal(a4, b4) and torch.equal(a5, b5)
    with pytest.raises(RuntimeError):
        enc.features(c4, normalized=False)
    b, s = synth.shelf_candidates(15, 300, 410, n_products=8)
    b = np.clip(b, 0, [410, 300, 410, 300]).astype(np.float32)
    gallery = dev(synth.unit_rows(np.random.RandomState(2), 300, 1024))
    seg = dev(np.array([0, len(s)], np.int32))
    outs = []
    for pad in (True, False):
        hp = pipeline.HotPath(enc, gallery, k=3, size=64, trunk_batch=8, channels_last=True, pad_channel=pad)
        outs.append(hp.step([img], dev(b), dev(s), seg, len(s)))
    assert torch.equal(outs[0]["embedding"], outs[1]["embedding"]) and torch.equal(outs[0]["topk_idx"], outs[1]["topk_idx"])
