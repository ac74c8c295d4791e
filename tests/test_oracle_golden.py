"""CPU: the oracle (oracle/oracle.c) against the golden vectors produced by the real reference (oracle/make_golden.py)."""
import numpy as np

import oracle
from cvpce_b200 import synth
from conftest import topk_check


def test_nms_matches_torchvision_golden(golden):
    for name, (boxes, scores, thr, per_img) in synth.nms_cases().items():
        keep = oracle.nms(boxes, scores, thr)
        assert np.array_equal(keep, golden["nms"][name + "/keep"]), name
        key = name + "/keep_postprocess"
        if key in golden["nms"].files:
            assert np.array_equal(oracle.postprocess_nms(boxes, scores, thr, per_img), golden["nms"][key]), name


def test_nms_known_answers():
    b, s, thr, _ = synth.nms_cases()["ties"]
    assert oracle.nms(b, s, thr).tolist() == [3, 0]           # stable: equal scores keep the lower index first
    b, s, thr, _ = synth.nms_cases()["zero_area"]
    assert oracle.nms(b, s, thr).tolist() == [0, 1, 2, 3]     # 0/0 = NaN does not suppress
    b, s, thr, _ = synth.nms_cases()["hand_iou"]
    assert oracle.nms(b, s, thr).tolist() == [0, 2, 4]        # IoU .81 suppressed, .64 kept, then 1/1.44=.694 vs box 0


def test_nms_threshold_is_compared_as_double():
    # IoU == fp32(1/3) exactly; the CPU op compares against the DOUBLE 1/3 (< fp32(1/3)) => suppressed
    b = np.array([[0, 0, 2, 1], [1, 0, 3, 1]], dtype=np.float32)
    s = np.array([0.9, 0.8], dtype=np.float32)
    assert oracle.nms(b, s, 1 / 3).tolist() == [0]
    assert oracle.nms(b, s, float(np.float32(1 / 3))).tolist() == [0, 1]


def test_crops_match_reference_golden(golden):
    img = synth.shelf_image(synth.CROP_IMAGE_SEED, *synth.CROP_IMAGE_HW)
    boxes = synth.crop_boxes()
    for key, S, aa in (("aa_S256", 256, True), ("aa_S224", 224, True), ("noaa_S256", 256, False)):
        out, valid = oracle.crop_resize(img, boxes, S, aa)
        assert valid.all()
        got = out.reshape(len(boxes), -1)[:, synth.sample_pixels(S)]
        np.testing.assert_allclose(got, golden["crops"][key], rtol=0, atol=1e-6)


def test_crop_int_truncation_and_empty():
    img = synth.shelf_image(1, 40, 50)
    a, _ = oracle.crop_resize(img, np.array([[10.7, 20.2, 20.9, 35.5]], dtype=np.float32), 32)
    b, _ = oracle.crop_resize(img, np.array([[10, 20, 20, 35]], dtype=np.float32), 32)
    assert np.array_equal(a, b)
    out, valid = oracle.crop_resize(img, np.array([[10, 10, 10, 10], [5, 5, 5, 30]], dtype=np.float32), 16)
    assert valid.tolist() == [False, True]
    np.testing.assert_allclose(out, 0.5, atol=1e-6)  # zero-width crop: the whole canvas is padding


def test_mac_matches_reference_golden(golden):
    f4, f5 = synth.feature_maps()
    out = oracle.region_pool_embed([f4, f5])
    np.testing.assert_allclose(out, golden["mac"]["mac_given_maps"], rtol=0, atol=1e-6)
    raw = oracle.region_pool_embed([f4, f5], final_l2=False)
    assert np.array_equal(raw, np.concatenate([f4.max(axis=(2, 3)), f5.max(axis=(2, 3))], 1))  # maxima are exact


def test_rmac_regions_and_embedding():
    r = oracle.rmac_regions(16, 16, 3)
    assert len(r) == 1 + 4 + 9 and r[0].tolist() == [0, 0, 16, 16]
    assert all(0 <= y0 < y1 <= 16 and 0 <= x0 < x1 <= 16 for y0, x0, y1, x1 in r.tolist())
    f4, f5 = synth.feature_maps()
    regions = [oracle.rmac_regions(f.shape[2], f.shape[3], 3) for f in (f4, f5)]
    out = oracle.region_pool_embed([f4, f5], regions, per_region_l2=True)
    np.testing.assert_allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-6)
    # numpy restatement of Tolias et al.: per-region max -> L2 -> sum, per map; concat; L2
    descs = []
    for f, rg in ((f4, regions[0]), (f5, regions[1])):
        d = 0
        for y0, x0, y1, x1 in rg:
            v = f[:, :, y0:y1, x0:x1].max(axis=(2, 3)).astype(np.float64)
            d = d + v / np.linalg.norm(v, axis=1, keepdims=True)
        descs.append(d)
    ref = np.concatenate(descs, 1)
    ref /= np.linalg.norm(ref, axis=1, keepdims=True)
    np.testing.assert_allclose(out, ref, atol=1e-6)


def test_nearest_neighbors_reference_kat(golden):
    a, q, expected = synth.nn_kat()
    assert np.array_equal(oracle.nearest_neighbors(a, q)[:, 0], expected)
    assert np.array_equal(golden["nn"]["kat_top1"], expected)


def test_nearest_neighbors_matches_reference_golden(golden):
    for name, (Q, G, D, k, seed) in synth.NN_CASES.items():
        g, q = synth.nn_case(Q, G, D, seed)
        sim, idx = oracle.sim_topk(q, g, k)
        rate = topk_check(q, g, golden["nn"][name], sim, idx)
        assert rate == 1.0, (name, rate)


def test_sim_topk_ties_and_padding():
    g = np.array([[1, 0], [1, 0], [0, 1], [1, 0]], dtype=np.float32)
    q = np.array([[1, 0]], dtype=np.float32)
    sim, idx = oracle.sim_topk(q, g, 3)
    assert idx.tolist() == [[0, 1, 3]]  # equal similarity: ascending index
    sim, idx = oracle.sim_topk(q, np.zeros((1, 2), dtype=np.float32), 1)
    assert sim[0, 0] == 0.0  # zero vector: norm clamped to eps => similarity 0 (distance 1)


def test_oracle_nms_and_box_iou_equal_the_compiled_torchvision_ops_on_random_cases():
    """The third-party ops the reference calls (torchvision.ops.nms / box_iou, CPU) are available wherever the tests run, so the
    restatement is also pinned on fresh random inputs: small integer coordinates make exact score ties, IoUs exactly at the
    threshold, zero-area and fully nested boxes common."""
    import torch
    import torchvision
    rs = np.random.RandomState(2024)
    for case in range(60):
        n = int(rs.choice([0, 1, 2, 7, 40, 150]))
        xy = rs.randint(0, 12, size=(n, 2)).astype(np.float32)
        wh = rs.randint(0, 7, size=(n, 2)).astype(np.float32)            # zero-width / zero-height boxes included
        boxes = np.concatenate([xy, xy + wh], axis=1)
        scores = (rs.randint(0, 8, size=n) / 8.0).astype(np.float32)     # many equal scores
        thr = float(rs.choice([0.25, 1 / 3, 0.5, 0.7]))
        want = torchvision.ops.nms(torch.from_numpy(boxes), torch.from_numpy(scores), thr).numpy()
        assert np.array_equal(oracle.nms(boxes, scores, thr), want), (case, n, thr)
        m = int(rs.choice([1, 3, 30]))
        other = np.concatenate([rs.randint(0, 12, size=(m, 2)), rs.randint(6, 16, size=(m, 2))], axis=1).astype(np.float32)
        want_iou = torchvision.ops.box_iou(torch.from_numpy(boxes), torch.from_numpy(other)).numpy()
        got_iou = oracle.box_iou(boxes, other)
        nan = np.isnan(want_iou)
        assert np.array_equal(np.isnan(got_iou), nan) and np.array_equal(got_iou[~nan], want_iou[~nan]), case
