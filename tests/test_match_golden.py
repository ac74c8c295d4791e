"""CPU: the oracle's box IoU / greedy matching against the reference's own known answers (test/metrics_test.py)
and the golden vectors recorded from the real reference (metrics.iou_matrices / check_matches, planograms.finalize_via_ransac)."""
import numpy as np

import oracle
from cvpce_b200 import synth


def test_iou_known_answers_of_the_reference_tests():
    # test/metrics_test.py:23-40 (test_iou_matrices): IoUs of PREDICTIONS[2] against TARGETS[2], sorted descending per row
    t, p = (np.array(x, dtype=np.float32) for x in (synth.METRICS_KAT_TARGETS[2], synth.METRICS_KAT_PREDICTIONS[2]))
    iou = oracle.box_iou(p, t)
    expected_sorted = np.array([[0.04, 0], [0.16, 0], [0.64, 0], [25 / 36, 1 / (25 + 36 - 1)], [0.36, 0]], dtype=np.float32)
    np.testing.assert_allclose(-np.sort(-iou, axis=1), expected_sorted, rtol=1e-5, atol=1e-8)
    assert np.argsort(-iou, axis=1, kind="stable").tolist() == [[0, 1], [0, 1], [0, 1], [0, 1], [1, 0]]
    # test/metrics_test.py:42-59 (test_iou_matrices_2)
    t, p = (np.array(x, dtype=np.float32) for x in (synth.METRICS_KAT_TARGETS[0], synth.METRICS_KAT_PREDICTIONS[0]))
    iou = oracle.box_iou(p, t)
    expected_sorted = np.array([[0.81, 0, 0], [0.64, 0, 0], [1, 0, 0],
                                [1 / 1.44, 0.1 / (1.44 + 1 - 0.1), 0.01 / (1.44 + 1 - 0.01)], [0, 0, 0]], dtype=np.float32)
    np.testing.assert_allclose(-np.sort(-iou, axis=1), expected_sorted, rtol=1e-5, atol=1e-7)


def test_check_matches_known_answer_of_the_reference_tests():
    # test/metrics_test.py:61-68 (test_check_matches): threshold 0.65 on case 0 -> true positives [1, 0, 0, 1, 0]
    t, p = (np.array(x, dtype=np.float32) for x in (synth.METRICS_KAT_TARGETS[0], synth.METRICS_KAT_PREDICTIONS[0]))
    tp, _ = oracle.iou_match(p, t, 0.65)
    assert tp.astype(int).tolist() == [1, 0, 0, 1, 0]
    # test/metrics_test.py:70-78 (tps_fps): recall_300 1, 3/4, 1/2 at the default 0.5 <=> 3, 3, 1 matched targets
    for i, n_used in enumerate((3, 3, 1)):
        t, p = (np.array(x, dtype=np.float32) for x in (synth.METRICS_KAT_TARGETS[i], synth.METRICS_KAT_PREDICTIONS[i]))
        tp, used = oracle.iou_match(p, t, 0.5)
        assert int(tp.sum()) == n_used


def test_iou_and_matches_equal_reference_golden(golden):
    g = golden["match"]
    for name, (targets, preds) in synth.match_cases().items():
        iou = oracle.box_iou(preds, targets)
        assert np.array_equal(iou.view(np.uint32), g[name + "/iou"].view(np.uint32)), name       # bit-equal to torchvision
        assert np.array_equal(-np.sort(-iou, axis=1), g[name + "/sorted_iou"]), name
        for thr in synth.MATCH_THRESHOLDS:
            tp, _ = oracle.iou_match(preds, targets, thr)
            assert np.array_equal(tp.astype(np.uint8), g[name + "/tp_%g" % thr]), (name, thr)


def label_ids(l1, l2):
    key = {l: i for i, l in enumerate(sorted(set(l1) | set(l2)))}
    return np.array([key[l] for l in l1], dtype=np.int64), np.array([key[l] for l in l2], dtype=np.int64)


def test_planogram_matching_equals_reference_golden(golden):
    g = golden["match"]
    for seed in (51, 52):
        b1, l1, b2, l2, _ = synth.planogram_case(seed)
        i1, i2 = label_ids(l1, l2)
        for thr in (0.5, 0.8):
            key = "planogram%d_%g" % (seed, thr)
            expected_positions = g[key + "/expected_positions"]
            assert expected_positions.shape == b1.shape
            me, ma = oracle.iou_match(expected_positions, b2, thr, i1, i2)
            assert np.array_equal(me.astype(np.uint8), g[key + "/matched_expected"]), key
            assert np.array_equal(ma.astype(np.uint8), g[key + "/matched_actual"]), key
            assert np.array_equal(np.nonzero(~me)[0], g[key + "/missing_expected"]), key


def test_threshold_is_compared_in_fp32_and_nan_claims():
    # IoU == fp32(0.7) < 0.7 (double): the reference compares a fp32 tensor element with the Python float in fp32 => a match
    a = np.array([[0, 0, 10, 10]], dtype=np.float32)
    b = np.array([[0, 0, 7, 10]], dtype=np.float32)
    assert oracle.box_iou(a, b)[0, 0] == np.float32(0.7)
    tp, _ = oracle.iou_match(a, b, 0.7)
    assert tp.tolist() == [True]
    # two empty boxes: 0/0 = NaN is not "< thr", sorts first, and claims
    z = np.array([[5, 5, 5, 5]], dtype=np.float32)
    tp, used = oracle.iou_match(z, z, 0.5)
    assert np.isnan(oracle.box_iou(z, z)[0, 0]) and tp.tolist() == [True] and used.tolist() == [True]
    # a row marks EVERY unmarked column above the threshold (no stop after the first hit): the second row finds nothing left
    rows = np.array([[0, 0, 10, 10], [0, 0, 10, 10]], dtype=np.float32)
    cols = np.array([[0, 0, 10, 10], [0, 0, 10, 9]], dtype=np.float32)
    tp, used = oracle.iou_match(rows, cols, 0.5)
    assert tp.tolist() == [True, False] and used.tolist() == [True, True]
