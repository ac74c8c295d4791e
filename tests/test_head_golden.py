"""CPU: the oracle's pre-NMS candidate selection + NMS (oracle.c: oracle_select_level, oracle_nms) against
RetinaNet.postprocess_detections of the reference's GaussianLayerNetwork (tests/golden/head.npz, oracle/make_golden.py:
golden_head).  SURVEY §8 f1."""
import numpy as np
import pytest

import oracle
from cvpce_b200 import synth
from conftest import rows_check

GLN_DETECTIONS_PER_IMG = 1000   # proposals.py:163-166
SCORE_THRESH = 0.05             # tv:retinanet.py RetinaNet.__init__ default, kept by proposals.gln


def test_det_sigmoid_is_within_ulps_of_torch():
    import torch
    x = np.concatenate([np.linspace(-100, 90, 4001), np.random.RandomState(0).standard_normal(4000) * 4]).astype(np.float32)
    mine = oracle.det_sigmoid(x).astype(np.float64)
    ref = torch.sigmoid(torch.from_numpy(x)).numpy().astype(np.float64)
    ok = ref > 1e-37
    assert (np.abs(mine[ok] - ref[ok]) / ref[ok]).max() < 4e-7


@pytest.mark.parametrize("name", [n for n in synth.HEAD_CASES if n != "ties"])
def test_candidates_match_torchvision(golden, name):
    logits, regression, anchors, sizes, C, topk = synth.head_case(name)
    res = oracle.select_candidates(logits, regression, anchors, sizes, SCORE_THRESH, topk)
    g = golden["head"]
    for b, (boxes, scores, labels, counts) in enumerate(res):
        pre = "%s/%d/" % (name, b)
        if pre + "cand_count" in g.files:
            assert len(scores) == int(g[pre + "cand_count"][0])
            assert abs(scores.astype(np.float64).sum() - float(g[pre + "cand_score_sum"][0])) < 1e-6 * len(scores)
            continue
        # levels are concatenated, each in descending score order: compare level by level
        assert len(scores) == len(g[pre + "cand_scores"])
        o = 0
        for c in counts:
            rows_check(boxes[o:o + c], scores[o:o + c], g[pre + "cand_boxes"][o:o + c], g[pre + "cand_scores"][o:o + c])
            if C == 1:
                assert np.array_equal(labels[o:o + c], g[pre + "cand_labels"][o:o + c])
            o += c


def test_ties_case_matches_torchvision_as_multiset(golden):
    """Equal scores: torch.topk's order (and, at the k-th score, its choice) is unspecified; the score sequence is not."""
    logits, regression, anchors, sizes, C, topk = synth.head_case("ties")
    boxes, scores, labels, counts = oracle.select_candidates(logits, regression, anchors, sizes, SCORE_THRESH, topk)[0]
    np.testing.assert_allclose(scores, golden["head"]["ties/0/cand_scores"], rtol=1e-6)
    assert counts[0] == topk and len(np.unique(scores[:topk])) < 12   # the first level really is tie-dominated


@pytest.mark.parametrize("name", ["dense_2img", "sparse", "fullsize"])
def test_detections_match_reference_gln_postprocess(golden, name):
    logits, regression, anchors, sizes, C, topk = synth.head_case(name)
    res = oracle.postprocess_detections(logits, regression, anchors, sizes, SCORE_THRESH, topk, 0.5, GLN_DETECTIONS_PER_IMG)
    for b, (boxes, scores, labels) in enumerate(res):
        pre = "%s/%d/" % (name, b)
        rows_check(boxes, scores, golden["head"][pre + "det_boxes"], golden["head"][pre + "det_scores"])
        assert np.array_equal(labels, golden["head"][pre + "det_labels"])


def test_selection_edge_cases():
    logits, regression, anchors, sizes, C, topk = synth.head_case("sparse")
    # threshold above every score: nothing selected
    res = oracle.select_candidates(logits, regression, anchors, sizes, 0.999999, topk)
    assert all(len(r[1]) == 0 for r in res)
    # topk = 1: the single best candidate of every level that has one
    res = oracle.select_candidates(logits, regression, anchors, sizes, SCORE_THRESH, 1)
    assert res[0][3].max() <= 1
    # NaN logits never pass (sigmoid(NaN) > thr is false)
    lg = [l.copy() for l in logits]
    lg[0][:] = np.nan
    res = oracle.select_candidates(lg, regression, anchors, sizes, SCORE_THRESH, topk)
    assert res[0][3][0] == 0
