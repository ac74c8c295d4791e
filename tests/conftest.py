import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    return {name: np.load(os.path.join(GOLDEN, name + ".npz")) for name in ("nms", "crops", "mac", "nn", "classify", "head", "match")}


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def topk_check(q, g, got_idx, oracle_sim, oracle_idx, tol=2e-6):
    """Tie-aware comparison of top-k indices: every returned index must be distinct per row and its exact (float64)
    cosine similarity must equal the oracle's similarity at the same rank within `tol`; where the oracle's similarity
    is separated from its neighbours by more than `tol`, the index itself must match. Returns the exact-match rate."""
    q = np.asarray(q, dtype=np.float64)
    g = np.asarray(g, dtype=np.float64)
    got_idx = np.asarray(got_idx)
    assert got_idx.shape == oracle_idx.shape, (got_idx.shape, oracle_idx.shape)
    qn = q / np.maximum(np.linalg.norm(q, axis=1, keepdims=True), 1e-8)
    gn = g / np.maximum(np.linalg.norm(g, axis=1, keepdims=True), 1e-8)
    Q, k = got_idx.shape
    for r in range(Q):
        assert len(set(got_idx[r].tolist())) == k, "duplicate index in row %d: %s" % (r, got_idx[r])
    sims = np.einsum("qkd,qd->qk", gn[got_idx], qn)
    np.testing.assert_allclose(sims, oracle_sim, rtol=0, atol=tol)
    pad = np.full((Q, 1), np.inf)
    gap_prev = np.abs(np.diff(np.concatenate([pad, oracle_sim], 1), axis=1))
    gap_next = np.abs(np.diff(np.concatenate([oracle_sim, -pad], 1), axis=1))
    clear = (gap_prev > 4 * tol) & (gap_next > 4 * tol)
    assert np.array_equal(got_idx[clear], oracle_idx[clear]), "index mismatch at clearly separated ranks"
    return float((got_idx == oracle_idx).mean())


def rows_check(got_boxes, got_scores, ref_boxes, ref_scores, score_rtol=1e-6, box_atol=2e-3, window=3):
    """Near-tie-aware comparison of score-descending detection rows against torchvision's: same count, scores equal in
    order within `score_rtol` (ATen's vectorised sigmoid/exp and the fixed-sequence ones differ by a few ulp), and every
    row's box equal within `box_atol` px to the reference row at the same position or -- where neighbouring scores are
    within `score_rtol` of each other, i.e. their order is decided by the last ulp -- at a position <= `window` away."""
    got_boxes, ref_boxes = np.asarray(got_boxes, dtype=np.float64), np.asarray(ref_boxes, dtype=np.float64)
    got_scores, ref_scores = np.asarray(got_scores, dtype=np.float64), np.asarray(ref_scores, dtype=np.float64)
    assert got_scores.shape == ref_scores.shape, (got_scores.shape, ref_scores.shape)
    np.testing.assert_allclose(got_scores, ref_scores, rtol=score_rtol, atol=0)
    n = len(ref_scores)
    bad = np.nonzero(np.abs(got_boxes - ref_boxes).max(axis=1) > box_atol)[0] if n else []
    for i in bad:
        ok = False
        for j in range(max(0, i - window), min(n, i + window + 1)):
            if abs(ref_scores[j] - got_scores[i]) <= 2 * score_rtol * abs(ref_scores[j]) and \
                    np.abs(got_boxes[i] - ref_boxes[j]).max() <= box_atol:
                ok = True
                break
        assert ok, "row %d differs from the reference beyond a near-tie swap: %s vs %s" % (i, got_boxes[i], ref_boxes[i])
    return len(bad)
