"""CPU oracle of the cvpce proposal -> product classification path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may import this package;
nothing under cvpce_b200/ does.  `oracle.c` is the plain-C restatement (numpy/ctypes wrappers below); `torch_port.py`
is the torch-CPU restatement of the reference's own Python call sequence used as the timed CPU baseline;
`shim.py` + `make_golden.py` import the real reference (only in the authoring container, where /root/reference
exists) to pin both against it and to write tests/golden/.
"""
import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_nms.restype = ctypes.c_int64
        _lib.oracle_crop_resize.restype = ctypes.c_int
        _lib.oracle_rmac_regions.restype = ctypes.c_int
    return _lib


def _p(a, t=ctypes.c_void_p):
    return a.ctypes.data_as(t)


def _threads():
    return max(1, int(os.environ.get("ORACLE_THREADS", os.cpu_count() or 1)))


def _parallel_rows(n, fn):
    """Run fn(lo, hi) over row chunks on host threads (ctypes releases the GIL; there is no OpenMP in the image)."""
    t = min(_threads(), max(1, n))
    if t == 1:
        fn(0, n)
        return
    step = (n + t - 1) // t
    with ThreadPoolExecutor(max_workers=t) as ex:
        list(ex.map(lambda lo: fn(lo, min(n, lo + step)), range(0, n, step)))


def nms(boxes, scores, iou_threshold):
    """torchvision.ops.nms restated (oracle.c: oracle_nms). boxes f32[N,4], scores f32[N] -> int64[K]."""
    boxes = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 4)
    scores = np.ascontiguousarray(scores, dtype=np.float32).reshape(-1)
    n = boxes.shape[0]
    keep = np.empty(max(n, 1), dtype=np.int64)
    k = lib().oracle_nms(_p(boxes), _p(scores), ctypes.c_int64(n), ctypes.c_double(float(iou_threshold)), _p(keep))
    return keep[:k].copy()


def postprocess_nms(boxes, scores, iou_threshold=0.5, detections_per_img=1000, min_score=None):
    """batched_nms(single class) + keep[:detections_per_img] (tv:retinanet.py:559-561) + optional confidence filter
    (cvpce/production.py:15). Returns kept indices."""
    keep = nms(boxes, scores, iou_threshold)[:detections_per_img]
    if min_score is not None:
        keep = keep[np.asarray(scores, dtype=np.float32)[keep] > np.float32(min_score)]
    return keep


def crop_resize(image, boxes, S=256, antialias=True, pad=0.5):
    """production.py:20 crop loop + datautils.resize_for_classification. image f32[3,H,W]; boxes f32[K,4] ->
    (crops f32[K,3,S,S], valid bool[K])."""
    image = np.ascontiguousarray(image, dtype=np.float32)
    boxes = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 4)
    _, H, W = image.shape
    K = boxes.shape[0]
    out = np.empty((K, 3, S, S), dtype=np.float32)
    valid = np.ones(K, dtype=bool)
    L = lib()

    def run(lo, hi):
        for i in range(lo, hi):
            rc = L.oracle_crop_resize(_p(image), ctypes.c_int64(H), ctypes.c_int64(W), ctypes.c_int64(H * W), ctypes.c_int64(W),
                                      _p(boxes[i]), ctypes.c_int(S), ctypes.c_int(1 if antialias else 0), ctypes.c_float(pad),
                                      _p(out[i]))
            valid[i] = rc == 0

    _parallel_rows(K, run)
    return out, valid


def rmac_regions(H, W, levels):
    buf = np.zeros((64, 4), dtype=np.int32)
    n = lib().oracle_rmac_regions(int(H), int(W), int(levels), _p(buf), 64)
    return buf[:n].copy()


def region_pool_embed(feats, regions=None, per_region_l2=False, final_l2=True, eps=1e-8):
    """MACVGG.forward tail (classification.py:47-51): feats = list of f32[B,C,H,W]; regions = None (MAC) or a list of
    int32[R_m,4] per map (R-MAC extension). Returns f32[B, sum C]."""
    feats = [np.ascontiguousarray(f, dtype=np.float32) for f in feats]
    B = feats[0].shape[0]
    n = len(feats)
    Cs = np.array([f.shape[1] for f in feats], dtype=np.int32)
    Hs = np.array([f.shape[2] for f in feats], dtype=np.int32)
    Ws = np.array([f.shape[3] for f in feats], dtype=np.int32)
    ptrs = (ctypes.c_void_p * n)(*[f.ctypes.data for f in feats])
    out = np.empty((B, int(Cs.sum())), dtype=np.float32)
    if regions is None:
        rg, nr = None, None
    else:
        nr = np.array([len(r) for r in regions], dtype=np.int32)
        rg = np.ascontiguousarray(np.concatenate([np.asarray(r, dtype=np.int32).reshape(-1, 4) for r in regions]), dtype=np.int32)
    lib().oracle_region_pool_embed(ptrs, _p(Cs), _p(Hs), _p(Ws), ctypes.c_int(n), ctypes.c_int64(B),
                                   _p(rg) if rg is not None else None, _p(nr) if nr is not None else None,
                                   ctypes.c_int(int(per_region_l2)), ctypes.c_int(int(final_l2)), ctypes.c_float(eps), _p(out))
    return out


def sim_topk(queries, gallery, k=1, normalize=True, eps=1e-8):
    """nearest_neighbors (classification.py:90-95) with exact (double) dot products.
    Returns (cosine similarity f64[Q,k], gallery index int64[Q,k]), ordered (similarity desc, index asc)."""
    q = np.ascontiguousarray(queries, dtype=np.float32)
    g = np.ascontiguousarray(gallery, dtype=np.float32)
    Q, D = q.shape
    G = g.shape[0]
    sim = np.empty((Q, k), dtype=np.float64)
    idx = np.empty((Q, k), dtype=np.int64)
    L = lib()
    _parallel_rows(Q, lambda lo, hi: L.oracle_sim_topk(_p(q[lo:hi]), _p(g), ctypes.c_int64(hi - lo), ctypes.c_int64(G),
                                                       ctypes.c_int64(D), ctypes.c_int(k), ctypes.c_int(int(normalize)),
                                                       ctypes.c_double(eps), _p(sim[lo:hi]), _p(idx[lo:hi])))
    return sim, idx


def sim_topk_f32(queries, gallery, k=1):
    q = np.ascontiguousarray(queries, dtype=np.float32)
    g = np.ascontiguousarray(gallery, dtype=np.float32)
    Q, D = q.shape
    G = g.shape[0]
    sim = np.empty((Q, k), dtype=np.float32)
    idx = np.empty((Q, k), dtype=np.int64)
    L = lib()
    _parallel_rows(Q, lambda lo, hi: L.oracle_sim_topk_f32(_p(q[lo:hi]), _p(g), ctypes.c_int64(hi - lo), ctypes.c_int64(G),
                                                           ctypes.c_int64(D), ctypes.c_int(k), _p(sim[lo:hi]), _p(idx[lo:hi])))
    return sim, idx


def nearest_neighbors(anchors, queries, k=1):
    """Same argument order and return convention as classification.nearest_neighbors: int64[Q,k]."""
    return sim_topk(queries, anchors, k=k, normalize=True)[1]


BBOX_XFORM_CLIP = float(np.log(1000.0 / 16))  # tv:models/detection/_utils.py BoxCoder default


def det_sigmoid(x):
    L = lib()
    L.oracle_det_sigmoid.restype = ctypes.c_float
    return np.array([L.oracle_det_sigmoid(ctypes.c_float(float(v))) for v in np.asarray(x, dtype=np.float32).reshape(-1)],
                    dtype=np.float32)


def select_candidates(logits, regression, anchors, image_sizes, score_thresh=0.05, topk=1000, weights=None):
    """Pre-NMS candidate selection of RetinaNet.postprocess_detections (tv:retinanet.py:527-557), oracle.c:
    oracle_select_level per (image, level).
    logits: list over levels of f32[B,A_l,C]; regression: list of f32[B,A_l,4]; anchors: list of f32[A_l,4];
    image_sizes: [(h,w)]*B.  Returns per image (boxes f32[N,4], scores f32[N], labels i64[N], level_counts i64[L])."""
    L = lib()
    L.oracle_select_level.restype = ctypes.c_int64
    B = logits[0].shape[0]
    C = logits[0].shape[2]
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
    out = []
    for b in range(B):
        bs, ss, ls, cnt = [], [], [], []
        for lg, rg, an in zip(logits, regression, anchors):
            A = lg.shape[1]
            lgb = np.ascontiguousarray(lg[b], dtype=np.float32)
            rgb = np.ascontiguousarray(rg[b], dtype=np.float32)
            anb = np.ascontiguousarray(an, dtype=np.float32)
            cap = max(1, min(int(topk), A * C))
            ob = np.empty((cap, 4), dtype=np.float32)
            os_ = np.empty(cap, dtype=np.float32)
            ol = np.empty(cap, dtype=np.int64)
            k = L.oracle_select_level(_p(lgb), _p(rgb), _p(anb), ctypes.c_int64(A), ctypes.c_int64(C),
                                      ctypes.c_float(float(score_thresh)), ctypes.c_int64(int(topk)),
                                      ctypes.c_int64(int(image_sizes[b][0])), ctypes.c_int64(int(image_sizes[b][1])),
                                      _p(w) if w is not None else None, ctypes.c_float(BBOX_XFORM_CLIP), _p(ob), _p(os_), _p(ol))
            bs.append(ob[:k])
            ss.append(os_[:k])
            ls.append(ol[:k])
            cnt.append(k)
        out.append((np.concatenate(bs), np.concatenate(ss), np.concatenate(ls), np.array(cnt, dtype=np.int64)))
    return out


def postprocess_detections(logits, regression, anchors, image_sizes, score_thresh=0.05, topk=1000, nms_thresh=0.5,
                           detections_per_img=300):
    """RetinaNet.postprocess_detections for a single-class head (tv:retinanet.py:508-571): selection, batched_nms,
    [:detections_per_img].  Returns per image (boxes, scores, labels)."""
    res = []
    for boxes, scores, labels, _ in select_candidates(logits, regression, anchors, image_sizes, score_thresh, topk):
        keep = nms(boxes, scores, nms_thresh)[:detections_per_img]
        res.append((boxes[keep], scores[keep], labels[keep]))
    return res


def box_iou(a, b):
    """tv:ops/boxes.py box_iou (cvpce/metrics.py:11-13, cvpce/planograms.py:224): f32 [Na,Nb]."""
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 4)
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, 4)
    out = np.empty((a.shape[0], b.shape[0]), dtype=np.float32)
    lib().oracle_box_iou(_p(a), ctypes.c_int(a.shape[0]), _p(b), ctypes.c_int(b.shape[0]), _p(out))
    return out


def iou_match(rows, cols, iou_threshold=0.5, row_labels=None, col_labels=None):
    """The reference's sorted greedy walk (cvpce/metrics.py:15-34; cvpce/planograms.py:214-232) -> (row_matched, col_matched) bool."""
    rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, 4)
    cols = np.ascontiguousarray(cols, dtype=np.float32).reshape(-1, 4)
    rl = None if row_labels is None else np.ascontiguousarray(row_labels, dtype=np.int64)
    cl = None if col_labels is None else np.ascontiguousarray(col_labels, dtype=np.int64)
    rm = np.zeros(rows.shape[0], dtype=np.uint8)
    cm = np.zeros(cols.shape[0], dtype=np.uint8)
    lib().oracle_iou_match(_p(rows), None if rl is None else _p(rl), ctypes.c_int(rows.shape[0]), _p(cols),
                           None if cl is None else _p(cl), ctypes.c_int(cols.shape[0]), ctypes.c_float(iou_threshold), _p(rm), _p(cm))
    return rm.astype(bool), cm.astype(bool)
