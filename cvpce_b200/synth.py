"""Seeded synthetic workloads shared by bench.py, the tests and the golden generator (host side, numpy only).

numpy RandomState only, so the same seed gives the same bytes in the authoring container and on the GPU box.
"""
from math import sqrt

import numpy as np

WEIGHT_SEED = 0            # torch.manual_seed before building the random-init VGG16 encoder
CROP_IMAGE_SEED = 7
CROP_IMAGE_HW = (384, 512)
SHELF_HW = (1536, 2048)    # BASELINE.json configs 2-4: 2048x1536 shelves


def shelf_image(seed, H, W):
    return np.random.RandomState(seed).rand(3, H, W).astype(np.float32)


def shelf_candidates(seed, H=SHELF_HW[0], W=SHELF_HW[1], n_products=200, duplicates=4):
    """Scored candidate boxes as a GLN would hand them to NMS: n_products true boxes on a jittered shelf grid
    (w in [60,400], h in [90,400] px), each with one primary candidate (jitter sigma 2 px, score U(0.6,1)) and
    `duplicates` worse ones (sigma 6 px, score U(0.05,0.6)). Scores are distinct fp32. Returns (boxes f32[N,4], scores f32[N])."""
    rs = np.random.RandomState(seed)
    cols = int(np.ceil(np.sqrt(n_products * W / H)))
    rows = int(np.ceil(n_products / cols))
    cw, ch = W / cols, H / rows
    boxes, scores = [], []
    for i in range(n_products):
        r, c = divmod(i, cols)
        w = min(rs.uniform(60, 400), cw * 1.15)
        h = min(rs.uniform(90, 400), ch * 1.15)
        cx = (c + 0.5) * cw + rs.uniform(-0.1, 0.1) * cw
        cy = (r + 0.5) * ch + rs.uniform(-0.1, 0.1) * ch
        base = np.array([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2])
        boxes.append(base + rs.normal(0, 2, 4))
        scores.append(rs.uniform(0.6, 1.0))
        for _ in range(duplicates):
            boxes.append(base + rs.normal(0, 6, 4))
            scores.append(rs.uniform(0.05, 0.6))
    boxes = np.asarray(boxes, dtype=np.float32)
    boxes[:, 0::2] = np.clip(boxes[:, 0::2], 0, W)
    boxes[:, 1::2] = np.clip(boxes[:, 1::2], 0, H)
    scores = np.asarray(scores, dtype=np.float32)
    # make scores distinct in fp32 (re-draw collisions)
    while len(np.unique(scores)) != len(scores):
        _, first = np.unique(scores, return_index=True)
        dup = np.setdiff1d(np.arange(len(scores)), first)
        scores[dup] = rs.uniform(0.05, 0.6, len(dup)).astype(np.float32)
    perm = rs.permutation(len(scores))
    return boxes[perm], scores[perm]


def random_boxes(seed, n, extent=1000.0, wh=(20.0, 140.0), tie_every=0):
    rs = np.random.RandomState(seed)
    xy = rs.uniform(0, extent, (n, 2))
    sz = rs.uniform(wh[0], wh[1], (n, 2))
    boxes = np.concatenate([xy, xy + sz], 1).astype(np.float32)
    scores = rs.uniform(0, 1, n).astype(np.float32)
    if tie_every:
        scores[::tie_every] = scores[0]
    return boxes, scores


def nms_cases():
    """name -> (boxes, scores, iou_threshold, detections_per_img)"""
    cases = {}
    cases["ties"] = (np.array([[0, 0, 10, 10]] * 3 + [[20, 20, 30, 30]] * 2, dtype=np.float32),
                     np.array([.5, .5, .5, .7, .7], dtype=np.float32), 0.5, 1000)
    cases["zero_area"] = (np.array([[5, 5, 5, 5], [5, 5, 5, 5], [0, 0, 10, 10], [1, 1, 1, 9]], dtype=np.float32),
                          np.array([.9, .8, .7, .6], dtype=np.float32), 0.5, 1000)
    cases["single"] = (np.array([[1, 2, 3, 4]], dtype=np.float32), np.array([.3], dtype=np.float32), 0.5, 1000)
    # hand-made IoU values 0.81 / 0.64 / 1/1.44 in the spirit of reference test/metrics_test.py:5-21
    cases["hand_iou"] = (np.array([[0, 0, 10, 10], [0, 0, 9, 9], [0, 0, 8, 8], [0, 0, 12, 12], [100, 100, 110, 110]], dtype=np.float32),
                         np.array([.9, .8, .7, .6, .5], dtype=np.float32), 0.65, 1000)
    b, s = random_boxes(11, 2000, tie_every=7)
    cases["random2000_ties"] = (b, s, 0.5, 1000)
    b, s = random_boxes(12, 5000)
    cases["random5000"] = (b, s, 0.5, 1000)
    b, s = random_boxes(13, 700, extent=300.0)
    cases["dense700_thr03"] = (b, s, 0.3, 200)
    b, s = shelf_candidates(100)
    cases["shelf200"] = (b, s, 0.5, 200)
    b, s = shelf_candidates(101, n_products=150)
    cases["shelf150"] = (b, s, 0.5, 1000)
    return cases


def crop_boxes():
    """Boxes on the CROP_IMAGE_HW image: up-sampling, down-sampling, 1-px strips, borders, whole image, fractional coords."""
    H, W = CROP_IMAGE_HW
    return np.array([
        [100.7, 200.2, 200.9, 350.5],   # survey KAT: ints (100,200,200,350)
        [0, 0, W, H],                   # whole image (down-sampling, L=512)
        [10, 10, 11, 300],              # 1-px wide
        [30, 50, 330, 51],              # 1-px tall
        [300.5, 100.5, 500.2, 380.9],   # down/up mix, L=280
        [50, 60, 90, 75],               # strong up-sampling
        [W - 1, H - 1, W, H],           # single pixel at the corner
        [5, 5, 405, 305],               # L=400 -> 256 down-sampling
        [0, 0, 256, 256],               # identity scale
        [17, 33, 274, 290],             # L=257
        [200, 0, 520, 384],             # x2 beyond the image: python slice clamps
        [123.9, 45.1, 223.2, 300.7],
    ], dtype=np.float32)


def sample_pixels(S, n=3072):
    return np.random.RandomState(S).choice(3 * S * S, n, replace=False)


def feature_maps():
    rs = np.random.RandomState(21)
    f4 = np.abs(rs.randn(5, 40, 12, 12)).astype(np.float32)
    f5 = np.abs(rs.randn(5, 24, 6, 6)).astype(np.float32)
    return f4, f5


def encoder_inputs(n=4, S=64, seed=31):
    return (np.random.RandomState(seed).rand(n, 3, S, S) * 2 - 1).astype(np.float32)


def nn_kat():
    """reference test/models/classification_test.py:8-25 (6 unit anchors in R^3, 7 queries, expected top-1)."""
    a = np.array([[1, 0, 0], [1 / sqrt(3)] * 3, [-1 / sqrt(3)] * 3, [-1, 0, 0], [1 / sqrt(2), 0, 1 / sqrt(2)],
                  [-1 / sqrt(2), 0, -1 / sqrt(2)]], dtype=np.float32)
    q = np.array([[1 / sqrt(1.01), 0.1 / sqrt(1.01), 0], [0.9 / sqrt(2.02), 0, 1.1 / sqrt(2.02)], [-1, 0, 0], [1, 0, 0],
                  [1 / sqrt(3)] * 3, [-1.1 / sqrt(2.02), 0, -0.9 / sqrt(2.02)], [-1, 0, 0]], dtype=np.float32)
    return a, q, np.array([0, 4, 3, 0, 1, 5, 3], dtype=np.int64)


# name -> (Q, G, D, k, seed)
NN_CASES = {
    "c1_like": (64, 180, 1024, 5, 41),      # BASELINE config 1/2 shape
    "d512": (48, 1000, 512, 5, 42),
    "small_odd": (7, 33, 100, 3, 43),       # D not a multiple of 64 -> fp32 path
    "k_gt_1_wide": (16, 300, 256, 10, 44),
}


def unit_rows(rs, n, d, nonneg=True):
    x = rs.randn(n, d)
    if nonneg:
        x = np.abs(x)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x.astype(np.float32)


def nn_case(Q, G, D, seed):
    rs = np.random.RandomState(seed)
    return unit_rows(rs, G, D), unit_rows(rs, Q, D)


def planted_queries(rs, gallery, Q, frac=0.1, noise=0.05):
    """C5-style queries: `frac` of them are noisy copies of random gallery rows (planted neighbours), the rest random."""
    G, D = gallery.shape
    q = rs.randn(Q, D).astype(np.float32)
    n_plant = int(Q * frac)
    planted = rs.randint(0, G, n_plant)
    q[:n_plant] = gallery[planted] + noise * rs.randn(n_plant, D).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.astype(np.float32), planted


def classify_case():
    rs = np.random.RandomState(51)
    gallery = (rs.rand(10, 3, 64, 64) * 2 - 1).astype(np.float32)
    queries = rs.rand(6, 3, 64, 64).astype(np.float32)
    # make three queries near-copies of gallery items so the expected labels are meaningful
    for qi, gi in ((0, 3), (2, 7), (5, 1)):
        queries[qi] = np.clip((gallery[gi] + 1) / 2 + 0.01 * rs.randn(3, 64, 64), 0, 1)
    return gallery, queries


class GalleryDataset:
    """Yields the (img[-1,1], raw img, categories, annotation) tuples Classifier.build_index expects (datautils.py:397-415)."""

    def __init__(self, images):
        import torch
        self.images = [torch.from_numpy(im) for im in images]

    def __len__(self):
        return len(self.images)

    def __getitem__(self, i):
        return self.images[i], self.images[i], [], "g%d" % i


# ---------------------------------------------------------------------------------------------------------------------
# detector-head outputs for the pre-NMS candidate selection (tv:retinanet.py:508-571)
# ---------------------------------------------------------------------------------------------------------------------
FPN_STRIDES = (8, 16, 32, 64, 128)       # P3..P7 of the GLN / RetinaNet FPN
FPN_SIZES = (32, 64, 128, 256, 512)      # tv:retinanet.py _default_anchorgen


def fpn_anchors(H, W):
    """RetinaNet-style anchors per FPN level: 9 per cell (3 scales x 3 aspect ratios), f32 [A_l,4] xyxy, cell-major with
    the 9 anchors innermost (the layout the head's permute/reshape produces)."""
    out = []
    for stride, size in zip(FPN_STRIDES, FPN_SIZES):
        fh, fw = -(-H // stride), -(-W // stride)
        scales = np.array([size * 2 ** (i / 3) for i in range(3)], dtype=np.float64)
        ratios = np.array([0.5, 1.0, 2.0])
        hr = np.sqrt(ratios)
        wr = 1.0 / hr
        ws = (wr[:, None] * scales[None, :]).reshape(-1)
        hs = (hr[:, None] * scales[None, :]).reshape(-1)
        base = np.round(np.stack([-ws, -hs, ws, hs], axis=1) / 2)
        sy, sx = np.meshgrid(np.arange(fh) * stride, np.arange(fw) * stride, indexing="ij")
        shifts = np.stack([sx.reshape(-1), sy.reshape(-1), sx.reshape(-1), sy.reshape(-1)], axis=1)
        out.append((shifts[:, None, :] + base[None, :, :]).reshape(-1, 4).astype(np.float32))
    return out


# name -> (seed, image sizes [(h,w)], padded batch size (H,W), num_classes, logit mean, logit std, quantise step, topk)
HEAD_CASES = {
    "dense_2img": (11, [(200, 264), (184, 248)], (200, 264), 1, -2.0, 2.0, 0.0, 1000),
    "sparse": (12, [(200, 264)], (200, 264), 1, -6.5, 1.5, 0.0, 1000),
    "multiclass": (13, [(136, 200)], (136, 200), 3, -2.5, 2.0, 0.0, 300),
    "ties": (14, [(200, 264)], (200, 264), 1, -1.0, 1.5, 1.0, 1000),        # 1.0-quantised logits: large tie groups
    "fullsize": (15, [(800, 1066)], (800, 1066), 1, -4.0, 1.5, 0.0, 1000),  # the GLN transform's 800 x <=1333 frame
}


def head_case(name):
    """Seeded head outputs of one case: (logits [f32 [B,A_l,C]], regression [f32 [B,A_l,4]], anchors [f32 [A_l,4]],
    image_sizes, num_classes, topk)."""
    seed, sizes, (H, W), C, mu, sd, q, topk = HEAD_CASES[name]
    rs = np.random.RandomState(seed)
    anchors = fpn_anchors(H, W)
    B = len(sizes)
    logits, regression = [], []
    for a in anchors:
        A = a.shape[0]
        lg = (mu + sd * rs.standard_normal((B, A, C))).astype(np.float32)
        if q > 0:
            lg = (np.round(lg / q) * q).astype(np.float32)
        rg = (0.4 * rs.standard_normal((B, A, 4))).astype(np.float32)
        # a few extreme deltas: exercise the bbox_xform_clip clamp and the clip to the image
        m = rs.random_sample((B, A)) < 0.01
        rg[m] *= 12.0
        logits.append(lg)
        regression.append(rg)
    return logits, regression, anchors, sizes, C, topk


GLN_FRAME = (800, 1066)  # tv GeneralizedRCNNTransform(min_size=800, max_size=1333) applied to a 1536 x 2048 shelf


def head_from_candidates(boxes, scores, image_hw, seed, frame=GLN_FRAME, background=(-6.0, 1.0)):
    """Head outputs of a 'trained' detector whose decoded, thresholded candidates are (up to fp32 rounding) the given scored
    boxes: every candidate is encoded (BoxCoder.encode_single, weights 1) against the best-shaped free anchor at its
    centre cell of the FPN level matching its size, with logit = logit(score); every other anchor gets a background logit
    N(background).  boxes are in original-image pixels (image_hw); the head works in the resized `frame`.
    Returns (logits f32 [A,1], regression f32 [A,4], anchors per level, box_scale (x ratio, y ratio) back to the image)."""
    H, W = image_hw
    fh, fw = frame
    sx, sy = fw / W, fh / H
    anchors = fpn_anchors(fh, fw)
    n_per = [a.shape[0] for a in anchors]
    off = np.concatenate([[0], np.cumsum(n_per)])
    rs = np.random.RandomState(seed)
    A = int(off[-1])
    logits = (background[0] + background[1] * rs.standard_normal(A)).astype(np.float32)
    regression = (0.1 * rs.standard_normal((A, 4))).astype(np.float32)
    used = set()
    b = np.asarray(boxes, dtype=np.float64) * [sx, sy, sx, sy]
    for (x1, y1, x2, y2), s in zip(b, np.asarray(scores, dtype=np.float64)):
        w, h = max(x2 - x1, 1e-3), max(y2 - y1, 1e-3)
        cx, cy = x1 + 0.5 * w, y1 + 0.5 * h
        lvl = int(np.argmin([abs(np.log(np.sqrt(w * h) / (sz * 2 ** (1 / 3)))) for sz in FPN_SIZES]))
        stride = FPN_STRIDES[lvl]
        gw = -(-fw // stride)
        gh = -(-fh // stride)
        ix, iy = min(gw - 1, max(0, int(round(cx / stride)))), min(gh - 1, max(0, int(round(cy / stride))))
        base = int(off[lvl]) + (iy * gw + ix) * 9
        cand = anchors[lvl][(iy * gw + ix) * 9:(iy * gw + ix) * 9 + 9].astype(np.float64)
        aw, ah = cand[:, 2] - cand[:, 0], cand[:, 3] - cand[:, 1]
        order = np.argsort(np.abs(np.log(w / aw)) + np.abs(np.log(h / ah)))
        slot = next((int(o) for o in order if base + int(o) not in used), None)
        if slot is None:
            continue  # all 9 anchors of the cell taken: drop this duplicate
        used.add(base + slot)
        ax1, ay1 = cand[slot, 0], cand[slot, 1]
        acx, acy = ax1 + 0.5 * aw[slot], ay1 + 0.5 * ah[slot]
        regression[base + slot] = [(cx - acx) / aw[slot], (cy - acy) / ah[slot], np.log(w / aw[slot]), np.log(h / ah[slot])]
        s = min(max(s, 1e-6), 1 - 1e-6)
        logits[base + slot] = np.log(s / (1 - s))
    return logits.reshape(A, 1), regression, anchors, (np.float32(W / fw), np.float32(H / fh))


# ---------------------------------------------------------------------------------------------------------------------
# box IoU / greedy matching: detection evaluation (cvpce/metrics.py) and planogram matching (cvpce/planograms.py)
# ---------------------------------------------------------------------------------------------------------------------
# the reference's own test data, test/metrics_test.py:5-21 (targets, predictions already in confidence order)
METRICS_KAT_TARGETS = [
    [[0, 0, 1, 1], [1, 0, 2, 1], [1, 1, 2, 2]],
    [[1, 1, 2, 2], [3, 1, 4, 2], [5, 1, 6, 2], [7, 1, 8, 2]],
    [[0, 0, 5, 5], [5, 5, 10, 10]],
]
METRICS_KAT_PREDICTIONS = [
    [[0, 0, .9, .9], [1.1, 0.1, 1.9, 0.9], [0, 0, 1, 1], [0.9, 0.9, 2.1, 2.1], [3, 3, 4, 4]],
    [[1, 0, 2, 1], [1, 1, 2, 2], [5, 1, 6, 2], [7, 1.1, 8, 1.9], [9, 9, 10, 10]],
    [[0, 0, 1, 1], [1, 1, 3, 3], [0.5, 0.5, 4.5, 4.5], [0, 0, 6, 6], [6, 6, 9, 9]],
]
MATCH_THRESHOLDS = (0.5, 0.65, 0.75)


def detection_eval_case(seed, n_targets, n_extra, duplicates=3, extent=1500.0):
    """(targets f32 [T,4], predictions f32 [P,4] in descending-confidence order): every target gets `duplicates` jittered
    predictions of varying quality plus `n_extra` unrelated boxes, shuffled by a random confidence."""
    rs = np.random.RandomState(seed)
    xy = rs.uniform(0, extent, size=(n_targets, 2))
    wh = rs.uniform(40, 220, size=(n_targets, 2))
    targets = np.concatenate([xy, xy + wh], axis=1).astype(np.float32)
    preds = []
    for d in range(duplicates):
        jit = rs.normal(0, 4.0 + 14.0 * d, size=(n_targets, 4))
        preds.append(targets + jit)
    xy = rs.uniform(0, extent, size=(n_extra, 2))
    wh = rs.uniform(40, 220, size=(n_extra, 2))
    preds.append(np.concatenate([xy, xy + wh], axis=1))
    preds = np.concatenate(preds).astype(np.float32)
    preds[:, 2:] = np.maximum(preds[:, 2:], preds[:, :2])  # keep x2 >= x1, y2 >= y1
    order = np.argsort(-rs.uniform(size=len(preds)), kind="stable")
    return targets, np.ascontiguousarray(preds[order])


def match_cases():
    """name -> (targets, predictions): the reference's three test cases, two random images and the empty-ish corners."""
    cases = {}
    for i, (t, p) in enumerate(zip(METRICS_KAT_TARGETS, METRICS_KAT_PREDICTIONS)):
        cases["kat%d" % i] = (np.array(t, dtype=np.float32), np.array(p, dtype=np.float32))
    cases["image150"] = detection_eval_case(41, 150, 300)
    cases["image40_dense"] = detection_eval_case(42, 40, 60, duplicates=5, extent=400.0)
    cases["one_target"] = detection_eval_case(43, 1, 6)
    return cases


def planogram_case(seed, n=60, n_labels=12, missing=9, extra=7):
    """A planogram (b1, l1) and a detection result (b2, l2) of the same shelf under a mild projective map: `missing` products
    are absent from the detections, `extra` detections have no planogram counterpart, a few detections carry a wrong label.
    Returns b1, l1, b2, l2 (labels as strings, like the reference's annotations) and the list of true (i, j) pairs."""
    rs = np.random.RandomState(seed)
    cols = 10
    b1 = []
    for i in range(n):
        r, c = divmod(i, cols)
        x, y = 40 + c * 130 + rs.uniform(-6, 6), 30 + r * 190 + rs.uniform(-6, 6)
        b1.append([x, y, x + rs.uniform(90, 120), y + rs.uniform(140, 175)])
    b1 = np.array(b1, dtype=np.float32)
    l1 = ["sku%02d" % rs.randint(n_labels) for _ in range(n)]
    Hm = np.array([[1.9, 0.03, 55.0], [-0.02, 1.85, 20.0], [1e-5, 2e-5, 1.0]])

    def proj(x, y):
        v = Hm @ np.array([x, y, 1.0])
        return v[:2] / v[2]
    present = np.sort(rs.permutation(n)[:n - missing])
    b2, l2, pairs = [], [], []
    for j, i in enumerate(present):
        p1, p2 = proj(*b1[i, :2]), proj(*b1[i, 2:])
        b2.append(np.concatenate([p1, p2]) + rs.normal(0, 3.0, size=4))
        l2.append(l1[i] if rs.uniform() > 0.08 else "sku%02d" % rs.randint(n_labels))
        pairs.append((int(i), j))
    for _ in range(extra):
        i = rs.randint(n)
        p1, p2 = proj(*b1[i, :2]), proj(*b1[i, 2:])
        b2.append(np.concatenate([p1, p2]) + rs.normal(0, 25.0, size=4))   # overlapping double detections
        l2.append(l1[i])
    b2 = np.array(b2, dtype=np.float32)
    solution = [pairs[t] for t in rs.permutation(len(pairs))[:14] if l2[pairs[t][1]] == l1[pairs[t][0]]]
    return b1, l1, b2, l2, solution


# ---------------------------------------------------------------------------------------------------------------------
# a14: a small synthetic evaluation set for the two entry points that drive the path
# (cvpce/classification_eval.py:6-56 eval_dihe, cvpce/detection_eval.py:6-49 evaluate_detections)
# ---------------------------------------------------------------------------------------------------------------------
EVAL_CROP_SIZE = 64          # CLASSIFICATION_IMAGE_SIZE used for the evaluation goldens (keeps the CPU reference run short)
EVAL_IMAGE_HW = (240, 320)
EVAL_THRESHOLDS = (0.5, 0.75)


def _resize_bilinear(img, h, w):
    """[3,H,W] -> [3,h,w], plain bilinear (numpy; only used to paint gallery items into the synthetic shelves)."""
    C, H, W = img.shape
    ys = np.clip((np.arange(h) + 0.5) * H / h - 0.5, 0, H - 1)
    xs = np.clip((np.arange(w) + 0.5) * W / w - 0.5, 0, W - 1)
    y0, x0 = np.floor(ys).astype(int), np.floor(xs).astype(int)
    y1, x1 = np.minimum(y0 + 1, H - 1), np.minimum(x0 + 1, W - 1)
    fy, fx = (ys - y0)[None, :, None], (xs - x0)[None, None, :]
    a = img[:, y0][:, :, x0] * (1 - fx) + img[:, y0][:, :, x1] * fx
    b = img[:, y1][:, :, x0] * (1 - fx) + img[:, y1][:, :, x1] * fx
    return (a * (1 - fy) + b * fy).astype(np.float32)


def eval_case(seed=61, n_images=4, n_products=9, n_gallery=12):
    """Gallery images [G,3,64,64] in [-1,1] (annotations "g<i>") and `n_images` synthetic shelves [3,H,W] in [0,1] on which
    `n_products` gallery items are painted (resized to their box, plus noise), with their boxes and labels.  About a fifth of
    the products are painted from a DIFFERENT item than their label says, and the box list of each shelf has one box that
    sticks out of the image (clip_boxes_to_image, classification_eval.py:23), so accuracy is neither 0 nor 1.
    Returns (gallery, annotations, shelves) with shelves = list of (image, boxes f32 [n,4], labels i64 [n])."""
    rs = np.random.RandomState(seed)
    H, W = EVAL_IMAGE_HW
    gallery = (rs.rand(n_gallery, 3, 64, 64) * 2 - 1).astype(np.float32)
    # smooth the items a little so that resizing keeps them recognisable to a random-init encoder
    for g in gallery:
        g[:] = (g + np.roll(g, 1, 1) + np.roll(g, 1, 2) + np.roll(np.roll(g, 1, 1), 1, 2)) / 4
    gallery = np.clip(gallery * 1.8, -1, 1).astype(np.float32)
    annotations = ["g%d" % i for i in range(n_gallery)]
    shelves = []
    for _ in range(n_images):
        img = (0.45 + 0.1 * rs.rand(3, H, W)).astype(np.float32)
        boxes, labels = [], []
        cols = 3
        for p in range(n_products):
            cw, ch = W // cols, H // ((n_products + cols - 1) // cols)
            bw, bh = int(rs.randint(40, cw - 8)), int(rs.randint(36, ch - 6))
            x1 = (p % cols) * cw + int(rs.randint(0, cw - bw))
            y1 = (p // cols) * ch + int(rs.randint(0, ch - bh))
            label = len(shelves) * n_products + p           # the first n_gallery products cover every class once
            label = label if label < n_gallery else int(rs.randint(0, n_gallery))
            painted = label if rs.rand() > 0.2 else int(rs.randint(0, n_gallery))
            patch = _resize_bilinear((gallery[painted] + 1) / 2, bh, bw) + 0.02 * rs.randn(3, bh, bw).astype(np.float32)
            img[:, y1:y1 + bh, x1:x1 + bw] = np.clip(patch, 0, 1)
            boxes.append([x1, y1, x1 + bw, y1 + bh])
            labels.append(label)
        boxes = np.asarray(boxes, dtype=np.float32)
        boxes[-1, 2:] += np.array([W, H], dtype=np.float32) * 0.1   # sticks out of the image on purpose
        shelves.append((img, boxes, np.asarray(labels, dtype=np.int64)))
    return gallery, annotations, shelves


def eval_detector_heads(shelves, seed=71):
    """Synthetic detector-head outputs for the evaluate_detections golden: every ground-truth box of a shelf gets one well
    scored candidate (jittered by ~2 px) and two low-scored duplicates, two thirds of the shelves' boxes only (missed products
    lower the recall), plus a few false positives.  One (logits, regression, anchors) triple per shelf, frame = image size."""
    rs = np.random.RandomState(seed)
    H, W = EVAL_IMAGE_HW
    heads = []
    for si, (_, boxes, _) in enumerate(shelves):
        gt = np.clip(boxes, 0, [W, H, W, H])
        keep = rs.rand(len(gt)) > 0.3
        cand, score = [], []
        for b in gt[keep]:
            cand.append(b + rs.randn(4) * 2.0)
            score.append(rs.uniform(0.6, 0.98))
            for _ in range(2):
                cand.append(b + rs.randn(4) * 5.0)
                score.append(rs.uniform(0.08, 0.5))
        for _ in range(3):
            x, y = rs.uniform(0, W - 60), rs.uniform(0, H - 60)
            cand.append(np.array([x, y, x + rs.uniform(30, 60), y + rs.uniform(30, 60)]))
            score.append(rs.uniform(0.3, 0.7))
        cand = np.clip(np.asarray(cand, dtype=np.float32), 0, [W, H, W, H]).astype(np.float32)
        lg, rg, anc, _ = head_from_candidates(cand, np.asarray(score, dtype=np.float32), (H, W), seed * 100 + si, frame=(H, W))
        heads.append((lg, rg, anc))
    return heads
