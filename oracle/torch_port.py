"""Torch-CPU restatement of the reference's own call sequence for the hot path.  TEST INFRASTRUCTURE / CPU BASELINE ONLY.

This is what `bench.py --impl reference` and bench.py's `cpu_baseline` time on the box's host cores (kind "port"): the
same torch / torchvision CPU ops the reference executes, in the same order, so the number is the reference's CPU cost
rather than the cost of oracle.c's scalar loops.  /root/reference cannot travel to the GPU box; tests/test_reference_pin.py
checks this port against the real reference where it is available.  Never imported by cvpce_b200/.
"""
import time

import torch
import torchvision
from torch import nn
from torchvision.models import vgg
from torchvision.transforms import functional as ttf

CLASSIFICATION_IMAGE_SIZE = 256  # cvpce/datautils.py:232


def postprocess_nms(boxes, scores, nms_thresh=0.5, detections_per_img=200):
    """tv:models/detection/retinanet.py:559-561 for the single-class GLN."""
    labels = torch.zeros(len(scores), dtype=torch.int64)
    return torchvision.ops.batched_nms(boxes, scores, labels, nms_thresh)[:detections_per_img]


def resize_for_classification(img, size=CLASSIFICATION_IMAGE_SIZE):
    """cvpce/datautils.py:234-239"""
    h, w = img.shape[1:]
    side = max(h, w)
    canvas = torch.full((3, side, side), 0.5)
    canvas[:, :h, :w] = img
    return ttf.resize(canvas, (size, size))


def crops_for_boxes(image, boxes, size=CLASSIFICATION_IMAGE_SIZE):
    """cvpce/production.py:16-20 (the per-box python loop)."""
    if not len(boxes):
        return torch.empty((0, 3, size, size))
    out = []
    for x1, y1, x2, y2 in boxes.to(dtype=torch.long):
        out.append(resize_for_classification(image[:, y1:y2, x1:x2], size))
    return torch.stack(out)


class MACVGGPort(nn.Module):
    """cvpce/models/classification.py:20-51 with vgg16 (no BN), as the production CLIs build it."""
    embedding_size = 1024

    def __init__(self):
        super().__init__()
        feats = vgg.VGG(vgg.make_layers(vgg.cfgs['D'], batch_norm=False), init_weights=True).features
        self.block1, self.block2 = feats[:23], feats[23:30]

    def forward(self, x, eps=1e-8):
        x = ttf.normalize(x, [0.485 * 2 - 1, 0.456 * 2 - 1, 0.406 * 2 - 1], [0.229 * 2, 0.224 * 2, 0.225 * 2])
        x = self.block1(x)
        d1 = x.amax(dim=(-2, -1))
        x = self.block2(x)
        d2 = x.amax(dim=(-2, -1))
        d = torch.cat((d1, d2), dim=1)
        return d / torch.linalg.norm(d, dim=1, keepdim=True).clamp(min=eps)


def nearest_neighbors(anchors, queries, k=1):
    """cvpce/models/classification.py:87-95 in the reference's own form (gathers two [Q,A,D] tensors, full argsort)."""
    ai = torch.arange(len(anchors))
    qi = torch.arange(len(queries))
    qm, am = torch.meshgrid(qi, ai, indexing="ij")
    dist = 1 - nn.functional.cosine_similarity(anchors[am], queries[qm], dim=-1)
    return dist.argsort(dim=-1)[:, :k]


def nearest_neighbors_chunked(anchors, queries, k=1, chunk=8):
    """Same result for shapes where the reference form cannot allocate Q*A*D*4 bytes twice: queries in small chunks."""
    return torch.cat([nearest_neighbors(anchors, queries[i:i + chunk], k) for i in range(0, len(queries), chunk)])


def _tick(stages, name, t0):
    """Accumulates wall seconds of one stage into `stages` (per-stage CPU timings)."""
    if stages is not None:
        stages[name] = stages.get(name, 0.0) + (time.perf_counter() - t0)


@torch.no_grad()
def classify(encoder, gallery_embedding, images, k=1, batch_size=32, nn_chunk=None, stages=None, return_embedding=False):
    """cvpce/production.py:57-74 (returns indices instead of annotation strings).  `stages`: optional dict that receives the
    seconds spent in the encoder and in nearest_neighbors."""
    res, embs = [], []
    for i in range(0, len(images), batch_size):
        t0 = time.perf_counter()
        batch = images[i:i + batch_size] * 2 - 1  # utils.scale_to_tanh
        emb = encoder(batch)
        _tick(stages, "encoder", t0)
        t0 = time.perf_counter()
        if nn_chunk:
            res.append(nearest_neighbors_chunked(gallery_embedding, emb, k, nn_chunk))
        else:
            res.append(nearest_neighbors(gallery_embedding, emb, k))
        _tick(stages, "nearest_neighbors", t0)
        embs.append(emb)
    idx = torch.cat(res) if res else torch.empty((0, k), dtype=torch.int64)
    if return_embedding:
        return idx, (torch.cat(embs) if embs else torch.empty((0, encoder.embedding_size)))
    return idx


@torch.no_grad()
def hot_path(encoder, gallery_embedding, image, cand_boxes, cand_scores, k=1, detections_per_img=200, max_crops=None,
             batch_size=32, nn_chunk=None, stages=None, return_embedding=False):
    """One shelf image through the whole path on the CPU: NMS -> crops -> encoder -> nearest neighbours."""
    t0 = time.perf_counter()
    keep = postprocess_nms(cand_boxes, cand_scores, 0.5, detections_per_img)
    boxes = cand_boxes[keep]
    _tick(stages, "nms", t0)
    if max_crops is not None:
        boxes = boxes[:max_crops]
    t0 = time.perf_counter()
    crops = crops_for_boxes(image, boxes)
    _tick(stages, "crop_loop", t0)
    res = classify(encoder, gallery_embedding, crops, k, batch_size, nn_chunk, stages, return_embedding)
    if return_embedding:
        return boxes, res[0], res[1]
    return boxes, res


# ---------------------------------------------------------------------------------------------------------------------
# detection-metric bookkeeping (cvpce/metrics.py:36-124): NOT on the hot path and not rebuilt in cvpce_b200 -- the product
# hands its matches to the reference's own _do_calculate.  The GPU box has no reference, so the a14 integration test passes
# this restatement instead; tests/test_reference_pin.py checks it against the live metrics._do_calculate.
# ---------------------------------------------------------------------------------------------------------------------
def _precision_and_recall(tp, fp, total_targets):
    """metrics.py:46-55"""
    tp, fp = tp.cumsum(0), fp.cumsum(0)
    precision = tp / (tp + fp)
    precision[torch.isnan(precision)] = 0
    recall = tp / total_targets if total_targets > 0 else torch.zeros_like(tp)
    return precision, recall


def _average_precision(precision, recall):
    """metrics.py:62-70: 11-point interpolated AP."""
    values = torch.zeros(11, dtype=torch.float)
    for i, r in enumerate(torch.linspace(0, 1, 11)):
        at = precision[recall >= r]
        if len(at) == 0:
            break
        values[i] = at.max()
    return values.mean()


def do_calculate(iou_thresholds, matches_for_threshold, sorted_confidences, total_targets):
    """metrics.py:36-44 (merge_matches) + :97-124 (_do_calculate)."""
    conf, order = torch.sort(torch.cat(sorted_confidences), descending=True)
    res = {}
    for t in iou_thresholds:
        d = matches_for_threshold[t]
        tp, fp = torch.cat(d["true_positives"])[order], torch.cat(d["false_positives"])[order]
        ar_300 = sum(d["recall_300"]) / len(d["recall_300"])
        p, r = _precision_and_recall(tp, fp, total_targets)
        f = 2 * p * r / (p + r)
        f[torch.isnan(f)] = 0
        if len(f) > 0:
            max_f, at = f.max(0)
            best = (max_f, p[at], r[at], conf[at])
        else:
            best = (0.0, 0.0, 0.0, 0.0)
        res[t] = {"raw": {"p": p, "r": r, "f": f, "c": conf}, "f": best[0], "p": best[1], "r": best[2], "c": best[3],
                  "ap": _average_precision(p, r), "ar_300": ar_300}
    return res
