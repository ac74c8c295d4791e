"""Host-side mirror of the detector post-processing the reference's GLN runs (cvpce/models/proposals.py:161-182
GaussianLayerNetwork inherits torchvision's RetinaNet.postprocess_detections unchanged, tv:retinanet.py:508-571):
candidate selection + NMS (a1) + [:detections_per_img] (a2) in two launches for the whole batch and ONE
host read (the per-image keep counts), instead of ~25 launches and 5 host syncs per image.
"""
import torch

from . import ops


def postprocess_detections_batch(head_outputs, anchors, image_shapes, score_thresh=0.05, topk_candidates=1000,
                                 nms_thresh=0.5, detections_per_img=300, min_score=None, box_scale=None):
    """Device-resident form: returns (boxes f32 [B*stride,4], scores f32 [B*stride], keep_counts i32 [B], stride); image b's
    detections are rows [b*stride, b*stride + keep_counts[b]), score-descending.  No host synchronisation.
    `min_score` folds the production confidence filter (production.py:15) in, `box_scale` f32 [B,2] the rescale to the
    original image (tv:transform.py:306-319)."""
    cls_logits, bbox_regression = head_outputs["cls_logits"], head_outputs["bbox_regression"]
    if cls_logits[0].shape[-1] != 1:
        raise NotImplementedError("postprocess_detections: the sm_100a NMS is class-agnostic (GLN has one class); "
                                  "got num_classes=%d" % cls_logits[0].shape[-1])
    level_anchors = _shared_anchors(anchors)
    boxes, scores, _, seg_counts, _, stride = ops.select_candidates(
        cls_logits, bbox_regression, level_anchors, image_shapes, score_thresh, topk_candidates, want_labels=False)
    _, counts, kept_boxes, kept_scores = ops.nms_segments(boxes, scores, stride, seg_counts, nms_thresh, min_score=min_score,
                                                          max_keep=detections_per_img, box_scale=box_scale,
                                                          want_boxes=True, want_scores=True)
    return kept_boxes, kept_scores, counts, stride


def _shared_anchors(anchors):
    """tv:retinanet.py:650-652 hands one list of per-level anchors PER IMAGE; the AnchorGenerator builds them from the padded
    batch shape, so every image has the same set: use image 0's (checked by shape; a [B,A,4] stack is accepted as is)."""
    if isinstance(anchors[0], (list, tuple)):
        first = anchors[0]
        for a in anchors[1:]:
            if len(a) != len(first) or any(x.shape != y.shape for x, y in zip(a, first)):
                raise RuntimeError("postprocess_detections: images of one batch must share their anchor shapes")
        return list(first)
    return list(anchors)


def postprocess_detections(head_outputs, anchors, image_shapes, score_thresh=0.05, topk_candidates=1000, nms_thresh=0.5,
                           detections_per_img=300):
    """Same arguments and result as RetinaNet.postprocess_detections(self, head_outputs, anchors, image_shapes) with the
    model's thresholds passed explicitly: a list of {'boxes','scores','labels'} per image."""
    boxes, scores, counts, stride = postprocess_detections_batch(head_outputs, anchors, image_shapes, score_thresh,
                                                                 topk_candidates, nms_thresh, detections_per_img)
    counts = ops.checked_counts(counts.cpu(), "postprocess_detections").tolist()  # the one host read
    out = []
    for b, n in enumerate(counts):
        out.append({"boxes": boxes[b * stride: b * stride + n], "scores": scores[b * stride: b * stride + n],
                    "labels": torch.zeros(n, dtype=torch.int64, device=boxes.device)})
    return out


def retinanet_postprocess_detections(self, head_outputs, anchors, image_shapes):
    """Bound replacement for RetinaNet.postprocess_detections (installed by cvpce_b200.patch.apply)."""
    return postprocess_detections(head_outputs, anchors, image_shapes, self.score_thresh, self.topk_candidates,
                                  self.nms_thresh, self.detections_per_img)
