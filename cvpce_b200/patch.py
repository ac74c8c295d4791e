"""Drop-in patchers: swap the hot-path call sites of an imported, unmodified reference (and of torchvision's RetinaNet
post-processing) for the sm_100a implementations.  Nothing in /root/reference is edited; see INTEGRATION.md.

    import cvpce, cvpce_b200.patch
    cvpce_b200.patch.apply(cvpce)      # idempotent; cvpce_b200.patch.revert() undoes it
"""
import sys

import torch

from . import _lib, classification, datautils, detection, metrics, ops, planograms, production

_saved = []


def _swap(obj, name, new):
    _saved.append((obj, name, getattr(obj, name)))
    setattr(obj, name, new)


def _make_batched_nms(original):
    """retinanet.box_ops IS torchvision.ops.boxes, a module shared by every torchvision detector of the process (Faster R-CNN
    RPN / RoIHeads, FCOS, SSD ...).  Only calls the sm_100a kernel is built for are taken over: CUDA boxes, at most
    CVPCE_NMS_MAX_N of them.  Anything else is not this library's call site and runs torchvision's own function unchanged
    (an RPN training step may pass 10k boxes; data-preparation code may call it on the host)."""
    def _batched_nms(boxes, scores, idxs, iou_threshold):
        if boxes.is_cuda and boxes.shape[0] <= _lib.NMS_MAX_N:
            return ops.batched_nms(boxes, scores, idxs, iou_threshold)
        return original(boxes, scores, idxs, iou_threshold)
    return _batched_nms


def _macvgg_forward(self, x, eps=1e-8):
    from torchvision.transforms import functional as ttf
    x = ttf.normalize(x, classification.NORM_MEAN, classification.NORM_STD)
    f4 = self.block1(x)
    f5 = self.block2(f4)
    return ops.region_pool_embed([f4, f5], eps=eps)


def _macresnet_forward(self, x, eps=1e-8):
    feats = []
    for b in self.blocks:
        x = b(x)
        feats.append(x)
    return ops.region_pool_embed(feats, eps=eps)


def _generate_proposals_and_images(self, image):
    boxes = self.generate_proposals(image)
    S = datautils.CLASSIFICATION_IMAGE_SIZE
    if not len(boxes):
        return boxes, torch.empty((0, 3, S, S))
    return boxes, datautils.crops_for_classification(image.to(boxes.device), boxes)


def _make_resize_for_classification(original):
    """datautils.resize_for_classification has two kinds of callers in the reference: the hot path's per-box crop loop
    (production.py:20, classification_eval.py:24, detection_eval.py:29 -- replaced wholesale by the batched kernel, and served
    here when a straggler passes a CUDA crop) and HOST-side data preparation that is not on the path at all: the gallery
    dataset's __getitem__ inside forked DataLoader workers (datautils.py:511, under Classifier.build_index) and
    PlanogramComparator.compare on a host image (production.py:111).  Host images keep running the reference's own function."""
    def _resize_for_classification(img):
        if img.is_cuda:
            return datautils.resize_for_classification(img)
        return original(img)
    return _resize_for_classification


def _classify(self, images, return_embedding=False):
    return production.Classifier.classify(self, images, return_embedding)


def apply(ref=None):
    """Patch the reference package `ref` (default: sys.modules['cvpce']) and torchvision's RetinaNet NMS call site."""
    if _saved:
        return
    ref = ref or sys.modules.get("cvpce")
    from torchvision.models.detection import retinanet
    _swap(retinanet.box_ops, "batched_nms", _make_batched_nms(retinanet.box_ops.batched_nms))   # tv:retinanet.py:560
    if ref is None:
        return
    gln = getattr(getattr(ref.models, "proposals", None), "GaussianLayerNetwork", None)
    if gln is not None:  # single-class detector: sigmoid/top-k/decode/clip + NMS in two launches (tv:retinanet.py:508-571)
        def _gln_postprocess(self, head_outputs, anchors, image_shapes):
            if not head_outputs["cls_logits"][0].is_cuda:
                raise RuntimeError("cvpce_b200.patch: postprocess_detections called with CPU tensors; no CPU fallback")
            if head_outputs["cls_logits"][0].shape[-1] != 1:  # multi-class heads keep torchvision's loop (+ patched NMS)
                return retinanet.RetinaNet.postprocess_detections(self, head_outputs, anchors, image_shapes)
            return detection.retinanet_postprocess_detections(self, head_outputs, anchors, image_shapes)
        _saved.append((gln, "postprocess_detections", None))
        gln.postprocess_detections = _gln_postprocess                            # proposals.py:161 (inherited method)
    rc, rp, rd = ref.models.classification, ref.production, ref.datautils
    _swap(rc, "nearest_neighbors", classification.nearest_neighbors)            # classification.py:90
    _swap(rp, "nearest_neighbors", classification.nearest_neighbors)            # production.py:6 (from-import alias)
    if hasattr(ref, "utils") and hasattr(ref.utils, "nearest_neighbors"):
        _swap(ref.utils, "nearest_neighbors", classification.nearest_neighbors) # utils.py:20
    _swap(rc.MACVGG, "forward", _macvgg_forward)                                # classification.py:38-51
    _swap(rc.MACResNet, "forward", _macresnet_forward)                          # classification.py:78-85
    _swap(rd, "resize_for_classification", _make_resize_for_classification(rd.resize_for_classification))  # datautils.py:234
    _swap(rp.ProposalGenerator, "generate_proposals_and_images", _generate_proposals_and_images)  # production.py:16
    # f3: detection-evaluation matching and the planogram box_iou call site
    rm = getattr(ref, "metrics", None)
    if rm is not None:
        do_calc = rm._do_calculate

        def _calculate_metrics(targets, predictions, confidences, iou_thresholds=(0.5,)):
            return metrics.calculate_metrics(targets, predictions, confidences, iou_thresholds, do_calculate=do_calc)
        _swap(rm, "iou_matrices", metrics.iou_matrices)                         # metrics.py:11
        _swap(rm, "_process_one", metrics.process_one)                          # metrics.py:80 (sync path; the async workers
        _swap(rm, "calculate_metrics", _calculate_metrics)                      # metrics.py:126  are forked CPU processes)
    rg = getattr(ref, "planograms", None)
    if rg is not None and hasattr(rg, "tvops"):
        _swap(rg, "tvops", planograms.TvopsProxy())                             # planograms.py:224
    # Classifier.classify keeps its signature; the patched version embeds all crops, searches once, syncs once
    rp.Classifier.embed = production.Classifier.embed
    rp.Classifier.prepared_gallery = production.Classifier.prepared_gallery
    rp.Classifier._prepared = None
    _swap(rp.Classifier, "classify", _classify)                                 # production.py:57


def revert():
    while _saved:
        obj, name, old = _saved.pop()
        if old is None:
            delattr(obj, name)  # the attribute was inherited before the patch
        else:
            setattr(obj, name, old)
