"""Host-side mirror of cvpce/detection_eval.py (evaluate_detections, mean_average_metrics): the entry point that drives
detector -> crop every detection -> classify -> per-class detection metrics.  Same arguments and return
value.  Per image the crop loop (detection_eval.py:29) is one launch, classify searches all detections at once, and the
matching of every class's images is one claim launch per threshold (cvpce_b200.metrics); precision / recall / AP bookkeeping
is the reference's own metrics._do_calculate (not on the path, not rebuilt) -- found in an imported `cvpce` package or passed
as `do_calculate`."""
import sys

import torch
from torch.utils import data as tdata

from . import datautils, metrics, production


def _collate(samples):
    images, targets = zip(*samples)
    return list(images), list(targets)


def _reference_do_calculate():
    ref = sys.modules.get("cvpce.metrics")
    if ref is None or not hasattr(ref, "_do_calculate"):
        raise RuntimeError("cvpce_b200.detection_eval: pass do_calculate= (the reference's cvpce.metrics._do_calculate: the AP "
                           "bookkeeping is not part of the hot path and is not rebuilt) or import cvpce.metrics first")
    return ref._do_calculate


def evaluate_detections(p_model, c_model, testset, trainset, thresholds=(0.5,), proposal_batch_size=2, classification_batch_size=16,
                        num_workers=8, load_classifier_index=None, device=torch.device('cuda'), do_calculate=None):
    """detection_eval.py:6-49.  testset yields (img f32 [3,H,W], {'boxes' f32 [n,4], 'labels' i64 [n]}) and has `int_to_ann` /
    `ann_to_int`; p_model(list of images) returns torchvision detection dicts.  Returns (per-class metrics, overall metrics)."""
    do_calculate = do_calculate or _reference_do_calculate()
    loader = tdata.DataLoader(testset, batch_size=proposal_batch_size, num_workers=num_workers, collate_fn=_collate, pin_memory=True)
    classifier = production.Classifier(c_model, trainset, device=device, emb_device=device, batch_size=classification_batch_size,
                                       num_workers=num_workers, load=load_classifier_index)
    n_classes = len(testset.int_to_ann)
    per_class = {c: ([], [], []) for c in range(n_classes)}   # targets, predictions, confidences
    overall = ([], [], [])
    with torch.no_grad():
        for images, batch_targets in loader:
            images = [im.to(device, non_blocking=True) for im in images]
            for img, r, t in zip(images, p_model(images), batch_targets):
                boxes, scores = r['boxes'], r['scores']
                crops = datautils.crops_for_classification(img, boxes)                      # detection_eval.py:29
                top1 = [a[0] for a in classifier.classify(crops)]                           # :30
                cls = torch.tensor([testset.ann_to_int.get(a, -1) for a in top1], dtype=torch.int64)
                boxes, scores = boxes.detach().cpu(), scores.detach().cpu()
                t_boxes, t_labels = t['boxes'].detach().cpu(), t['labels'].detach().cpu()
                for c in sorted(set(cls.tolist()) | set(t_labels.tolist())):
                    group = (t_boxes[t_labels == c], boxes[cls == c], scores[cls == c])
                    for dst in ((overall,) if c == -1 else (overall, per_class[c])):
                        for lst, item in zip(dst, group):
                            lst.append(item)
    def strip(res):
        return {thr: {k: v for k, v in itm.items() if k != 'raw'} for thr, itm in res.items()}
    res = {c: strip(metrics.calculate_metrics(*per_class[c], thresholds, do_calculate=do_calculate)) for c in range(n_classes)}
    return res, strip(metrics.calculate_metrics(*overall, thresholds, do_calculate=do_calculate))


def mean_average_metrics(per_class_metrics, thresholds):
    """detection_eval.py:51-55."""
    n = len(per_class_metrics)
    return {t: {'map': sum(d[t]['ap'] for d in per_class_metrics.values()) / n,
                'mar300': sum(d[t]['ar_300'] for d in per_class_metrics.values()) / n} for t in thresholds}
