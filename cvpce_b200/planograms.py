"""Host-side mirror of the expected-vs-detected matching inside the reference's planograms.finalize_via_ransac
(cvpce/planograms.py:210-236) on the sm_100a claim kernel.  Graph matching and the RANSAC homography (planograms.py:1-208)
are CPU graph / OpenCV code outside the hot path and stay in the reference.
"""
import torch
import torchvision

from . import ops


def labels_to_ids(l1, *ln):
    """utils.labels_to_tensors (cvpce/utils.py:290-294): shared integer ids for lists of label strings. Only equality
    matters downstream, so ids are assigned in first-seen order instead of set order."""
    key, conv = [], {}
    for lbl in [l1, *ln]:
        for l in lbl:
            if l not in conv:
                conv[l] = len(key)
                key.append(l)
    return (*(torch.tensor([conv[l] for l in lbl], dtype=torch.long) for lbl in [l1, *ln]), key)


def match_expected(expected_positions, detected_boxes, expected_labels, detected_labels, iou_threshold=0.5):
    """planograms.py:210-236: which projected planogram boxes found a detection of their own label with IoU >= threshold
    (every such unclaimed detection is claimed, in planogram order), and which detections were claimed.

    expected_positions f32 [E,4], detected_boxes f32 [D,4] (any device; the arithmetic runs on the current CUDA device);
    labels: lists of hashable labels or integer tensors.  Returns (matched_expected bool [E], matched_actual bool [D],
    missing_expected i64 [M]) on the device of `expected_positions`."""
    if not torch.cuda.is_available():
        raise RuntimeError("cvpce_b200.planograms: no CUDA device (cvpce_b200 has no CPU fallback)")
    dev = torch.device("cuda", torch.cuda.current_device())
    if not torch.is_tensor(expected_labels):
        expected_labels, detected_labels, _ = labels_to_ids(list(expected_labels), list(detected_labels))
    me, ma = ops.iou_match(expected_positions.to(dev), detected_boxes.to(dev), iou_threshold,
                           row_labels=expected_labels.to(dev), col_labels=detected_labels.to(dev))
    me, ma = me.to(expected_positions.device), ma.to(expected_positions.device)
    return me, ma, torch.where(~me)[0]


class TvopsProxy:
    """Stands in for the `tvops` name inside cvpce/planograms.py (and cvpce/metrics.py): `box_iou` runs on the sm_100a kernel
    (result returned on the inputs' device), every other attribute is torchvision.ops'."""

    @staticmethod
    def box_iou(boxes1, boxes2):
        if not torch.cuda.is_available():
            raise RuntimeError("cvpce_b200.patch: box_iou needs a CUDA device; no CPU fallback")
        dev = torch.device("cuda", torch.cuda.current_device())
        return ops.box_iou(boxes1.to(dev), boxes2.to(dev)).to(boxes1.device)

    def __getattr__(self, name):
        return getattr(torchvision.ops, name)
