"""Synthetic evaluation set + stand-in detector for the a14 integration goldens (eval_dihe, evaluate_detections).
TEST INFRASTRUCTURE ONLY: used by oracle/make_golden.py (with the live reference, CPU) and by tests/ (with the mirrors, GPU)."""
import numpy as np
import torch

from cvpce_b200 import synth


class DiheTestSet:
    """Iterates (img f32 [3,H,W] in [0,1], target annotation strings, integer boxes [n,4]) like the reference's Grocery Products
    test set does for classification_eval.eval_dihe (classification_eval.py:19)."""

    def __init__(self, annotations, shelves):
        self.items = [(torch.from_numpy(img), [annotations[l] for l in labels], torch.from_numpy(boxes).to(torch.int64))
                      for img, boxes, labels in shelves]

    def __iter__(self):
        return iter(self.items)

    def __len__(self):
        return len(self.items)


class DetectionTestSet(torch.utils.data.Dataset):
    """(img, {'boxes', 'labels'}) samples with the int_to_ann / ann_to_int tables detection_eval.evaluate_detections reads."""

    def __init__(self, annotations, shelves):
        H, W = synth.EVAL_IMAGE_HW
        self.int_to_ann = list(annotations)
        self.ann_to_int = {a: i for i, a in enumerate(annotations)}
        self.items = [(torch.from_numpy(img), {"boxes": torch.from_numpy(np.clip(boxes, 0, [W, H, W, H]).astype(np.float32)),
                                               "labels": torch.from_numpy(labels)}) for img, boxes, labels in shelves]

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        return self.items[i]


class StubDetector(torch.nn.Module):
    """Stands in for the GLN network body: hands seeded head outputs (synth.eval_detector_heads) to a post-processing function
    with RetinaNet.postprocess_detections' signature -- the reference's own (inherited by GaussianLayerNetwork) for the golden,
    cvpce_b200.detection's for the GPU test.  Images must arrive in data-set order."""

    def __init__(self, heads, postprocess):
        super().__init__()
        self.heads, self.postprocess, self.cursor = heads, postprocess, 0

    def forward(self, images):
        out = []
        for img in images:
            lg, rg, anc = self.heads[self.cursor]
            self.cursor += 1
            dev = img.device
            n_per = [a.shape[0] for a in anc]
            logits = list(torch.from_numpy(lg).to(dev)[None].split(n_per, 1))
            regression = list(torch.from_numpy(rg).to(dev)[None].split(n_per, 1))
            anchors = [[torch.from_numpy(a).to(dev) for a in anc]]
            out += self.postprocess({"cls_logits": logits, "bbox_regression": regression}, anchors, [tuple(img.shape[-2:])])
        return out


def flatten_metrics(per_class, overall, thresholds):
    """{class: {thr: {...}}}, {thr: {...}} -> dict of float arrays [n_classes + 1, n_thresholds] per metric (last row = overall)."""
    keys = ("f", "p", "r", "c", "ap", "ar_300")
    rows = [per_class[c] for c in sorted(per_class)] + [overall]
    return {k: np.array([[float(r[t][k]) for t in thresholds] for r in rows], dtype=np.float64) for k in keys}
