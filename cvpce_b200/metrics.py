"""Host-side mirror of the reference's detection-evaluation matching (cvpce/metrics.py:11-34, 80-95) on the sm_100a kernels.

The reference builds the [P,T] IoU matrix, sorts every row and walks it in two nested Python loops per image and threshold;
here the IoU matrix is one launch (`iou_matrices`) and the claim pass is one launch for a whole list of images
(`process_batch`), with no matrix and no sort.  Precision / recall / AP bookkeeping (metrics.py:36-78, 97-139) is not on the
path and stays in the reference: `calculate_metrics` hands the per-image results to the reference's `_do_calculate`.

The reference calls these with CPU tensors (detection_eval.py:34-36 moves everything to the host first); tensors are copied
to the current CUDA device for the arithmetic and results come back on the caller's device -- there is no CPU arithmetic
path and a machine without a CUDA device raises.
"""
import torch

from . import ops


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("cvpce_b200.metrics: no CUDA device (cvpce_b200 has no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def iou_matrices(targets, sorted_predictions):
    """metrics.py:11-13: (IoU rows sorted descending, their column indices) of predictions (rows) against targets."""
    dev = _device()
    ious = ops.box_iou(sorted_predictions.to(dev), targets.to(dev))
    s, i = torch.sort(ious, dim=1, descending=True)
    return s.to(sorted_predictions.device), i.to(sorted_predictions.device)


def match_detections(targets, sorted_predictions, iou_threshold=0.5):
    """iou_matrices + check_matches (metrics.py:11-34) fused: (true_positive f32 [P], false_positive f32 [P])."""
    dev = _device()
    tp, _ = ops.iou_match(sorted_predictions.to(dev), targets.to(dev), iou_threshold)
    tp = tp.to(device=sorted_predictions.device, dtype=torch.float32)
    return tp, 1.0 - tp


def _recall_300(tp, n_targets):
    # metrics.py:46-55 + :91: recall after each prediction, read at the 300th (or last) one; 0 for an image without predictions
    if tp.numel() == 0:
        return 0
    r = tp.cumsum(0) / n_targets if n_targets > 0 else torch.zeros_like(tp)
    return r[:300][-1]


def process_batch(targets, predictions, confidences, iou_thresholds=(0.5,)):
    """metrics._process_one (metrics.py:80-95) for a list of images at once: per image
    ({thr: {'true_positives', 'false_positives', 'recall_300'}}, confidences sorted descending, number of targets).
    One sort per image on the caller's device (as the reference does), then ONE claim launch per threshold for all images."""
    dev = _device()
    n = len(targets)
    if n == 0:
        return []
    sorted_conf, sorted_pred = [], []
    for p, c in zip(predictions, confidences):
        c, order = torch.sort(c, descending=True)               # metrics.py:81-82
        sorted_conf.append(c)
        sorted_pred.append(p[order].reshape(-1, 4))
    rows = torch.cat(sorted_pred).to(dev, torch.float32)
    cols = torch.cat([t.reshape(-1, 4) for t in targets]).to(dev, torch.float32)
    n_rows = [int(p.shape[0]) for p in sorted_pred]
    n_cols = [int(t.reshape(-1, 4).shape[0]) for t in targets]
    ro = torch.tensor([0] + n_rows, dtype=torch.int64).cumsum(0).to(torch.int32).to(dev)
    co = torch.tensor([0] + n_cols, dtype=torch.int64).cumsum(0).to(torch.int32).to(dev)
    per_thr = {}
    for t in iou_thresholds:
        tp, _ = ops.iou_match(rows, cols, t, row_offsets=ro, col_offsets=co, max_cols=max(n_cols))
        per_thr[t] = tp.to(torch.float32).cpu().split(n_rows)    # one D2H per threshold
    out = []
    for i in range(n):
        m = {}
        for t in iou_thresholds:
            tp = per_thr[t][i].to(sorted_conf[i].device)
            m[t] = {"true_positives": tp, "false_positives": 1.0 - tp, "recall_300": _recall_300(tp, n_cols[i])}
        out.append((m, sorted_conf[i], n_cols[i]))
    return out


def process_one(target, prediction, confidence, iou_thresholds):
    """Drop-in for metrics._process_one (metrics.py:80-95)."""
    return process_batch([target], [prediction], [confidence], iou_thresholds)[0]


def calculate_metrics(targets, predictions, confidences, iou_thresholds=(0.5,), do_calculate=None):
    """Drop-in for metrics.calculate_metrics (metrics.py:126-139): matching for ALL images in one launch per threshold, then
    the reference's own `_do_calculate` (pass it as `do_calculate`; cvpce_b200.patch does) for precision / recall / AP."""
    if do_calculate is None:
        raise RuntimeError("cvpce_b200.metrics.calculate_metrics: pass the reference's metrics._do_calculate (AP bookkeeping is "
                           "not part of the hot path and is not rebuilt)")
    matches_for_threshold = {t: {"true_positives": [], "false_positives": [], "recall_300": []} for t in iou_thresholds}
    sorted_confidences = []
    total_targets = 0
    for matches, conf, n_t in process_batch(list(targets), list(predictions), list(confidences), iou_thresholds):
        sorted_confidences.append(conf)
        total_targets += n_t
        for t in iou_thresholds:
            for k in ("true_positives", "false_positives", "recall_300"):
                matches_for_threshold[t][k].append(matches[t][k])
    return do_calculate(iou_thresholds, matches_for_threshold, sorted_confidences, total_targets)
