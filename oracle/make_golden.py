"""Writes tests/golden/*.npz by running the UNMODIFIED reference (through oracle/shim.py) and the compiled third-party ops
it calls (torchvision.ops.nms / batched_nms, F.interpolate) on seeded synthetic inputs.

Run in the authoring container only (needs /root/reference):   python -m oracle.make_golden
Inputs are regenerated from numpy RandomState seeds by oracle/synth.py, so only the small outputs are stored.
"""
import os
import sys
import warnings

import numpy as np
import torch
import torchvision

from . import shim
from cvpce_b200 import synth

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def golden_nms():
    cases = {}
    for name, (boxes, scores, thr, per_img) in synth.nms_cases().items():
        b, s = torch.from_numpy(boxes), torch.from_numpy(scores)
        keep = torchvision.ops.nms(b, s, thr)
        # the RetinaNet call site: batched_nms with a single class, then [:detections_per_img]  (tv:retinanet.py:559-561)
        keep_b = torchvision.ops.batched_nms(b, s, torch.zeros(len(s), dtype=torch.int64), thr)[:per_img]
        cases[name + "/keep"] = keep.numpy()
        if torch.equal(keep[:per_img], keep_b):
            cases[name + "/keep_postprocess"] = keep_b.numpy()
        else:
            # batched_nms' "vanilla" branch (CPU, numel > 4000) re-sorts the kept indices by score with an UNSTABLE sort:
            # same set, same score sequence, but equal scores may come in another index order. Not stored as golden.
            assert name == "random2000_ties" and torch.equal(s[keep[:per_img]], s[keep_b])
    np.savez_compressed(os.path.join(OUT, "nms.npz"), **cases)
    print("nms:", {k: v.shape for k, v in cases.items()})


def golden_crops(ref):
    out = {}
    img = torch.from_numpy(synth.shelf_image(synth.CROP_IMAGE_SEED, *synth.CROP_IMAGE_HW))
    boxes = torch.from_numpy(synth.crop_boxes())
    for S in (256, 224):
        ref.datautils.CLASSIFICATION_IMAGE_SIZE = S
        pix = synth.sample_pixels(S)
        vals = []
        for x1, y1, x2, y2 in boxes.to(dtype=torch.long):  # production.py:20
            r = ref.datautils.resize_for_classification(img[:, y1:y2, x1:x2])
            vals.append(r.reshape(-1)[pix].numpy())
        out["aa_S%d" % S] = np.stack(vals)
    ref.datautils.CLASSIFICATION_IMAGE_SIZE = 256
    # thesis-era semantics (torchvision 0.9: no antialias): same canvas, F.interpolate(antialias=False)
    pix = synth.sample_pixels(256)
    vals = []
    for x1, y1, x2, y2 in boxes.to(dtype=torch.long):
        c = img[:, y1:y2, x1:x2]
        _, h, w = c.shape
        L = max(h, w)
        canvas = torch.full((3, L, L), 0.5)
        canvas[:, :h, :w] = c
        r = torch.nn.functional.interpolate(canvas[None], size=(256, 256), mode="bilinear", align_corners=False, antialias=False)[0]
        vals.append(r.reshape(-1)[pix].numpy())
    out["noaa_S256"] = np.stack(vals)
    np.savez_compressed(os.path.join(OUT, "crops.npz"), **out)
    print("crops:", {k: v.shape for k, v in out.items()})


def golden_mac(ref):
    out = {}
    # (1) pooling + concat + L2 of MACVGG.forward on given feature maps: the trunk blocks are swapped for stand-ins that
    # return seeded maps, so the reference's own forward (classification.py:46-51) does the pooling.
    f4, f5 = synth.feature_maps()
    enc = ref.models.classification.MACVGG(batch_norm=False)

    class Give(torch.nn.Module):
        def __init__(self, t):
            super().__init__()
            self.t = t

        def forward(self, x):
            return self.t

    enc.block1, enc.block2 = Give(torch.from_numpy(f4)), Give(torch.from_numpy(f5))
    out["mac_given_maps"] = enc(torch.zeros(f4.shape[0], 3, 8, 8)).numpy()
    # (2) the whole encoder with seeded random-init VGG16 weights (no BN, as production uses): embeddings of seeded inputs
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = ref.models.classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    x = torch.from_numpy(synth.encoder_inputs())
    with torch.no_grad():
        out["macvgg_embeddings"] = enc(x).numpy()
    w0 = enc.block1[0].weight.detach().reshape(-1)
    out["macvgg_weight_probe"] = torch.cat([w0[:8], enc.block2[-2].weight.detach().reshape(-1)[:8]]).numpy()
    # (3) MACResNet (classification.py:53-85) over a seeded random-init resnet50 (BatchNorm in eval mode), descriptor
    # layers (2, 3): 512 + 1024 channels -> 1536-d.  The reference's own factory (classification.py:111-121) passes
    # constructor arguments the installed torchvision no longer accepts; the class itself is used unchanged.
    from torchvision.models import resnet
    torch.manual_seed(synth.WEIGHT_SEED + 1)
    enc = ref.models.classification.MACResNet(resnet.resnet50(weights=None), [2, 3]).eval().requires_grad_(False)
    with torch.no_grad():
        out["macresnet_embeddings"] = enc(x).numpy()
    out["macresnet_weight_probe"] = enc.blocks[0][0][0].weight.detach().reshape(-1)[:8].numpy()
    np.savez_compressed(os.path.join(OUT, "mac.npz"), **out)
    print("mac:", {k: v.shape for k, v in out.items()})


def golden_nn(ref):
    out = {}
    nn = ref.models.classification.nearest_neighbors
    a, q, expected = synth.nn_kat()
    got = nn(torch.from_numpy(a), torch.from_numpy(q))[:, 0].numpy()
    assert np.array_equal(got, expected), (got, expected)  # reference test/models/classification_test.py:7-28
    out["kat_top1"] = got
    for name, (Q, G, D, k, seed) in synth.NN_CASES.items():
        g, qq = synth.nn_case(Q, G, D, seed)
        out[name] = nn(torch.from_numpy(g), torch.from_numpy(qq), k).numpy()
    np.savez_compressed(os.path.join(OUT, "nn.npz"), **out)
    print("nn:", {k: v.shape for k, v in out.items()})


def golden_classify(ref):
    """Classifier.build_index + classify (production.py:36-74) on a seeded toy gallery, CPU devices."""
    out = {}
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = ref.models.classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    gallery, queries = synth.classify_case()
    cpu = torch.device("cpu")
    clf = ref.production.Classifier(enc, synth.GalleryDataset(gallery), device=cpu, emb_device=cpu, batch_size=4,
                                    num_workers=0, k=3)
    res, emb = clf.classify(torch.from_numpy(queries), return_embedding=True)
    out["gallery_embedding"] = clf.embedding.numpy()
    out["query_embedding"] = emb.numpy()
    out["labels"] = np.array([[int(a[1:]) for a in row] for row in res], dtype=np.int64)  # annotations are "g<i>"
    np.savez_compressed(os.path.join(OUT, "classify.npz"), **out)
    print("classify:", {k: v.shape for k, v in out.items()}, out["labels"].tolist())


def golden_head(ref):
    """RetinaNet.postprocess_detections as inherited by the reference's GaussianLayerNetwork (proposals.py:161-182;
    tv:retinanet.py:508-571) on seeded head outputs.  The arguments of the batched_nms call inside it are recorded too:
    they are the pre-NMS candidates (boxes, scores, labels after sigmoid / threshold / top-k / decode / clip / cat)."""
    from torchvision.models.detection import retinanet
    import cvpce.models.proposals as proposals
    out = {}
    for name, (seed, sizes, hw, C, mu, sd, q, topk) in synth.HEAD_CASES.items():
        logits, regression, anchors, sizes, C, topk = synth.head_case(name)
        model = proposals.gln(num_classes=C, pretrained_backbone=False).eval()
        model.topk_candidates = topk
        B = len(sizes)
        head = {"cls_logits": [torch.from_numpy(l) for l in logits], "bbox_regression": [torch.from_numpy(r) for r in regression]}
        anc = [[torch.from_numpy(a) for a in anchors] for _ in range(B)]
        captured = []
        real = retinanet.box_ops.batched_nms

        def spy(boxes, scores, idxs, thr):
            captured.append((boxes.clone(), scores.clone(), idxs.clone()))
            return real(boxes, scores, idxs, thr)

        retinanet.box_ops.batched_nms = spy
        try:
            with torch.no_grad():
                dets = model.postprocess_detections(head, anc, [tuple(s) for s in sizes])
        finally:
            retinanet.box_ops.batched_nms = real
        for b in range(B):
            cb, cs, cl = captured[b]
            pre = "%s/%d/" % (name, b)
            if name == "fullsize":   # keep the fixture small: candidates as count + per-level sums only
                out[pre + "cand_count"] = np.array([len(cs)], dtype=np.int64)
                out[pre + "cand_score_sum"] = np.array([cs.double().sum().item()])
            else:
                out[pre + "cand_boxes"] = cb.numpy()
                out[pre + "cand_scores"] = cs.numpy()
                out[pre + "cand_labels"] = cl.numpy()
            out[pre + "det_boxes"] = dets[b]["boxes"].numpy()
            out[pre + "det_scores"] = dets[b]["scores"].numpy()
            out[pre + "det_labels"] = dets[b]["labels"].numpy()
    np.savez_compressed(os.path.join(OUT, "head.npz"), **out)
    print("head:", {k: v.shape for k, v in out.items() if k.endswith(("cand_scores", "det_scores", "cand_count"))})


def golden_match(ref):
    """the reference's detection-evaluation matching (metrics.iou_matrices + check_matches) and the
    expected-vs-detected matching inside planograms.finalize_via_ransac (run for real, cv2 RANSAC included; the projected
    positions it matches on are stored because they come out of cv2)."""
    import cvpce.metrics as metrics
    import cvpce.planograms as planograms
    out = {}
    for name, (targets, preds) in synth.match_cases().items():
        t, p = torch.from_numpy(targets), torch.from_numpy(preds)
        ious, idx = metrics.iou_matrices(t, p)                         # metrics.py:11-13
        out[name + "/iou"] = torchvision.ops.box_iou(p, t).numpy()
        out[name + "/sorted_iou"] = ious.numpy()
        for thr in synth.MATCH_THRESHOLDS:
            tp, fp = metrics.check_matches(ious, idx, thr)             # metrics.py:15-34
            assert torch.equal(tp + fp, torch.ones_like(tp))
            out[name + "/tp_%g" % thr] = tp.numpy().astype(np.uint8)
    for seed in (51, 52):
        b1, l1, b2, l2, solution = synth.planogram_case(seed)
        for thr in (0.5, 0.8):
            res = planograms.finalize_via_ransac(solution, torch.from_numpy(b1), torch.from_numpy(b2), l1, l2, iou_threshold=thr,
                                                 return_matched_actual=True, return_expected_positions=True)
            matched_expected, matched_actual, expected_positions, missing_expected = res[0], res[1], res[2], res[3]
            key = "planogram%d_%g" % (seed, thr)
            out[key + "/expected_positions"] = expected_positions.numpy()
            out[key + "/matched_expected"] = matched_expected.numpy().astype(np.uint8)
            out[key + "/matched_actual"] = matched_actual.numpy().astype(np.uint8)
            out[key + "/missing_expected"] = missing_expected.numpy()
    np.savez_compressed(os.path.join(OUT, "match.npz"), **out)
    print("match:", {k: int(v.sum()) for k, v in out.items() if "/tp_" in k or "/matched_" in k})


def run_reference_evals(ref):
    """The reference's two evaluation entry points on the synthetic evaluation set, CPU:
    cvpce/classification_eval.py:6-56 eval_dihe and cvpce/detection_eval.py:6-49 evaluate_detections run UNMODIFIED; only what
    pins them to a GPU is overridden -- Classifier's device defaults (production.py:23) and SKU110KBatch.cuda
    (datautils.py:124-127) -- plus CLASSIFICATION_IMAGE_SIZE = 64 to keep the CPU run short.  The network body of the detector
    is a stand-in (a random-init GLN detects nothing) that feeds seeded head outputs to the reference's own
    GaussianLayerNetwork.postprocess_detections.  Returns the dict stored as tests/golden/eval.npz."""
    import functools
    import types
    import cvpce.classification_eval as ce
    import cvpce.detection_eval as de
    import cvpce.models.proposals as proposals
    from . import eval_harness as eh
    cpu = torch.device("cpu")
    out = {}
    gallery, annotations, shelves = synth.eval_case()
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = ref.models.classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    predicted = []

    class SpyClassifier(ref.production.Classifier):
        def classify(self, images, return_embedding=False):
            res = super().classify(images, return_embedding)
            predicted.append(res[0] if return_embedding else res)
            return res

    saved = (ce.production, de.production, ref.datautils.CLASSIFICATION_IMAGE_SIZE, ref.datautils.SKU110KBatch.cuda)
    stand_in = types.SimpleNamespace(Classifier=functools.partial(SpyClassifier, device=cpu, emb_device=cpu))
    try:
        ce.production = de.production = stand_in
        ref.datautils.CLASSIFICATION_IMAGE_SIZE = synth.EVAL_CROP_SIZE
        ref.datautils.SKU110KBatch.cuda = lambda self, device=None, non_blocking=True: self
        # ---- eval_dihe
        k = (1, 3)
        acc = ce.eval_dihe(enc, synth.GalleryDataset(gallery), eh.DiheTestSet(annotations, shelves), batch_size=4, num_workers=0,
                           k=k, verbose=False)
        out["dihe/accuracy"] = np.array([acc[x] for x in k], dtype=np.float64)
        out["dihe/predicted"] = np.array([[annotations.index(a) for a in row] for per_img in predicted for row in per_img], dtype=np.int64)
        # ---- evaluate_detections
        predicted.clear()
        gln = proposals.gln(num_classes=1, pretrained_backbone=False).eval()
        gln.detections_per_img = 200
        heads = synth.eval_detector_heads(shelves)
        det = eh.StubDetector(heads, gln.postprocess_detections)
        per_class, overall = de.evaluate_detections(det, enc, eh.DetectionTestSet(annotations, shelves), synth.GalleryDataset(gallery),
                                                    thresholds=synth.EVAL_THRESHOLDS, proposal_batch_size=2,
                                                    classification_batch_size=4, num_workers=0)
        for name, arr in eh.flatten_metrics(per_class, overall, synth.EVAL_THRESHOLDS).items():
            out["det/" + name] = arr
        out["det/predicted_top1"] = np.array([annotations.index(row[0]) for per_img in predicted for row in per_img], dtype=np.int64)
        mam = de.mean_average_metrics(per_class, synth.EVAL_THRESHOLDS)
        out["det/map_mar300"] = np.array([[float(mam[t]["map"]), float(mam[t]["mar300"])] for t in synth.EVAL_THRESHOLDS])
    finally:
        ce.production, de.production, ref.datautils.CLASSIFICATION_IMAGE_SIZE, ref.datautils.SKU110KBatch.cuda = saved
    return out


def golden_eval(ref):
    out = run_reference_evals(ref)
    np.savez_compressed(os.path.join(OUT, "eval.npz"), **out)
    print("eval: dihe accuracy", out["dihe/accuracy"].tolist(), "detections classified", len(out["det/predicted_top1"]),
          "mAP / mAR300 per threshold", out["det/map_mar300"].tolist())


def main():
    if not shim.available():
        sys.exit("reference tree not present; golden vectors can only be regenerated in the authoring container")
    os.makedirs(OUT, exist_ok=True)
    warnings.simplefilter("ignore")
    ref = shim.load()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    golden_nms()
    golden_crops(ref)
    golden_mac(ref)
    golden_nn(ref)
    golden_classify(ref)
    golden_head(ref)
    golden_match(ref)
    golden_eval(ref)
    with open(os.path.join(OUT, "VERSIONS.txt"), "w") as f:
        f.write("torch %s\ntorchvision %s\nnumpy %s\nreference commit 7509e7d (laitalaj/cvpce)\n" %
                (torch.__version__, torchvision.__version__, np.__version__))


if __name__ == "__main__":
    main()
