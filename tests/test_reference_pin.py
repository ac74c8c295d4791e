"""CPU, authoring container only (skipped where /root/reference is absent, e.g. on the GPU box): pins the oracle and the
torch port against the UNMODIFIED reference on fresh random inputs, beyond the committed golden vectors."""
import warnings

import numpy as np
import pytest
import torch

import oracle
from oracle import shim, torch_port
from cvpce_b200 import synth

pytestmark = pytest.mark.skipif(not shim.available(), reason="reference tree not present")


@pytest.fixture(scope="module")
def ref():
    warnings.simplefilter("ignore")
    return shim.load()


def test_crop_loop_matches_reference(ref):
    img = torch.from_numpy(synth.shelf_image(3, 200, 260))
    boxes = torch.tensor([[10.9, 20.1, 120.5, 180.7], [0, 0, 260, 200], [100, 50, 101, 150], [30, 30, 90, 60]])
    want = torch.stack([ref.datautils.resize_for_classification(img[:, y1:y2, x1:x2]) for x1, y1, x2, y2 in boxes.to(dtype=torch.long)])
    got, _ = oracle.crop_resize(img.numpy(), boxes.numpy(), 256, True)
    np.testing.assert_allclose(got, want.numpy(), rtol=0, atol=1e-6)
    assert torch.equal(torch_port.crops_for_boxes(img, boxes), want)


def test_encoder_and_nn_port_match_reference(ref):
    torch.manual_seed(0)
    enc_ref = ref.models.classification.macvgg_embedder("vgg16", pretrained=False).eval()
    torch.manual_seed(0)
    enc_port = torch_port.MACVGGPort().eval()
    x = torch.from_numpy(synth.encoder_inputs(2, 64, 5))
    with torch.no_grad():
        assert torch.equal(enc_ref(x), enc_port(x))
    g, q = synth.nn_case(20, 90, 128, 77)
    want = ref.models.classification.nearest_neighbors(torch.from_numpy(g), torch.from_numpy(q), 4)
    assert torch.equal(torch_port.nearest_neighbors(torch.from_numpy(g), torch.from_numpy(q), 4), want)
    assert torch.equal(torch_port.nearest_neighbors_chunked(torch.from_numpy(g), torch.from_numpy(q), 4, 7), want)
    assert np.array_equal(oracle.nearest_neighbors(g, q, 4), want.numpy())


def test_reference_own_tests_pass_through_shim(ref, tmp_path):
    """The reference's hot-path test (test/models/classification_test.py) run unmodified, via a `thesis` alias."""
    import importlib.util
    import sys
    sys.modules["thesis"] = ref
    sys.modules["thesis.models"] = ref.models
    sys.modules["thesis.models.classification"] = ref.models.classification
    spec = importlib.util.spec_from_file_location("ref_classification_test", shim.REFERENCE_ROOT + "/test/models/classification_test.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.test_nearest_neigbors()


def test_matching_oracle_equals_reference_on_fresh_inputs(ref):
    """oracle box_iou / iou_match against the live metrics.iou_matrices + check_matches (metrics.py:11-34) on cases that
    are not in the golden file, and the reference's own metrics tests run unmodified through the shim."""
    import importlib.util
    import sys
    import cvpce.metrics as metrics
    for seed, thr in ((91, 0.5), (92, 0.3), (93, 0.9)):
        t, p = synth.detection_eval_case(seed, 25, 40, duplicates=4, extent=500.0)
        ious, idx = metrics.iou_matrices(torch.from_numpy(t), torch.from_numpy(p))
        tp, fp = metrics.check_matches(ious, idx, thr)
        assert np.array_equal(-np.sort(-oracle.box_iou(p, t), axis=1), ious.numpy())
        got, _ = oracle.iou_match(p, t, thr)
        assert np.array_equal(got, tp.numpy().astype(bool)), (seed, thr)
    sys.modules["thesis"] = ref
    sys.modules["thesis.metrics"] = metrics
    spec = importlib.util.spec_from_file_location("ref_metrics_test", shim.REFERENCE_ROOT + "/test/metrics_test.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.test_iou_matrices()
    mod.test_iou_matrices_2()
    mod.test_check_matches()


def test_patch_swaps_and_restores_every_call_site(ref):
    """cvpce_b200.patch on the UNMODIFIED reference: every hot-path call site is swapped, the swapped hot-path functions refuse
    CPU tensors loudly (no CPU fallback), the two shared helpers that the reference also calls from host-side data preparation
    (resize_for_classification in dataset workers, torchvision's batched_nms in other detectors) hand host tensors to the
    ORIGINAL function, and revert() puts back exactly the original objects."""
    import cvpce.metrics as rmetrics
    import cvpce.planograms as rplan
    import cvpce.models.proposals as rprop
    from torchvision.models.detection import retinanet
    from cvpce_b200 import patch
    rc, rp, rd = ref.models.classification, ref.production, ref.datautils
    sites = [(rc, "nearest_neighbors"), (rp, "nearest_neighbors"), (ref.utils, "nearest_neighbors"), (rc.MACVGG, "forward"),
             (rc.MACResNet, "forward"), (rd, "resize_for_classification"), (rp.ProposalGenerator, "generate_proposals_and_images"),
             (rp.Classifier, "classify"), (rmetrics, "iou_matrices"), (rmetrics, "_process_one"), (rmetrics, "calculate_metrics"),
             (rplan, "tvops"), (retinanet.box_ops, "batched_nms")]
    before = [getattr(o, n) for o, n in sites]
    gln_own = "postprocess_detections" in vars(rprop.GaussianLayerNetwork)
    patch.apply(ref)
    try:
        patch.apply(ref)  # idempotent
        after = [getattr(o, n) for o, n in sites]
        assert all(a is not b for a, b in zip(after, before)), [n for (o, n), a, b in zip(sites, after, before) if a is b]
        assert "postprocess_detections" in vars(rprop.GaussianLayerNetwork)
        b = torch.tensor([[0., 0., 10., 10.], [1., 1., 11., 11.]])
        with pytest.raises(RuntimeError):
            rc.nearest_neighbors(torch.rand(4, 8), torch.rand(2, 8), 1)
        with pytest.raises(RuntimeError):
            rc.MACVGG.forward(rc.macvgg_embedder('vgg16', pretrained=False).eval(), torch.rand(1, 3, 64, 64))
        # host-side callers of the two shared helpers keep the reference's / torchvision's own arithmetic, bit for bit
        sc, lab = torch.tensor([.9, .8]), torch.zeros(2, dtype=torch.int64)
        assert torch.equal(retinanet.box_ops.batched_nms(b, sc, lab, 0.5), before[sites.index((retinanet.box_ops, "batched_nms"))](b, sc, lab, 0.5))
        host_img = torch.rand(3, 20, 30, generator=torch.Generator().manual_seed(1))
        assert torch.equal(rd.resize_for_classification(host_img), before[sites.index((rd, "resize_for_classification"))](host_img))
        # ... so the gallery dataset's __getitem__ (datautils.py:511, run in DataLoader workers under Classifier.build_index)
        # and PlanogramComparator.compare on a host image (production.py:111) work with the patch applied
        if not torch.cuda.is_available():   # the GPU-backed metrics / box_iou replacements say so instead of computing on the host
            with pytest.raises(RuntimeError):
                rmetrics.iou_matrices(b, b)
            with pytest.raises(RuntimeError):
                rplan.tvops.box_iou(b, b)
        assert rplan.tvops.clip_boxes_to_image is torchvision_ops().clip_boxes_to_image   # everything else passes through
    finally:
        patch.revert()
    assert all(getattr(o, n) is b for (o, n), b in zip(sites, before))
    assert ("postprocess_detections" in vars(rprop.GaussianLayerNetwork)) == gln_own
    # the reference computes again after the revert
    assert rc.nearest_neighbors(torch.eye(3), torch.eye(3)[[2, 0]], 1).flatten().tolist() == [2, 0]


def torchvision_ops():
    import torchvision
    return torchvision.ops


def test_eval_entry_points_reproduce_the_golden(ref):
    """a14: tests/golden/eval.npz IS what the unmodified eval_dihe / evaluate_detections return on the synthetic evaluation set
    (re-run here against the live reference), and the AP-bookkeeping restatement the GPU test hands to the mirror agrees with
    the reference's metrics._do_calculate."""
    import os
    import cvpce.metrics as metrics
    from oracle import make_golden, torch_port
    want = np.load(os.path.join(os.path.dirname(__file__), "golden", "eval.npz"))
    got = make_golden.run_reference_evals(ref)
    assert sorted(got) == sorted(want.files)
    for k in want.files:
        np.testing.assert_array_equal(got[k], want[k], err_msg=k)
    rs = np.random.RandomState(5)
    for n_img in (1, 4):
        thr = (0.5, 0.75)
        m = {t: {"true_positives": [], "false_positives": [], "recall_300": []} for t in thr}
        conf, total = [], 0
        for _ in range(n_img):
            n = int(rs.randint(1, 30))
            c = torch.sort(torch.from_numpy(rs.rand(n).astype(np.float32)), descending=True).values
            conf.append(c)
            nt = int(rs.randint(0, 12))
            total += nt
            for t in thr:
                tp = torch.from_numpy((rs.rand(n) < 0.5).astype(np.float32))
                m[t]["true_positives"].append(tp)
                m[t]["false_positives"].append(1 - tp)
                m[t]["recall_300"].append(tp.cumsum(0)[:300][-1] / nt if nt else torch.tensor(0.0))
        a = metrics._do_calculate(thr, m, conf, total)
        b = torch_port.do_calculate(thr, m, conf, total)
        for t in thr:
            for key in ("f", "p", "r", "c", "ap", "ar_300"):
                assert float(a[t][key]) == float(b[t][key]), (t, key)
            for key in ("p", "r", "f", "c"):
                assert torch.equal(a[t]["raw"][key], b[t]["raw"][key])
