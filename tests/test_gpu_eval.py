"""a14: the two entry points that drive the hot path -- eval_dihe (classification_eval.py:6-56) and
evaluate_detections (detection_eval.py:6-49) -- through the cvpce_b200 mirrors on the GPU, against tests/golden/eval.npz,
which oracle/make_golden.py wrote by running the UNMODIFIED reference functions on the same synthetic evaluation set on the CPU
(tests/test_reference_pin.py re-runs that comparison against the live reference where it is present)."""
import numpy as np
import pytest
import torch

from cvpce_b200 import synth

pytestmark = pytest.mark.gpu
GOLDEN = np.load(__import__("os").path.join(__import__("os").path.dirname(__file__), "golden", "eval.npz"))


@pytest.fixture()
def setup(cuda):
    from cvpce_b200 import classification, datautils
    from oracle import eval_harness as eh
    saved = (torch.backends.cudnn.allow_tf32, datautils.CLASSIFICATION_IMAGE_SIZE)
    torch.backends.cudnn.allow_tf32 = False
    datautils.CLASSIFICATION_IMAGE_SIZE = synth.EVAL_CROP_SIZE
    gallery, annotations, shelves = synth.eval_case()
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(cuda)
    yield enc, gallery, annotations, shelves, eh
    torch.backends.cudnn.allow_tf32, datautils.CLASSIFICATION_IMAGE_SIZE = saved


def _spy(production, annotations, sink):
    class Spy(production.Classifier):
        def classify(self, images, return_embedding=False):
            res = super().classify(images, return_embedding)
            sink.extend([annotations.index(a) for a in row] for row in (res[0] if return_embedding else res))
            return res
    return Spy


def test_eval_dihe_mirror_reproduces_the_reference_accuracy(setup, monkeypatch):
    from cvpce_b200 import classification_eval, production
    enc, gallery, annotations, shelves, eh = setup
    predicted = []
    monkeypatch.setattr(classification_eval.production, "Classifier", _spy(production, annotations, predicted))
    acc = classification_eval.eval_dihe(enc, synth.GalleryDataset(gallery), eh.DiheTestSet(annotations, shelves), batch_size=4,
                                        num_workers=0, k=(1, 3), verbose=False)
    assert np.array_equal(np.array(predicted), GOLDEN["dihe/predicted"])          # every crop's top-3 annotations
    assert [acc[1], acc[3]] == GOLDEN["dihe/accuracy"].tolist()
    assert enc.block1[0].weight.requires_grad                                      # classification_eval.py:47 restores it


def test_evaluate_detections_mirror_reproduces_the_reference_metrics(setup, monkeypatch):
    from cvpce_b200 import detection, detection_eval, production
    from oracle import torch_port
    enc, gallery, annotations, shelves, eh = setup
    predicted = []
    monkeypatch.setattr(detection_eval.production, "Classifier", _spy(production, annotations, predicted))

    def postprocess(head_outputs, anchors, image_shapes):    # the GLN's thresholds (proposals.py:162-166, detections_per_img 200)
        return detection.postprocess_detections(head_outputs, anchors, image_shapes, 0.05, 1000, 0.5, 200)

    det = eh.StubDetector(synth.eval_detector_heads(shelves), postprocess)
    per_class, overall = detection_eval.evaluate_detections(det, enc, eh.DetectionTestSet(annotations, shelves), synth.GalleryDataset(gallery),
                                                            thresholds=synth.EVAL_THRESHOLDS, proposal_batch_size=2,
                                                            classification_batch_size=4, num_workers=0, do_calculate=torch_port.do_calculate)
    assert np.array_equal(np.array([p[0] for p in predicted]), GOLDEN["det/predicted_top1"])   # same detections, same classes
    got = eh.flatten_metrics(per_class, overall, synth.EVAL_THRESHOLDS)
    for k, arr in got.items():
        np.testing.assert_allclose(arr, GOLDEN["det/" + k], rtol=1e-6, atol=1e-7, err_msg=k)
    mam = detection_eval.mean_average_metrics(per_class, synth.EVAL_THRESHOLDS)
    np.testing.assert_allclose([[mam[t]["map"], mam[t]["mar300"]] for t in synth.EVAL_THRESHOLDS], GOLDEN["det/map_mar300"], rtol=1e-6)
