"""Host-side mirror of cvpce/classification_eval.py (eval_dihe): the entry point that drives crop -> embed -> search per test
image.  Same arguments and return value; the per-box python crop loop (classification_eval.py:24) is one
launch of the crop kernel per image and Classifier.classify searches all crops of the image at once."""
from collections import Counter

import torch
from torchvision import ops as tvops

from . import datautils, production


def eval_dihe(encoder, sampleset, testset, batch_size, num_workers, k=(1,), verbose=True, device=torch.device('cuda'),
              emb_device=None, load=None):
    """classification_eval.py:6-56.  testset yields (img f32 [3,H,W] in [0,1], target annotations, boxes [n,4]).
    Returns {knn: accuracy}."""
    emb_device = device if emb_device is None else emb_device
    encoder.requires_grad_(False)
    classifier = production.Classifier(encoder, sampleset, device=device, emb_device=emb_device, batch_size=batch_size,
                                       num_workers=num_workers, k=max(k), load=load)
    total = 0
    correct = Counter()
    per_annotation = Counter()
    missed = Counter()
    mistaken_for = {}
    for i, (img, target_anns, boxes) in enumerate(testset):
        if verbose and i % 10 == 0:
            print(f'{i}...')
        boxes = tvops.clip_boxes_to_image(boxes, (img.shape[1], img.shape[2]))       # classification_eval.py:23
        crops = datautils.crops_for_classification(img.to(device), boxes)            # :24, all boxes in one launch
        predicted = classifier.classify(crops)                                       # :25
        total += len(target_anns)
        for want, got in zip(target_anns, predicted):
            per_annotation[want] += 1
            for knn in k:
                correct[knn] += want in got[:knn]
            if want != got[0]:
                missed[want] += 1
                mistaken_for.setdefault(want, Counter())[got[0]] += 1
    encoder.requires_grad_(True)
    accuracy = {knn: correct[knn] / total for knn in k}
    if verbose:
        print(f'Total annotations: {total}, Correctly guessed: {dict(correct)}, Accuracy: {accuracy}')
        worst = sorted(((n / per_annotation[a], n, a) for a, n in missed.items()), reverse=True)[:10]
        print('Most missed: ' + ', '.join(f'{a} ({n}, {p * 100} %)' for p, n, a in worst))
        for _, n, a in worst[:3]:
            print(f'{a}: commonly mistaken for ' + ', '.join(f'{b} ({m})' for b, m in mistaken_for[a].most_common(3)))
    return accuracy
