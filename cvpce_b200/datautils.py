"""Host-side mirror of the one cvpce/datautils.py helper on the hot path."""
import torch

from . import ops

CLASSIFICATION_IMAGE_SIZE = 256  # datautils.py:232

# which resize the installed torchvision performs for ttf.resize on tensors (antialias default changed in 0.17);
# the reference pins torchvision 0.9 (plain bilinear). The oracle as run on the box decides: installed => True.
ANTIALIAS = True


def resize_for_classification(img):
    """datautils.py:234-239 for one already-sliced crop: img f32[3,h,w] CUDA view -> f32[3,256,256] on the same device."""
    _, h, w = img.shape
    if img.stride(2) != 1:
        img = img.contiguous()
    table = ops.make_image_table([img])
    box = torch.tensor([[0.0, 0.0, float(w), float(h)]], device=img.device)
    return ops.crop_resize(table, 1, box, None, CLASSIFICATION_IMAGE_SIZE, ANTIALIAS)[0]


def crops_for_classification(image, boxes, size=None, antialias=None):
    """The whole crop loop of production.py:20 / classification_eval.py:24 / detection_eval.py:29 in one launch:
    image f32[3,H,W] CUDA, boxes f32[K,4] (any device) -> f32[K,3,S,S] CUDA. No per-box host synchronisation."""
    size = CLASSIFICATION_IMAGE_SIZE if size is None else size
    antialias = ANTIALIAS if antialias is None else antialias
    table = ops.make_image_table([image])
    return ops.crop_resize(table, 1, boxes.to(image.device, torch.float32), None, size, antialias)
