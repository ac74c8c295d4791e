"""Gate for the numeric mode bench.py's headline runs in.

bench.py's default line times HotPath with the FAST trunk (channels-last + torch.cudnn_convolution_relu + the library's
max-pool), 4-channel crops (CVPCE_NHWC4), bench.py's trunk batch (bench.TRUNK_BATCH), S = 256, the tcgen05 search with exact fp32 re-score and --
torch's GPU default -- TF32 convolutions.  Every other parity test switches TF32 off.  Here the step is built exactly as
bench.py:run_ours_shelves builds it and compared with the reference's torch CPU path in strict fp32
(oracle/torch_port.py: torchvision nms, python crop loop + antialiased resize, VGG16 fp32, amax / cat / norm):

* NMS boxes: bit-equal, in order (the library's kernels are fp32 in every mode);
* embeddings: the north-star's "within 1e-3 relative of the fp32 reference", measured as max_r ||e_r - e_r^cpu|| / ||e_r^cpu||
  (rows are unit vectors); the element-wise relative error over entries > 1e-3 is recorded next to it, not asserted at 1e-3:
  an entry of a unit 1024-vector is ~0.03 and carries the absolute error of the whole row;
* top-1 / top-k: a gallery row is a unit vector, so a row's similarity moves by at most ||e_r - e_r^cpu||: wherever the fp32
  gap between consecutive ranks exceeds TWICE that, the index must be identical; the overall agreement rate must meet the
  bar stated below.
The same comparison is repeated with fp32 convolutions (--strict-fp32, SURVEY a7 "parity mode keeps fp32").
The measured numbers are written to gpurun_out/headline_mode_check.json (copied to profiles/ by hand).
"""
import json
import os

import numpy as np
import pytest
import torch

import bench
from cvpce_b200 import synth

pytestmark = pytest.mark.gpu

TOP1_BAR_TF32 = 0.97     # agreement of top-1 indices with the fp32 CPU path over ALL rows (near-ties included)
TOP1_BAR_FP32 = 0.995
EMB_BAR_TF32 = 1e-3      # max_r ||de_r|| / ||e_r||
EMB_BAR_FP32 = 1e-4


def _cpu_reference(images, boxes, scores, seg, gallery, k):
    """The reference's torch CPU path, shelf by shelf: kept boxes, fp32 embeddings."""
    from oracle import torch_port
    from cvpce_b200 import classification
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    port = torch_port.MACVGGPort().eval().requires_grad_(False)
    port.block1.load_state_dict(enc.block1.state_dict())   # the same random-init weights as the GPU encoder below
    port.block2.load_state_dict(enc.block2.state_dict())
    kept, embs = [], []
    for i, img in enumerate(images):
        b = torch.from_numpy(boxes[seg[i]:seg[i + 1]])
        s = torch.from_numpy(scores[seg[i]:seg[i + 1]])
        kb, _, emb = torch_port.hot_path(port, gallery[:64], img, b, s, k=1, detections_per_img=200, return_embedding=True)
        kept.append(kb)
        embs.append(emb)
    return port, torch.cat(kept), torch.cat(embs)


def _compare(out, want_boxes, emb_cpu, gallery, k, emb_bar, top1_bar):
    assert torch.equal(out["boxes"].cpu(), want_boxes), "NMS keep list differs from torchvision's CPU op"
    e = out["embedding"].cpu().double()
    c = emb_cpu.double()
    de = (e - c).norm(dim=1)
    rel_norm = float((de / c.norm(dim=1)).max())
    big = c.abs() > 1e-3
    rel_elem = float(((e - c).abs()[big] / c.abs()[big]).max())
    g = gallery.double()
    sims = c @ g.T
    top = sims.sort(dim=1, descending=True)
    got = out["topk_idx"].cpu()
    gaps = top.values[:, :k] - top.values[:, 1:k + 1]          # gap below rank j
    gaps_above = torch.cat([torch.full((len(c), 1), float("inf"), dtype=torch.float64), gaps[:, :-1]], 1)
    clear = (gaps > 2 * de[:, None]) & (gaps_above > 2 * de[:, None])
    assert torch.equal(got[clear], top.indices[:, :k][clear]), "index differs where the fp32 ranking is decided by more than the embedding error"
    top1 = float((got[:, 0] == top.indices[:, 0]).double().mean())
    lists = float((got == top.indices[:, :k]).all(dim=1).double().mean())
    score_err = float((out["topk_sim"].cpu().double()[:, 0] - sims.gather(1, got[:, :1])[:, 0]).abs().max())
    stats = {"crops": int(len(c)), "emb_rel_norm_max": rel_norm, "emb_rel_elementwise_max_over_entries_gt_1e-3": rel_elem,
             "emb_abs_max": float((e - c).abs().max()), "top1_agreement": top1, "top%d_list_agreement" % k: lists,
             "rows_with_clear_top1": float(clear[:, 0].double().mean()), "top1_score_abs_err_max": score_err}
    assert rel_norm <= emb_bar, stats
    assert top1 >= top1_bar, stats
    assert score_err <= 2 * float(de.max()) + 1e-6, stats
    return stats


def test_headline_numeric_mode_against_fp32_cpu_path(cuda):
    from cvpce_b200 import classification, pipeline
    dev = torch.device("cuda", 0)
    H, W = synth.SHELF_HW
    n_shelves, products, k, G = 2, 150, 5, 8192
    imgs_d, boxes, scores, sizes = [], [], [], []
    for i in range(n_shelves):
        gen = torch.Generator(device=dev).manual_seed(100 + i)          # bench.py:build_shelves
        imgs_d.append(torch.rand((3, H, W), generator=gen, device=dev))
        b, s = synth.shelf_candidates(100 + i, H, W, products)
        boxes.append(b)
        scores.append(s)
        sizes.append(len(s))
    boxes, scores = np.concatenate(boxes), np.concatenate(scores)
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    gen = torch.Generator(device=dev).manual_seed(3)                    # bench.py:build_gallery
    gallery = torch.randn((G, 1024), generator=gen, device=dev).abs_()
    gallery = gallery / gallery.norm(dim=1, keepdim=True)
    _, want_boxes, emb_cpu = _cpu_reference([im.cpu() for im in imgs_d], boxes, scores, seg, gallery.cpu(), k)
    assert len(want_boxes) >= 256

    results = {}
    saved = torch.backends.cudnn.allow_tf32
    try:
        for mode, tf32, emb_bar, top1_bar in (("headline_tf32", True, EMB_BAR_TF32, TOP1_BAR_TF32), ("strict_fp32", False, EMB_BAR_FP32, TOP1_BAR_FP32)):
            torch.backends.cudnn.allow_tf32 = tf32
            torch.manual_seed(synth.WEIGHT_SEED)                         # bench.py:build_encoder
            enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(dev)
            enc.enable_fast_trunk()
            hp = pipeline.HotPath(enc, gallery, k=k, trunk_batch=bench.TRUNK_BATCH, trunk_dtype=torch.float32, channels_last=True, n_streams=1, pad_channel=True)
            out = hp.step(imgs_d, torch.from_numpy(boxes).to(dev), torch.from_numpy(scores).to(dev), torch.from_numpy(seg).to(dev), max(sizes))
            torch.cuda.synchronize()
            results[mode] = _compare(out, want_boxes, emb_cpu, gallery.cpu(), k, emb_bar, top1_bar)
            results[mode]["allow_tf32"] = tf32
    finally:
        torch.backends.cudnn.allow_tf32 = saved
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", "headline_mode_check.json"), "w") as f:
        json.dump({"config": "HotPath as bench.py builds it: fast trunk, NHWC4 crops, trunk batch %d, S=256, tcgen05 + exact re-score, "
                             "%d shelves x ~%d products, gallery %d x 1024, top-%d; reference = oracle/torch_port.py, strict fp32 on the CPU" % (bench.TRUNK_BATCH, n_shelves, products, G, k),
                   "bars": {"emb_rel_norm_tf32": EMB_BAR_TF32, "emb_rel_norm_fp32": EMB_BAR_FP32, "top1_tf32": TOP1_BAR_TF32, "top1_fp32": TOP1_BAR_FP32},
                   "results": results}, f, indent=1)
