"""Tensor-in / tensor-out wrappers over the C ABI (include/cvpce_b200.h).

torch is plumbing here: it owns device memory and the current stream; all arithmetic happens in libcvpce_b200.so.
CPU tensors are rejected -- there is no CPU fallback.
"""
import ctypes
import math
import struct

import torch

from . import _lib
from ._lib import BF16, F16, F32, NCHW, NHWC, NHWC4  # noqa: F401

_DTYPE_CODE = {torch.float32: F32, torch.bfloat16: BF16, torch.float16: F16}


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _need_cuda(name, *tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("%s: expected CUDA tensors (cvpce_b200 has no CPU fallback), got device %s" % (name, t.device))


def threshold_as_cpu_op(iou_threshold):
    """torchvision's CPU nms compares the fp32 IoU with a DOUBLE threshold; `ovr > thr_double` is equivalent to
    `ovr > f` with f the largest fp32 <= thr_double. (Equal to float(thr) for 0.5, the only value cvpce uses.)"""
    f = struct.unpack("f", struct.pack("f", iou_threshold))[0]
    if f > iou_threshold:  # rounded up: step one ulp down
        bits = struct.unpack("I", struct.pack("f", f))[0]
        bits = bits - 1 if f > 0 else bits + 1
        f = struct.unpack("f", struct.pack("I", bits))[0]
    return f


# ---------------------------------------------------------------------------------------------------------------------
# NMS  (tv:ops/boxes.py:20-48, tv:models/detection/retinanet.py:559-561, cvpce/production.py:13-15)
# ---------------------------------------------------------------------------------------------------------------------
def nms_batched(boxes, scores, seg_offsets, max_n, iou_threshold=0.5, min_score=None, max_keep=0, box_scale=None,
                want_boxes=False, want_scores=False):
    """Greedy NMS over a CSR batch of images in one launch.

    boxes f32[total,4], scores f32[total], seg_offsets i32[n_images+1] (device), max_n = host upper bound of candidates per image.
    Returns (keep i64[total] local indices laid out per segment, keep_counts i32[n_images], kept_boxes|None, kept_scores|None).
    No host synchronisation.  keep_counts[b] = -1 flags a segment longer than max_n: pass the host copy of the counts through
    checked_counts() before using them."""
    _need_cuda("nms_batched", boxes, scores, seg_offsets, box_scale)
    boxes = boxes.contiguous().float()
    scores = scores.contiguous().float()
    seg_offsets = seg_offsets.contiguous().to(torch.int32)
    n_images = seg_offsets.numel() - 1
    total = boxes.shape[0]
    keep = torch.empty(total, dtype=torch.int64, device=boxes.device)
    counts = torch.empty(n_images, dtype=torch.int32, device=boxes.device)
    kept_boxes = torch.empty((total, 4), dtype=torch.float32, device=boxes.device) if want_boxes else None
    kept_scores = torch.empty(total, dtype=torch.float32, device=boxes.device) if want_scores else None
    if box_scale is not None:
        box_scale = box_scale.contiguous().float()
    lib = _lib.load()
    if total == 0:  # no candidate in any image: torch hands out null pointers for empty tensors, nothing to launch
        counts.zero_()
        return keep, counts, kept_boxes, kept_scores
    rc = lib.cvpce_nms_batched(_ptr(boxes), _ptr(scores), _ptr(seg_offsets), n_images, int(max_n),
                               threshold_as_cpu_op(float(iou_threshold)),
                               float("-inf") if min_score is None else float(min_score), int(max_keep),
                               _ptr(keep), _ptr(counts), _ptr(box_scale), _ptr(kept_boxes), _ptr(kept_scores), _stream())
    _lib.check(rc, "cvpce_nms_batched")
    return keep, counts, kept_boxes, kept_scores


def nms_segments(boxes, scores, seg_stride, seg_counts, iou_threshold=0.5, min_score=None, max_keep=0, box_scale=None,
                 want_boxes=False, want_scores=False):
    """nms_batched over fixed-stride segments (image b = rows [b*seg_stride, b*seg_stride + seg_counts[b])): the layout
    select_candidates writes.  Same returns as nms_batched; no host synchronisation."""
    _need_cuda("nms_segments", boxes, scores, seg_counts, box_scale)
    if boxes.dtype != torch.float32 or scores.dtype != torch.float32 or not boxes.is_contiguous() or not scores.is_contiguous():
        raise RuntimeError("nms_segments: boxes / scores must be contiguous float32")
    seg_counts = seg_counts.contiguous().to(torch.int32)
    n_images = seg_counts.numel()
    total = boxes.shape[0]
    if total < n_images * seg_stride:
        raise RuntimeError("nms_segments: boxes has %d rows, needs n_images*seg_stride = %d" % (total, n_images * seg_stride))
    keep = torch.empty(total, dtype=torch.int64, device=boxes.device)
    counts = torch.empty(n_images, dtype=torch.int32, device=boxes.device)
    kept_boxes = torch.empty((total, 4), dtype=torch.float32, device=boxes.device) if want_boxes else None
    kept_scores = torch.empty(total, dtype=torch.float32, device=boxes.device) if want_scores else None
    if box_scale is not None:
        box_scale = box_scale.contiguous().float()
    rc = _lib.load().cvpce_nms_segments(_ptr(boxes), _ptr(scores), int(seg_stride), _ptr(seg_counts), n_images,
                                        threshold_as_cpu_op(float(iou_threshold)),
                                        float("-inf") if min_score is None else float(min_score), int(max_keep),
                                        _ptr(keep), _ptr(counts), _ptr(box_scale), _ptr(kept_boxes), _ptr(kept_scores), _stream())
    _lib.check(rc, "cvpce_nms_segments")
    return keep, counts, kept_boxes, kept_scores


# ---------------------------------------------------------------------------------------------------------------------
# box IoU + greedy IoU claim  (tv:ops/boxes.py box_iou; cvpce/metrics.py:11-34; cvpce/planograms.py:214-232)
# ---------------------------------------------------------------------------------------------------------------------
MATCH_MAX_COLS = 16384


def _boxes_f32(name, b):
    if b.dim() != 2 or b.shape[1] != 4:
        raise RuntimeError("%s: boxes must be [N,4], got %s" % (name, tuple(b.shape)))
    return b.contiguous().float()


def box_iou(boxes_a, boxes_b):
    """torchvision.ops.box_iou(boxes_a [Na,4], boxes_b [Nb,4]) -> f32 [Na,Nb], bit-equal to torchvision's fp32 result."""
    _need_cuda("box_iou", boxes_a, boxes_b)
    a, b = _boxes_f32("box_iou", boxes_a), _boxes_f32("box_iou", boxes_b)
    out = torch.empty((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
    rc = _lib.load().cvpce_box_iou(_ptr(a), a.shape[0], _ptr(b), b.shape[0], _ptr(out), _stream())
    _lib.check(rc, "cvpce_box_iou")
    return out


def iou_match(row_boxes, col_boxes, iou_threshold=0.5, row_labels=None, col_labels=None, row_offsets=None, col_offsets=None,
              max_cols=None):
    """The reference's greedy claim pass (metrics.check_matches / the matching loop of planograms.finalize_via_ransac) without
    the IoU matrix or its sort: rows in index order mark every still-unmarked column with !(iou < thr) (and equal label).

    One problem by default; pass i32 CSR `row_offsets` / `col_offsets` (device) plus the host bound `max_cols` for a batch of
    images in one launch.  Returns (row_matched bool [R], col_matched bool [C]).  No host synchronisation."""
    _need_cuda("iou_match", row_boxes, col_boxes, row_labels, col_labels, row_offsets, col_offsets)
    r, c = _boxes_f32("iou_match", row_boxes), _boxes_f32("iou_match", col_boxes)
    dev = r.device
    if (row_labels is None) != (col_labels is None):
        raise RuntimeError("iou_match: give both label tensors or neither")
    if row_labels is not None:
        row_labels, col_labels = row_labels.contiguous().to(torch.int64), col_labels.contiguous().to(torch.int64)
        if row_labels.numel() != r.shape[0] or col_labels.numel() != c.shape[0]:
            raise RuntimeError("iou_match: one label per box expected")
    if (row_offsets is None) != (col_offsets is None):
        raise RuntimeError("iou_match: give both offset tensors or neither")
    if row_offsets is None:
        row_offsets = torch.tensor([0, r.shape[0]], dtype=torch.int32, device=dev)
        col_offsets = torch.tensor([0, c.shape[0]], dtype=torch.int32, device=dev)
        max_cols = c.shape[0]
    elif max_cols is None:
        raise RuntimeError("iou_match: max_cols (host upper bound of columns per problem) is required with offsets")
    row_offsets, col_offsets = row_offsets.contiguous().to(torch.int32), col_offsets.contiguous().to(torch.int32)
    n_problems = row_offsets.numel() - 1
    if col_offsets.numel() - 1 != n_problems:
        raise RuntimeError("iou_match: row_offsets and col_offsets describe different numbers of problems")
    row_m = torch.zeros(r.shape[0], dtype=torch.uint8, device=dev)
    col_m = torch.zeros(c.shape[0], dtype=torch.uint8, device=dev)
    if r.shape[0] and c.shape[0]:
        rc = _lib.load().cvpce_iou_match(_ptr(r), _ptr(row_labels), _ptr(row_offsets), _ptr(c), _ptr(col_labels), _ptr(col_offsets),
                                         n_problems, int(max_cols), float(iou_threshold), _ptr(row_m), _ptr(col_m), _stream())
        _lib.check(rc, "cvpce_iou_match")
    return row_m.bool(), col_m.bool()


# ---------------------------------------------------------------------------------------------------------------------
# pre-NMS candidate selection  (tv:models/detection/retinanet.py:527-557, tv:models/detection/_utils.py:183-224)
# ---------------------------------------------------------------------------------------------------------------------
BBOX_XFORM_CLIP = math.log(1000.0 / 16)


def select_candidates(cls_logits, bbox_regression, anchors, image_sizes, score_thresh=0.05, topk=1000, weights=None,
                      bbox_xform_clip=BBOX_XFORM_CLIP, want_labels=True):
    """Sigmoid, threshold, per-level top-k, box decode, clip and concatenation over FPN levels for a whole batch, one launch.

    cls_logits: list over levels of f32 [B,A_l,C] (views of the concatenated head output are fine: only the last two dims
    must be contiguous); bbox_regression: list of f32 [B,A_l,4]; anchors: list of f32 [A_l,4] (one set shared by the
    batch) or of [B,A_l,4]; image_sizes: i32 [B,2] device tensor or a list of (h, w).
    Returns (boxes f32 [B*L*topk,4], scores f32 [B*L*topk], labels i64|None, seg_counts i32 [B], level_counts i32 [B,L],
    seg_stride): image b's candidates are rows [b*seg_stride, b*seg_stride + seg_counts[b])."""
    n_levels = len(cls_logits)
    if n_levels == 0 or n_levels != len(bbox_regression) or n_levels != len(anchors):
        raise RuntimeError("select_candidates: need the same, non-zero number of levels in logits / regression / anchors")
    _need_cuda("select_candidates", *cls_logits, *bbox_regression, *anchors)
    device = cls_logits[0].device
    B, _, C = cls_logits[0].shape
    if not isinstance(image_sizes, torch.Tensor):
        image_sizes = torch.tensor([[int(h), int(w)] for h, w in image_sizes], dtype=torch.int32).to(device, non_blocking=True)
    image_sizes = image_sizes.contiguous().to(torch.int32)
    _need_cuda("select_candidates", image_sizes)
    levels = (_lib.FpnLevel * n_levels)()
    hold = []
    for l, (lg, rg, an) in enumerate(zip(cls_logits, bbox_regression, anchors)):
        if lg.dtype != torch.float32 or rg.dtype != torch.float32:
            raise RuntimeError("select_candidates: head outputs must be float32")
        A = lg.shape[1]
        if lg.shape != (B, A, C) or rg.shape != (B, A, 4) or an.shape[-2:] != (A, 4):
            raise RuntimeError("select_candidates: level %d shapes %s / %s / %s do not agree" % (l, tuple(lg.shape), tuple(rg.shape), tuple(an.shape)))
        if A > 0 and (lg.stride(2) != 1 or (lg.stride(1) != C and A > 1)):
            lg = lg.contiguous()
        if A > 0 and (rg.stride(2) != 1 or (rg.stride(1) != 4 and A > 1) or rg.data_ptr() % 16 or (rg.stride(0) % 4 and B > 1)):
            rg = rg.contiguous()
        an = an.float()
        if not an.is_contiguous() or an.data_ptr() % 16:
            an = an.contiguous().clone()
        hold += [lg, rg, an]
        levels[l] = _lib.FpnLevel(lg.data_ptr(), rg.data_ptr(), an.data_ptr(), lg.stride(0) if B > 1 else 0,
                                  rg.stride(0) if B > 1 else 0, an.stride(0) if an.dim() == 3 and B > 1 else 0, A)
    seg_stride = n_levels * int(topk)
    boxes = torch.empty((B * seg_stride, 4), dtype=torch.float32, device=device)
    scores = torch.empty(B * seg_stride, dtype=torch.float32, device=device)
    labels = torch.empty(B * seg_stride, dtype=torch.int64, device=device) if want_labels else None
    seg_counts = torch.empty(B, dtype=torch.int32, device=device)
    level_counts = torch.empty((B, n_levels), dtype=torch.int32, device=device)
    w = None if weights is None else (ctypes.c_float * 4)(*[float(v) for v in weights])
    rc = _lib.load().cvpce_select_candidates(levels, n_levels, B, C, float(score_thresh), int(topk), _ptr(image_sizes),
                                             ctypes.cast(w, ctypes.c_void_p) if w is not None else None,
                                             float(bbox_xform_clip), _ptr(boxes), _ptr(scores), _ptr(labels),
                                             _ptr(seg_counts), _ptr(level_counts), _stream())
    _lib.check(rc, "cvpce_select_candidates")
    del hold
    return boxes, scores, labels, seg_counts, level_counts, seg_stride


def checked_counts(counts_host, what="nms"):
    """The kernels report a segment with more candidates than the caller's max_n bound (or than CVPCE_NMS_MAX_N) by writing
    keep_counts[b] = -1.  Call this on the HOST copy of the counts (the step's one D2H read): a wrong bound raises instead of
    silently dropping or garbling detections."""
    bad = (counts_host < 0).nonzero()
    if bad.numel():
        raise RuntimeError("%s: image(s) %s hold more candidates than max_n / CVPCE_NMS_MAX_N=%d allows (keep_counts = -1)"
                           % (what, bad.flatten().tolist()[:8], _lib.NMS_MAX_N))
    return counts_host


def nms(boxes, scores, iou_threshold):
    """Drop-in for torchvision.ops.nms(boxes, scores, iou_threshold) -> int64[K] (one D2H read of the count)."""
    _need_cuda("nms", boxes, scores)
    n = boxes.shape[0]
    if n == 0:
        return torch.empty(0, dtype=torch.int64, device=boxes.device)
    seg = torch.tensor([0, n], dtype=torch.int32, device=boxes.device)
    keep, counts, _, _ = nms_batched(boxes, scores, seg, n, iou_threshold)
    return keep[: int(checked_counts(counts.cpu(), "nms")[0])]


def batched_nms(boxes, scores, idxs, iou_threshold):
    """Drop-in for torchvision.ops.batched_nms (tv:ops/boxes.py:51-103): per-class NMS through the coordinate-offset
    trick torchvision itself uses on CUDA (single-class GLN input: the offset is 0)."""
    _need_cuda("batched_nms", boxes, scores, idxs)
    if boxes.numel() == 0:
        return torch.empty((0,), dtype=torch.int64, device=boxes.device)
    max_coordinate = boxes.max()
    offsets = idxs.to(boxes) * (max_coordinate + torch.tensor(1).to(boxes))
    return nms(boxes + offsets[:, None], scores, iou_threshold)


# ---------------------------------------------------------------------------------------------------------------------
# 2x2 max pooling between the trunk's convolutions  (tv:models/vgg.py make_layers 'M'; classification.py:31-37)
# ---------------------------------------------------------------------------------------------------------------------
def maxpool2x2_supported(x):
    per16 = 4 if x.dtype == torch.float32 else 8
    return (x.is_cuda and x.dim() == 4 and x.dtype in _DTYPE_CODE and x.shape[1] % per16 == 0
            and x.shape[2] >= 2 and x.shape[3] >= 2 and x.is_contiguous(memory_format=torch.channels_last))


def maxpool2x2(x):
    """nn.MaxPool2d(kernel_size=2, stride=2) on a channels_last CUDA tensor [B,C,H,W] -> channels_last [B,C,H//2,W//2]."""
    _need_cuda("maxpool2x2", x)
    if not maxpool2x2_supported(x):
        raise RuntimeError("maxpool2x2: needs a channels_last float32 / bfloat16 / float16 CUDA tensor with 16-byte channel groups")
    B, C, H, W = x.shape
    out = torch.empty((B, C, H // 2, W // 2), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    rc = _lib.load().cvpce_maxpool2x2_nhwc(_ptr(x), _DTYPE_CODE[x.dtype], B, H, W, C, _ptr(out), _stream())
    _lib.check(rc, "cvpce_maxpool2x2_nhwc")
    return out


# ---------------------------------------------------------------------------------------------------------------------
# crop + pad + resize  (cvpce/production.py:20, cvpce/datautils.py:234-239, utils.py:280-281, classification.py:41-44)
# ---------------------------------------------------------------------------------------------------------------------
def make_image_table(images):
    """images: list of CUDA f32 [3,H,W] tensors (pixel stride 1) or one [n,3,H,W] tensor -> i64[n,5] device table.
    The caller must keep `images` alive while the table is used."""
    if isinstance(images, torch.Tensor):
        images = [images] if images.dim() == 3 else list(images.unbind(0))
    rows = []
    for im in images:
        _need_cuda("make_image_table", im)
        if im.dtype != torch.float32 or im.dim() != 3 or im.shape[0] != 3 or im.stride(2) != 1:
            raise RuntimeError("make_image_table: images must be float32 [3,H,W] with unit pixel stride")
        rows.append([im.data_ptr(), im.shape[1], im.shape[2], im.stride(0), im.stride(1)])
    return torch.tensor(rows, dtype=torch.int64).to(images[0].device, non_blocking=True)


def crop_resize(image_table, n_images, boxes, box_image_idx=None, size=256, antialias=True, pad_value=0.5, affine=None,
                out_dtype=torch.float32, channels_last=False, return_valid=False, out=None, pad_channel=False):
    """All crops of a batch in one launch. boxes f32[K,4] (device) -> [K,3,S,S] (or a channels_last view of [K,S,S,3]).
    pad_channel (with channels_last): [K,S,S,4] with a zero 4th channel, returned as a channels_last [K,4,S,S] view -- what
    classification.MACVGG's fast trunk feeds to the first convolution with a zero-padded weight.
    `out`: optional preallocated PHYSICAL buffer ([K,S,S,3] when channels_last, else [K,3,S,S]; contiguous), e.g. a row slice
    of a larger crop tensor that is filled group by group."""
    _need_cuda("crop_resize", image_table, boxes, box_image_idx, out)
    boxes = boxes.contiguous().float()
    K = boxes.shape[0]
    S = int(size)
    if out_dtype not in _DTYPE_CODE:
        raise RuntimeError("crop_resize: out_dtype must be float32, bfloat16 or float16")
    dt = _DTYPE_CODE[out_dtype]
    if pad_channel and not channels_last:
        raise RuntimeError("crop_resize: pad_channel needs channels_last")
    shape = (K, S, S, 4 if pad_channel else 3) if channels_last else (K, 3, S, S)
    if out is None:
        out = torch.empty(shape, dtype=out_dtype, device=boxes.device)
    elif tuple(out.shape) != shape or out.dtype != out_dtype or not out.is_contiguous():
        raise RuntimeError("crop_resize: out must be a contiguous %s tensor of shape %s" % (out_dtype, shape))
    valid = torch.empty(K, dtype=torch.int32, device=boxes.device) if return_valid else None
    if box_image_idx is not None:
        box_image_idx = box_image_idx.contiguous().to(torch.int32)
    aff = None
    if affine is not None:
        aff = (ctypes.c_float * 6)(*[float(v) for v in affine])
    if K > 0:
        rc = _lib.load().cvpce_crop_resize(_ptr(image_table), int(n_images), _ptr(boxes), _ptr(box_image_idx), K, S,
                                           1 if antialias else 0, float(pad_value),
                                           ctypes.cast(aff, ctypes.c_void_p) if aff is not None else None, dt,
                                           (NHWC4 if pad_channel else NHWC) if channels_last else NCHW, _ptr(out), _ptr(valid), _stream())
        _lib.check(rc, "cvpce_crop_resize")
    if channels_last:
        out = out.permute(0, 3, 1, 2)  # logical NCHW, physical NHWC (torch.channels_last)
    return (out, valid) if return_valid else out


# ---------------------------------------------------------------------------------------------------------------------
# region max-pool embedding  (cvpce/models/classification.py:46-51, :79-85)
# ---------------------------------------------------------------------------------------------------------------------
def rmac_regions(H, W, levels):
    buf = (ctypes.c_int32 * (4 * 64))()
    n = _lib.load().cvpce_rmac_regions(int(H), int(W), int(levels), ctypes.cast(buf, ctypes.c_void_p), 64)
    return [tuple(buf[4 * i: 4 * i + 4]) for i in range(min(n, 64))]


_POOL_WS = {}  # (device index, stream handle) -> zero-initialised workspace of the balanced MAC kernel (its counters stay zero between launches)


def _pool_workspace(device, stream, need):
    key = (device.index, stream.value or 0)
    ws = _POOL_WS.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros((max(need, 1 << 20),), dtype=torch.uint8, device=device)
        _POOL_WS[key] = ws
    return ws


def region_pool_embed(feats, regions=None, per_region_l2=False, final_l2=True, eps=1e-8, out=None):
    """feats: list of CUDA [B,C,H,W] maps (f32 or bf16, any strides) -> f32[B, sum C].
    regions: None (MAC: whole map) or a list (per map) of lists of (y0,x0,y1,x1).
    Dense channels-last MAC batches take the balanced streaming kernel through a cached per-(device, stream) workspace."""
    _need_cuda("region_pool_embed", *feats)
    B = feats[0].shape[0]
    dtype = feats[0].dtype
    if dtype not in _DTYPE_CODE or any(f.dtype != dtype or f.shape[0] != B or f.dim() != 4 for f in feats):
        raise RuntimeError("region_pool_embed: maps must be 4-d, share batch size and dtype (float32, bfloat16; float16 for dense channels-last MAC)")
    D = sum(f.shape[1] for f in feats)
    if out is None:
        out = torch.empty((B, D), dtype=torch.float32, device=feats[0].device)
    maps = (_lib.FeatureMap * len(feats))()
    for i, f in enumerate(feats):
        maps[i] = _lib.FeatureMap(f.data_ptr(), f.shape[1], f.shape[2], f.shape[3], f.stride(0), f.stride(1), f.stride(2), f.stride(3))
    rg = nr = None
    if regions is not None:
        flat = [int(v) for per_map in regions for r in per_map for v in r]
        rg = (ctypes.c_int32 * len(flat))(*flat)
        nr = (ctypes.c_int32 * len(regions))(*[len(r) for r in regions])
    lib = _lib.load()
    stream = _stream()
    ws, ws_bytes = None, 0
    if regions is None and not per_region_l2 and B > 0:
        ws_bytes = int(lib.cvpce_region_pool_workspace_bytes(maps, len(feats), _DTYPE_CODE[dtype], B))
        if ws_bytes > 0:
            ws = _pool_workspace(feats[0].device, stream, ws_bytes)
    rc = lib.cvpce_region_pool_embed_ws(maps, len(feats), _DTYPE_CODE[dtype], B,
                                        ctypes.cast(rg, ctypes.c_void_p) if rg is not None else None,
                                        ctypes.cast(nr, ctypes.c_void_p) if nr is not None else None,
                                        1 if per_region_l2 else 0, 1 if final_l2 else 0, float(eps), _ptr(out),
                                        out.stride(0), _ptr(ws) if ws is not None else None, ws_bytes, stream)
    _lib.check(rc, "cvpce_region_pool_embed")
    return out


def l2norm_rows_(x, eps=1e-8):
    _need_cuda("l2norm_rows_", x)
    if x.dtype != torch.float32 or x.dim() != 2 or x.stride(1) != 1:
        raise RuntimeError("l2norm_rows_: expected float32 [B,D] with unit column stride")
    rc = _lib.load().cvpce_l2norm_rows(_ptr(x), x.shape[0], x.shape[1], x.stride(0), float(eps), _stream())
    _lib.check(rc, "cvpce_l2norm_rows")
    return x


# ---------------------------------------------------------------------------------------------------------------------
# cosine similarity + top-k  (cvpce/models/classification.py:87-95)
# ---------------------------------------------------------------------------------------------------------------------
def rows_prepare(x, normalize=True, eps=1e-8, want_bf16=True):
    """f32[R,D] -> (bf16[R,D] of x/max(||x||,eps) (or a plain cast), inv_norm f32[R])."""
    _need_cuda("rows_prepare", x)
    x = x.contiguous().float()
    R, D = x.shape
    xb = torch.empty((R, D), dtype=torch.bfloat16, device=x.device) if want_bf16 else None
    inv = torch.empty(R, dtype=torch.float32, device=x.device)
    rc = _lib.load().cvpce_rows_prepare(_ptr(x), R, D, 1 if normalize else 0, float(eps), _ptr(xb), _ptr(inv), _stream())
    _lib.check(rc, "cvpce_rows_prepare")
    return xb, inv


def rows_prepare_scatter(x, peer_f32_ptrs, peer_bf16_ptrs, peer_inv_ptrs, normalize=True, eps=1e-8):
    """rows_prepare of this rank's block x f32[R,D] whose three results are stored by the kernel into every rank's query
    buffers: the pointer lists are raw device addresses (ints) of row 0 of THIS block inside each rank's [Q,D] fp32, [Q,D]
    bf16 and [Q] inverse-norm buffers (peer memory over NVLink).  The caller synchronises the ranks before searching."""
    _need_cuda("rows_prepare_scatter", x)
    x = x.contiguous().float()
    R, D = x.shape
    n = len(peer_f32_ptrs)
    if n < 1 or len(peer_bf16_ptrs) != n or len(peer_inv_ptrs) != n:
        raise RuntimeError("rows_prepare_scatter: need one fp32, one bf16 and one inverse-norm pointer per peer")
    arrs = [(ctypes.c_void_p * n)(*[int(p) for p in ptrs]) for ptrs in (peer_f32_ptrs, peer_bf16_ptrs, peer_inv_ptrs)]
    rc = _lib.load().cvpce_rows_prepare_scatter(_ptr(x), R, D, 1 if normalize else 0, float(eps),
                                                *[ctypes.cast(a, ctypes.c_void_p) for a in arrs], n, _stream())
    _lib.check(rc, "cvpce_rows_prepare_scatter")


def sim_topk_f32(queries, gallery, k=1, idx_base=0, normalize=True, eps=1e-8):
    """Exact-fp32 CUDA-core path: (similarity f32[Q,k], index i64[Q,k])."""
    _need_cuda("sim_topk_f32", queries, gallery)
    q = queries.contiguous().float()
    g = gallery.contiguous().float()
    Q, D = q.shape
    G = g.shape[0]
    s = torch.empty((Q, k), dtype=torch.float32, device=q.device)
    i = torch.empty((Q, k), dtype=torch.int64, device=q.device)
    rc = _lib.load().cvpce_sim_topk_f32(_ptr(q), _ptr(g), Q, G, D, int(k), int(idx_base), 1 if normalize else 0, float(eps),
                                        _ptr(s), _ptr(i), _stream())
    _lib.check(rc, "cvpce_sim_topk_f32")
    return s, i


def sim_topk_workspace_bytes(Q, G, D, k):
    return int(_lib.load().cvpce_sim_topk_workspace_bytes(int(Q), int(G), int(D), int(k)))


def sim_topk(q_bf16, g_bf16, k=1, idx_base=0, q_f32=None, g_f32=None, q_inv_norm=None, g_inv_norm=None, workspace=None):
    """tcgen05 path: bf16 operands [Q,D], [G,D]; with q_f32/g_f32 the candidates are re-scored in exact fp32 (parity mode).
    Returns (similarity f32[Q,k], index i64[Q,k])."""
    _need_cuda("sim_topk", q_bf16, g_bf16, q_f32, g_f32, q_inv_norm, g_inv_norm, workspace)
    if q_bf16.dtype != torch.bfloat16 or g_bf16.dtype != torch.bfloat16 or not q_bf16.is_contiguous() or not g_bf16.is_contiguous():
        raise RuntimeError("sim_topk: operands must be contiguous bfloat16 [rows, D]")
    Q, D = q_bf16.shape
    G = g_bf16.shape[0]
    need = sim_topk_workspace_bytes(Q, G, D, k)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty(max(need, 16), dtype=torch.uint8, device=q_bf16.device)
    s = torch.empty((Q, k), dtype=torch.float32, device=q_bf16.device)
    i = torch.empty((Q, k), dtype=torch.int64, device=q_bf16.device)
    for t in (q_f32, g_f32):
        if t is not None and (t.dtype != torch.float32 or not t.is_contiguous()):
            raise RuntimeError("sim_topk: q_f32 / g_f32 must be contiguous float32")
    rc = _lib.load().cvpce_sim_topk(_ptr(q_bf16), _ptr(g_bf16), Q, G, D, int(k), int(idx_base), _ptr(q_f32), _ptr(g_f32),
                                    _ptr(q_inv_norm), _ptr(g_inv_norm), _ptr(s), _ptr(i), _ptr(workspace),
                                    workspace.numel() * workspace.element_size(), _stream())
    _lib.check(rc, "cvpce_sim_topk")
    return s, i


def sim_topk_scatter(q_bf16, g_bf16, k, idx_base, peer_score_ptrs, peer_idx_ptrs, q_f32=None, g_f32=None, q_inv_norm=None,
                     g_inv_norm=None, workspace=None):
    """sim_topk whose [Q,k] result block is stored by the kernel into every peer's gather buffer: peer_*_ptrs are raw device
    addresses (ints) of THIS rank's [Q,k] slot inside each rank's [W,Q,k] score / index buffers (peer memory over NVLink).
    The caller synchronises the ranks (symmetric-memory barrier) before merging."""
    _need_cuda("sim_topk_scatter", q_bf16, g_bf16, q_f32, g_f32, q_inv_norm, g_inv_norm, workspace)
    if q_bf16.dtype != torch.bfloat16 or g_bf16.dtype != torch.bfloat16 or not q_bf16.is_contiguous() or not g_bf16.is_contiguous():
        raise RuntimeError("sim_topk_scatter: operands must be contiguous bfloat16 [rows, D]")
    Q, D = q_bf16.shape
    G = g_bf16.shape[0]
    n = len(peer_score_ptrs)
    if n != len(peer_idx_ptrs) or n < 1:
        raise RuntimeError("sim_topk_scatter: need one score and one index pointer per peer")
    need = sim_topk_workspace_bytes(Q, G, D, k)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty(max(need, 16), dtype=torch.uint8, device=q_bf16.device)
    ps = (ctypes.c_void_p * n)(*[int(p) for p in peer_score_ptrs])
    pi = (ctypes.c_void_p * n)(*[int(p) for p in peer_idx_ptrs])
    rc = _lib.load().cvpce_sim_topk_scatter(_ptr(q_bf16), _ptr(g_bf16), Q, G, D, int(k), int(idx_base), _ptr(q_f32), _ptr(g_f32),
                                            _ptr(q_inv_norm), _ptr(g_inv_norm), ctypes.cast(ps, ctypes.c_void_p),
                                            ctypes.cast(pi, ctypes.c_void_p), n, _ptr(workspace),
                                            workspace.numel() * workspace.element_size(), _stream())
    _lib.check(rc, "cvpce_sim_topk_scatter")


def topk_merge(scores, idx):
    """scores f32[W,Q,k], idx i64[W,Q,k] (global indices) -> merged (f32[Q,k], i64[Q,k])."""
    _need_cuda("topk_merge", scores, idx)
    scores = scores.contiguous().float()
    idx = idx.contiguous().to(torch.int64)
    W, Q, k = scores.shape
    s = torch.empty((Q, k), dtype=torch.float32, device=scores.device)
    i = torch.empty((Q, k), dtype=torch.int64, device=scores.device)
    rc = _lib.load().cvpce_topk_merge(_ptr(scores), _ptr(idx), W, Q, k, _ptr(s), _ptr(i), _stream())
    _lib.check(rc, "cvpce_topk_merge")
    return s, i


def gpu_kernel_launch_count():
    """Not tracked by the library; bench.py counts launches from the calls it makes."""
    return math.nan
