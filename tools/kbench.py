#!/usr/bin/env python
"""Per-kernel micro-benchmarks at the BASELINE shapes (CUDA events on the launching stream, L2 flushed between
iterations). Development tool: prints one JSON line per kernel; bench.py stays the contract benchmark.

    python tools/kbench.py [crop] [nms] [select] [pool] [prep] [sim] [--shelves 32] [--iters 5]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torchvision

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cvpce_b200 import ops, synth  # noqa: E402

HBM = 6540.2
TENSOR = 1680.3
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        _p = json.load(f)
    HBM, TENSOR = float(_p["hbm_gbs"]), float(_p["bf16_tflops"])
except Exception:
    pass

_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    _flush.zero_()


def timeit(fn, iters=5, warmup=3):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(min(ts))


def report(name, ms, best, nbytes=None, flops=None, **extra):
    d = {"kernel": name, "ms_median": round(ms, 4), "ms_best": round(best, 4)}
    if nbytes is not None:
        d["GBs"] = round(nbytes / ms / 1e6, 1)
        d["frac_hbm"] = round(nbytes / ms / 1e6 / HBM, 4)
    if flops is not None:
        d["TFLOPs"] = round(flops / ms / 1e9, 1)
        d["frac_tensor"] = round(flops / ms / 1e9 / TENSOR, 4)
    d.update(extra)
    print(json.dumps(d), flush=True)


def shelves(n, products=200):
    H, W = synth.SHELF_HW
    boxes, scores, sizes = [], [], []
    for i in range(n):
        b, s = synth.shelf_candidates(100 + i, H, W, products)
        boxes.append(b)
        scores.append(s)
        sizes.append(len(s))
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    return np.concatenate(boxes), np.concatenate(scores), seg, max(sizes)


def bench_nms(args):
    dev = "cuda"
    b, s, seg, mx = shelves(args.shelves)
    bd, sd, segd = torch.from_numpy(b).to(dev), torch.from_numpy(s).to(dev), torch.from_numpy(seg).to(dev)
    ms, best = timeit(lambda: ops.nms_batched(bd, sd, segd, mx, 0.5, None, 200, want_boxes=True), args.iters)
    _, counts, _, _ = ops.nms_batched(bd, sd, segd, mx, 0.5, None, 200)
    K = int(counts.sum())
    report("nms c3 (%d images x ~%d)" % (args.shelves, mx), ms, best, 20.0 * len(s) + 8.0 * K, kept=K)
    # stress shape: 5000 random candidates x 512 images (the real pre-NMS candidate count)
    n_img, n = args.stress_images, 5000
    bb, ss = [], []
    for i in range(8):
        x, y = synth.random_boxes(50 + i, n)
        bb.append(x)
        ss.append(y)
    bb = np.concatenate([bb[i % 8] for i in range(n_img)])
    ss = np.concatenate([ss[i % 8] for i in range(n_img)])
    seg = (np.arange(n_img + 1) * n).astype(np.int32)
    bd, sd, segd = torch.from_numpy(bb).to(dev), torch.from_numpy(ss).to(dev), torch.from_numpy(seg).to(dev)
    ms, best = timeit(lambda: ops.nms_batched(bd, sd, segd, n, 0.5, None, 1000), args.iters)
    _, counts, _, _ = ops.nms_batched(bd, sd, segd, n, 0.5, None, 1000)
    K = int(counts.sum())
    report("nms stress (%d images x %d)" % (n_img, n), ms, best, 20.0 * len(ss) + 8.0 * K, kept=K)


def bench_select(args):
    """f1 + a1: candidate selection and the whole detector post-processing at the GLN frame (800 x 1066, 5 FPN levels)."""
    from cvpce_b200 import detection
    dev = "cuda"
    H, W = 800, 1066
    anchors = [torch.from_numpy(a).to(dev) for a in synth.fpn_anchors(H, W)]
    n_per = [a.shape[0] for a in anchors]
    A = sum(n_per)
    for B in (1, args.shelves):
        g = torch.Generator(device=dev).manual_seed(5)
        for mu, tag in ((-4.6, "trained-like"), (-1.0, "dense")):
            lg = (mu + 1.5 * torch.randn((B, A, 1), device=dev, generator=g))
            rg = 0.3 * torch.randn((B, A, 4), device=dev, generator=g)
            lgs, rgs = list(lg.split(n_per, dim=1)), list(rg.split(n_per, dim=1))
            sizes = torch.tensor([[H, W]] * B, dtype=torch.int32, device=dev)
            _, _, _, cnt, _, _ = ops.select_candidates(lgs, rgs, anchors, sizes, 0.05, 1000)
            ncand = int(cnt.sum())
            nbytes = 4.0 * A * B + 56.0 * ncand
            ms, best = timeit(lambda: ops.select_candidates(lgs, rgs, anchors, sizes, 0.05, 1000, want_labels=False), args.iters)
            report("select_candidates B=%d %s (%d anchors/img, %d cand)" % (B, tag, A, ncand), ms, best, nbytes)
            head = {"cls_logits": lgs, "bbox_regression": rgs}
            ms, best = timeit(lambda: detection.postprocess_detections_batch(head, anchors, sizes, 0.05, 1000, 0.5, 1000), args.iters)
            report("postprocess (select+nms) B=%d %s" % (B, tag), ms, best, nbytes + 20.0 * ncand)


def bench_maxpool(args):
    """The four 2x2 max-pools of the VGG16 trunk for one 128-crop trunk batch, ours vs at::max_pool2d (channels-last)."""
    dev = "cuda"
    B = 128
    for C, S in ((64, 256), (128, 128), (256, 64), (512, 32)):
        x = torch.randn((B, C, S, S), device=dev).contiguous(memory_format=torch.channels_last)
        nb = 1.25 * x.numel() * 4
        ms, best = timeit(lambda: ops.maxpool2x2(x), args.iters)
        report("maxpool2x2 nhwc f32 B=%d C=%d %dx%d" % (B, C, S, S), ms, best, nb)
        ms, best = timeit(lambda: torch.nn.functional.max_pool2d(x, 2, 2), args.iters)
        report("  torch max_pool2d (at::max_pool_forward_nhwc) same shape", ms, best, nb)


def bench_crop(args):
    dev = "cuda"
    H, W = synth.SHELF_HW
    b, s, seg, mx = shelves(args.shelves)
    bd, sd, segd = torch.from_numpy(b).to(dev), torch.from_numpy(s).to(dev), torch.from_numpy(seg).to(dev)
    keep, counts, kb, _ = ops.nms_batched(bd, sd, segd, mx, 0.5, None, 200, want_boxes=True)
    counts_h = counts.cpu().numpy()
    boxes = torch.cat([kb[int(seg[i]): int(seg[i]) + int(counts_h[i])] for i in range(args.shelves)])
    idx = torch.repeat_interleave(torch.arange(args.shelves, device=dev), torch.from_numpy(counts_h).to(dev).long()).int()
    images = [torch.rand((3, H, W), device=dev) for _ in range(args.shelves)]
    table = ops.make_image_table(images)
    K = boxes.shape[0]
    bl = boxes.long()
    src = float(((bl[:, 2] - bl[:, 0]).clamp(min=0) * (bl[:, 3] - bl[:, 1]).clamp(min=0)).sum())
    from cvpce_b200 import classification
    aff = classification.tanh_normalize_affine()
    for S in args.sizes:
        for dt, cl, p4 in ((torch.float32, False, False), (torch.float32, True, False), (torch.float32, True, True), (torch.bfloat16, True, False)):
            esz = 4 if dt == torch.float32 else 2
            nbytes = 12.0 * src + (4.0 if p4 else 3.0) * esz * S * S * K  # bytes the kernel really moves (NHWC4 writes 16 B / pixel)
            ms, best = timeit(lambda: ops.crop_resize(table, args.shelves, boxes, idx, S, True, 0.5, aff, dt, cl, pad_channel=p4), args.iters)
            report("crop_resize S=%d %s %s K=%d" % (S, "f32" if esz == 4 else "bf16", ("nhwc4" if p4 else "nhwc") if cl else "nchw", K), ms, best, nbytes,
                   frac_hbm_survey_bytes=round((12.0 * src + 3.0 * esz * S * S * K) / ms / 1e6 / HBM, 4))
    ms, best = timeit(lambda: ops.crop_resize(table, args.shelves, boxes, idx, 256, False, 0.5, aff), args.iters)
    report("crop_resize S=256 f32 nchw plain-bilinear K=%d" % K, ms, best, 12.0 * src + 12.0 * 256 * 256 * K)


def bench_pool(args):
    dev = "cuda"
    for B in (128, 256):
        for cl in (False, True):
            f4 = torch.rand((B, 512, 32, 32), device=dev)
            f5 = torch.rand((B, 512, 16, 16), device=dev)
            if cl:
                f4, f5 = f4.contiguous(memory_format=torch.channels_last), f5.contiguous(memory_format=torch.channels_last)
            out = torch.empty((B, 1024), device=dev)
            nb = 4.0 * (f4.numel() + f5.numel()) + 4.0 * 1024 * B
            ms, best = timeit(lambda: ops.region_pool_embed([f4, f5], out=out), args.iters)
            report("region_pool MAC B=%d %s" % (B, "nhwc" if cl else "nchw"), ms, best, nb)
            if cl:  # the shape inside the step: MAC over pool4(relu4_3) and relu5_3, both 16 x 16
                p4 = torch.rand((B, 512, 16, 16), device=dev).contiguous(memory_format=torch.channels_last)
                nb2 = 4.0 * (p4.numel() + f5.numel()) + 4.0 * 1024 * B
                ms, best = timeit(lambda: ops.region_pool_embed([p4, f5], out=out), args.iters)
                report("region_pool MAC B=%d nhwc, pooled f4 (in-step shape)" % B, ms, best, nb2)
            regs = [ops.rmac_regions(32, 32, 3), ops.rmac_regions(16, 16, 3)]
            ms, best = timeit(lambda: ops.region_pool_embed([f4, f5], regions=regs, per_region_l2=True, out=out), args.iters)
            report("region_pool R-MAC L=3 B=%d %s" % (B, "nhwc" if cl else "nchw"), ms, best, nb)


def bench_rmac(args):
    """R-MAC (3 levels, per-region L2) on the VGG maps: the cluster kernel (rmac_kernel=1) against the persistent TMA-staged kernel
    (rmac_kernel=2) over batch sizes, layouts, dtypes and ring depths."""
    from cvpce_b200 import _lib
    dev = "cuda"
    for B in (32, 64, 128, 256):
        for cl in (True, False):
            for dt in (torch.float32, torch.bfloat16):
                f4 = torch.rand((B, 512, 32, 32), device=dev).to(dt)
                f5 = torch.rand((B, 512, 16, 16), device=dev).to(dt)
                if cl:
                    f4, f5 = f4.contiguous(memory_format=torch.channels_last), f5.contiguous(memory_format=torch.channels_last)
                out = torch.empty((B, 1024), device=dev)
                nb = f4.element_size() * (f4.numel() + f5.numel()) + 4.0 * 1024 * B
                regs = [ops.rmac_regions(32, 32, 3), ops.rmac_regions(16, 16, 3)]
                name = "R-MAC L=3 B=%d %s %s" % (B, "nhwc" if cl else "nchw", "f32" if dt == torch.float32 else "bf16")
                for kern, stages in ((1, 0), (2, 0), (2, 4), (4, 0)):
                    if stages and not (B in (128, 256) and dt == torch.float32):
                        continue
                    if kern == 4 and not cl:
                        continue
                    with _lib.tuning(rmac_kernel=kern, rmac_stages=stages):
                        ms, best = timeit(lambda: ops.region_pool_embed([f4, f5], regions=regs, per_region_l2=True, out=out), args.iters)
                    report(name + (" cluster kernel" if kern == 1 else (" staged kernel, warp-specialised" if kern == 4 else " staged kernel" + (" %d stages" % stages if stages else ""))), ms, best, nb)
                del f4, f5
    # the fast trunk's hand-off: pool4(relu4_3) and relu5_3, both 16 x 16, channels-last f32
    for B in (128, 256):
        p4 = torch.rand((B, 512, 16, 16), device=dev).contiguous(memory_format=torch.channels_last)
        f5 = torch.rand((B, 512, 16, 16), device=dev).contiguous(memory_format=torch.channels_last)
        out = torch.empty((B, 1024), device=dev)
        nb = 4.0 * (p4.numel() + f5.numel()) + 4.0 * 1024 * B
        regs = [ops.rmac_regions(16, 16, 3), ops.rmac_regions(16, 16, 3)]
        for kern in (1, 2, 4):
            with _lib.tuning(rmac_kernel=kern):
                ms, best = timeit(lambda: ops.region_pool_embed([p4, f5], regions=regs, per_region_l2=True, out=out), args.iters)
            report("R-MAC L=3 B=%d nhwc f32, two 16x16 maps (fast-trunk hand-off) %s" % (B, {1: "cluster kernel", 2: "staged kernel", 4: "staged kernel, warp-specialised"}[kern]), ms, best, nb)


def bench_prep(args):
    dev = "cuda"
    for R, D in ((6400, 1024), (125000, 512), (1000000, 512)):
        x = torch.randn((R, D), device=dev)
        ms, best = timeit(lambda: ops.rows_prepare(x, True), args.iters)
        report("rows_prepare %dx%d" % (R, D), ms, best, 6.0 * R * D + 4.0 * R)


def bench_sim(args):
    dev = "cuda"
    shapes = [(200, 180, 1024, 1), (6400, 8192, 1024, 5), (25000, 125000, 512, 5), (100000, 125000, 512, 5)]
    if args.quick:
        shapes = shapes[:3]
    if args.shape:
        shapes = [tuple(int(v) for v in args.shape.split(","))]
    for Q, G, D, k in shapes:
        q = torch.randn((Q, D), device=dev)
        g = torch.randn((G, D), device=dev)
        q /= q.norm(dim=1, keepdim=True)
        g /= g.norm(dim=1, keepdim=True)
        qb, qi = ops.rows_prepare(q, True)
        gb, gi = ops.rows_prepare(g, True)
        ws = torch.empty(ops.sim_topk_workspace_bytes(Q, G, D, k) + 16, dtype=torch.uint8, device=dev)
        fl = 2.0 * Q * G * D
        ms, best = timeit(lambda: ops.sim_topk(qb, gb, k, 0, q, g, qi, gi, workspace=ws), args.iters)
        report("sim_topk exact Q=%d G=%d D=%d k=%d" % (Q, G, D, k), ms, best, flops=fl)
        ms, best = timeit(lambda: ops.sim_topk(qb, gb, k, 0, workspace=ws), args.iters)
        report("sim_topk bf16-only Q=%d G=%d D=%d k=%d" % (Q, G, D, k), ms, best, flops=fl)


def bench_match(args):
    """f3: torchvision-exact box IoU matrix and the greedy claim pass of detection evaluation (C4 density: 512 images x
    (1000 detections vs 150 targets)); the oracle's literal sort-and-walk restatement on one host core beside it."""
    import time
    import oracle
    dev = "cuda"
    for na, nb in ((1000, 150), (8192, 8192), (20000, 16384)):
        a, _ = synth.random_boxes(1, na, extent=1500.0)
        b, _ = synth.random_boxes(2, nb, extent=1500.0)
        ad, bd = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
        ms, best = timeit(lambda: ops.box_iou(ad, bd), args.iters)
        report("box_iou %d x %d" % (na, nb), ms, best, 16.0 * (na + nb) + 4.0 * na * nb)
        ms, best = timeit(lambda: torchvision.ops.box_iou(ad, bd), args.iters)
        report("  torchvision.ops.box_iou (CUDA, eager) same shape", ms, best, 16.0 * (na + nb) + 4.0 * na * nb)
    n_img = args.stress_images
    rows, cols = [], []
    for i in range(n_img):
        t, p = synth.detection_eval_case(700 + i, 150, 1000 - 150 * 3)
        rows.append(p)
        cols.append(t)
    ro = torch.from_numpy(np.cumsum([0] + [len(r) for r in rows]).astype(np.int32)).to(dev)
    co = torch.from_numpy(np.cumsum([0] + [len(c) for c in cols]).astype(np.int32)).to(dev)
    rd, cd = torch.from_numpy(np.concatenate(rows)).to(dev), torch.from_numpy(np.concatenate(cols)).to(dev)
    ms, best = timeit(lambda: ops.iou_match(rd, cd, 0.5, row_offsets=ro, col_offsets=co, max_cols=150), args.iters)
    t0 = time.perf_counter()
    for i in range(8):
        oracle.iou_match(rows[i], cols[i], 0.5)
    cpu_ms = (time.perf_counter() - t0) / 8 * 1e3
    report("iou_match %d images x (1000 x 150), one launch" % n_img, ms, best, 16.0 * (rd.shape[0] + cd.shape[0]) + rd.shape[0] + cd.shape[0],
           images_per_s=round(n_img / ms * 1e3, 1), oracle_cpu_ms_per_image=round(cpu_ms, 3))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("which", nargs="*", default=["nms", "select", "crop", "pool", "prep", "sim"])
    ap.add_argument("--shelves", type=int, default=32)
    ap.add_argument("--stress-images", type=int, default=512)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--sizes", type=int, nargs="*", default=[256, 224])
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--shape", default=None, help="sim: one shape Q,G,D,k instead of the list")
    ap.add_argument("--tune", nargs="*", default=[], help="name=value overrides passed to cvpce_tuning_set")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    from cvpce_b200 import _lib
    for t in args.tune:
        name, value = t.split("=")
        _lib.tuning_set(name, int(value))
    if args.tune:
        print(json.dumps({"tuning": args.tune}), flush=True)
    for w in args.which:
        {"nms": bench_nms, "select": bench_select, "maxpool": bench_maxpool, "match": bench_match, "crop": bench_crop, "pool": bench_pool, "rmac": bench_rmac, "prep": bench_prep, "sim": bench_sim}[w](args)


if __name__ == "__main__":
    main()
