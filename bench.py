#!/usr/bin/env python
"""bench.py -- proposal crops classified / second through the B200-native hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2|c4|c5]

A "step" is one pass of the whole hot path over one batch of synthetic shelf images:
scored candidate boxes -> NMS -> crop/pad/resize -> VGG16 trunk (torch, not rebuilt: SURVEY §8 a7) -> MAC pool + L2 ->
cosine top-k against the gallery.  Default workload = BASELINE.json configs[2] ("c3": 32 shelves of 2048x1536 with ~200
products each per GPU, gallery 8192 x 1024): the largest single-GPU configuration; with N GPUs every rank processes its
own 32 shelves (images are independent: weak scaling, no collective in the data path; configs[3] at N=8 -> 256 shelves).
`value` is measured with inputs resident in HBM; `e2e` through the same public call with pinned HOST buffers (H2D of the
shelf images and candidates, D2H of the result indices inside the timed region).

--impl reference : the reference's own CPU implementation of the path (oracle/torch_port.py: the same torch/torchvision CPU
ops in the same order; /root/reference itself cannot travel to the GPU box), all host threads, each step a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# crops per cuDNN call of the torch trunk (and per pooling launch).  Measured on one box, alternating runs (profiles/r4wx_trunk_batch_sweep_in_step.jsonl):
# 256 -> 14.32-14.51k crops/s with the MAC launch at 0.70 of HBM; 512 -> 14.49-14.71k, 0.80; 800 -> 14.64-14.69k, 0.855;
# 1600 -> 14.71k, 0.91.  800 = the crops of one upload group of the end-to-end leg (4 shelves x 200) and 6400 / 8.
TRUNK_BATCH = 800
METRIC = "proposal crops classified/sec"
UNIT = "crops/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--shelves", type=int, default=None, help="shelf images per GPU per step (c3: 32, c2: 1, c4: 512 / n_gpus)")
    ap.add_argument("--products", type=int, default=200)
    ap.add_argument("--gallery", type=int, default=None)
    ap.add_argument("--k", type=int, default=5)
    ap.add_argument("--trunk", default="f32", choices=["f32", "f16", "bf16"],
                    help="dtype of the torch VGG16 trunk: f32 = torch's GPU default (TF32 convolutions, the headline); f16 = half weights / "
                         "activations with fp32 accumulation (same 10-bit mantissa as TF32: same measured embedding deviation, "
                         "tools/f16_probe.py); bf16 = 7-bit mantissa, does not meet the 1e-3 tolerance")
    ap.add_argument("--channels-last", action="store_true")
    ap.add_argument("--plain-trunk", action="store_true", help="call the torch trunk module by module in NCHW (the reference's way) "
                    "instead of channels-last + torch.cudnn_convolution_relu")
    ap.add_argument("--trunk-batch", type=int, default=TRUNK_BATCH)
    ap.add_argument("--streams", type=int, default=1, help="CUDA streams the trunk batches alternate between")
    ap.add_argument("--cudnn-benchmark", action="store_true")
    ap.add_argument("--strict-fp32", action="store_true", help="torch.backends.cudnn.allow_tf32 = False for the torch trunk (the default "
                    "is torch's own GPU default, TF32 convolutions: tools/tf32_check.py measures what that costs in accuracy)")
    ap.add_argument("--sim", default="tc", choices=["tc", "f32"], help="similarity kernel: tcgen05 or CUDA-core fp32")
    ap.add_argument("--queries", type=int, default=100000, help="c5 only")
    ap.add_argument("--c5-chunks", type=int, default=4, help="c5 e2e leg: query chunks (upload of chunk c+1 overlaps the search of chunk c)")
    ap.add_argument("--no-other-configs", action="store_true", help="headline only: skip the c2 / c1 / strict-fp32 / c5 legs reported under other_configs")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="c5, N > 1: how the per-rank top-k blocks reach the other "
                    "ranks: stored by the selection kernel into every peer's gather buffer over NVLink, or an NCCL all-gather")
    ap.add_argument("--from-head", action="store_true", help="start one stage earlier (SURVEY f1): synthetic GLN head outputs "
                    "(800x1066 frame, 5 FPN levels) -> candidate selection -> NMS -> ... instead of scored candidate boxes")
    ap.add_argument("--no-pad-channel", action="store_true", help="3-channel crops (the fast trunk's default is 4-channel crops with a "
                    "zero 4th channel so that cuDNN skips its channel-padding pass before conv1_1)")
    ap.add_argument("--e2e-serial", action="store_true", help="e2e leg: upload all images before the step instead of overlapping "
                    "the upload with the step's own compute")
    ap.add_argument("--upload-group", type=int, default=4, help="e2e leg: shelf images per upload group")
    ap.add_argument("--e2e-prefetch", type=int, default=0, help="e2e leg: 1 = the next step's images are uploaded (second staging buffer) behind this step's compute; 0 = every step uploads its own images first")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rmac-levels", type=int, default=0, help="0: MAC pooling (the reference's mode, the headline); L > 0: R-MAC with L scales "
                    "(north-star extension: regions + per-region L2 over relu4_3 and relu5_3, the persistent TMA-staged kernel)")
    ap.add_argument("--no-extension-modes", action="store_true", help="skip the short extra pass that also reports the f16-trunk "
                    "extension mode next to the headline (N = 1, default trunk only)")
    ap.add_argument("--cpu-sample-crops", type=int, default=32)
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------ helpers
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.proc = None
        self.lines = []
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm": float(p["hbm_gbs"]), "tensor_burst": float(p["bf16_tflops"]), "tensor_sustained": float(p["bf16_tflops_sustained"]),
                "source": "MEASURED_PEAKS.json"}
    except Exception:
        return {"hbm": 6650.0, "tensor_burst": 1590.0, "tensor_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def ncu_traffic(kernel, algorithmic_bytes):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed `ncu --set full` capture of
    this command (profiles/ncu_traffic.json, written by tools/ncu_summary.py traffic); None when no capture matches the
    workload (the capture's algorithmic bytes per launch must agree within 5 %)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f).get(kernel)
        if t and abs(t["algorithmic_bytes_per_launch"] - algorithmic_bytes) <= 0.05 * algorithmic_bytes:
            return t["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    elif torch.cuda.is_available():
        torch.cuda.set_device(0)
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms, world):
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return ms


def sum_over_ranks(v, world):
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())
    return v


# ----------------------------------------------------------------------------------------------------- workload (c2/c3)
def build_shelves(args, rank, n_shelves, device):
    from cvpce_b200 import synth
    H, W = synth.SHELF_HW
    images, boxes, scores, sizes = [], [], [], []
    for i in range(n_shelves):
        gi = rank * n_shelves + i
        gen = torch.Generator(device=device).manual_seed(100 + gi)
        images.append(torch.rand((3, H, W), generator=gen, device=device))
        b, s = synth.shelf_candidates(100 + gi, H, W, args.products)
        boxes.append(b)
        scores.append(s)
        sizes.append(len(s))
    seg = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    return images, np.concatenate(boxes), np.concatenate(scores), seg, max(sizes)


def build_encoder(args, device):
    from cvpce_b200 import classification, synth
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(device)
    if args.trunk == "bf16":
        enc = enc.to(torch.bfloat16)
    elif args.trunk == "f16":
        enc = enc.to(torch.float16)
    if not args.plain_trunk:
        enc.enable_fast_trunk()
        args.channels_last = True
    elif args.channels_last:
        enc = enc.to(memory_format=torch.channels_last)
    return enc


def build_gallery(G, D, device, seed=3):
    gen = torch.Generator(device=device).manual_seed(seed)
    g = torch.randn((G, D), generator=gen, device=device).abs_()
    return g / g.norm(dim=1, keepdim=True)


def run_ours_shelves(args, world, rank, local):
    from cvpce_b200 import _lib, pipeline, classification
    device = torch.device("cuda", local)
    _lib.load()
    peaks = measured_peaks()
    if args.workload == "c4":
        # BASELINE.json configs[3]: 512 SKU-110K-density shelves (~150 boxes each) split data-parallel over the GPUs
        # (strong scaling: the 512 images are fixed); gallery as in c3 (BASELINE leaves it open)
        n_shelves = args.shelves or max(1, 512 // world)
        if args.products == 200:
            args.products = 150
    else:
        n_shelves = args.shelves or (32 if args.workload == "c3" else 1)
    G = args.gallery or (8192 if args.workload in ("c3", "c4") else 180)
    enc = build_encoder(args, device)
    gallery = build_gallery(G, enc.embedding_size, device)
    images, boxes_h, scores_h, seg_h, max_n = build_shelves(args, rank, n_shelves, device)
    hp = pipeline.HotPath(enc, gallery, k=args.k, trunk_batch=args.trunk_batch,
                          trunk_dtype={"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[args.trunk], channels_last=args.channels_last,
                          n_streams=args.streams, pad_channel=not args.no_pad_channel, rmac_levels=args.rmac_levels)
    if args.sim == "f32":  # CUDA-core similarity instead of the tcgen05 kernel (debug / cross-check)
        from cvpce_b200 import ops
        ops_sim = ops.sim_topk

        def sim_f32(qb, gb, k, base, q32, g32, qi, gi, workspace=None):
            return ops.sim_topk_f32(q32, g32, k, base, normalize=True)
        ops.sim_topk = sim_f32
    boxes_d = torch.from_numpy(boxes_h).to(device)
    scores_d = torch.from_numpy(scores_h).to(device)
    seg_d = torch.from_numpy(seg_h).to(device)
    head_h = None
    if args.from_head:
        # head outputs whose decoded candidates are the same scored boxes (synth.head_from_candidates), kept as one
        # concatenated tensor per output split into per-level views, like torchvision's RetinaNet does
        from cvpce_b200 import synth
        lgs, rgs = [], []
        for i in range(n_shelves):
            lg, rg, anc, scale = synth.head_from_candidates(boxes_h[seg_h[i]:seg_h[i + 1]], scores_h[seg_h[i]:seg_h[i + 1]],
                                                            synth.SHELF_HW, 1000 + rank * n_shelves + i)
            lgs.append(lg)
            rgs.append(rg)
        n_per = [a.shape[0] for a in anc]
        head_h = (torch.from_numpy(np.stack(lgs)).pin_memory(), torch.from_numpy(np.stack(rgs)).pin_memory())
        anchors_d = [torch.from_numpy(a).to(device) for a in anc]
        frame_d = torch.tensor([list(synth.GLN_FRAME)] * n_shelves, dtype=torch.int32, device=device)
        scale_d = torch.tensor([[float(scale[0]), float(scale[1])]] * n_shelves, dtype=torch.float32, device=device)
        head_d = (head_h[0].to(device), head_h[1].to(device))

    def run_step(imgs, cand=None, head=None, uploads=None):
        if args.from_head:
            lg, rg = head if head is not None else head_d
            return hp.step_from_head(imgs, list(lg.split(n_per, dim=1)), list(rg.split(n_per, dim=1)), anchors_d, frame_d, scale_d,
                                     uploads=uploads)
        b, sc, sg = cand if cand is not None else (boxes_d, scores_d, seg_d)
        return hp.step(imgs, b, sc, sg, max_n, uploads=uploads)

    # -------- device-resident timing ("value")
    for _ in range(args.warmup):
        out = run_step(images)
    K_crops = int(out["counts"].sum())
    pad4_crops = bool(args.channels_last and not args.no_pad_channel and args.trunk == "f32" and not args.plain_trunk)
    hp.launches = 0
    hp.timer = pipeline.StageTimer()
    sampler = ClockSampler(local) if rank == 0 else None
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = run_step(images)
    e1.record()
    barrier(world)
    ms_rank = e0.elapsed_time(e1)
    per_rank = gather_per_rank(ms_rank / args.steps, world)
    ms_total = max_over_ranks(ms_rank, world)
    clocks = sampler.stop() if sampler else None
    launches = hp.launches
    stages = hp.timer.summary()
    hp.timer = pipeline._NoTimer()
    total_crops = sum_over_ranks(float(K_crops), world)
    value = total_crops * args.steps / (ms_total / 1e3)

    # -------- end-to-end timing: pinned host inputs, H2D + D2H inside the timed region
    images_h = [im.cpu().pin_memory() for im in images]
    boxes_p, scores_p, seg_p = torch.from_numpy(boxes_h).pin_memory(), torch.from_numpy(scores_h).pin_memory(), torch.from_numpy(seg_h).pin_memory()
    h2d = sum(im.numel() * 4 for im in images_h)
    h2d += (head_h[0].numel() + head_h[1].numel()) * 4 if args.from_head else (boxes_p.numel() + scores_p.numel() + seg_p.numel()) * 4
    del images

    def e2e_upload():
        # the step's images go up on the library's copy stream in groups (HotPath.upload_images) and are consumed group by
        # group, so the transfer overlaps convolutions; --e2e-serial uploads everything first
        if args.e2e_serial:
            return [im.to(device, non_blocking=True) for im in images_h], None
        return hp.upload_images(images_h, args.upload_group)

    def e2e_step(up=None, prefetch=False):
        imgs, ups = up if up is not None else e2e_upload()
        # double-buffered staging: the NEXT step's images start moving now, behind this step's compute (a serving loop that
        # knows its next batch, like the reference's DataLoader prefetch); every step's H2D copy is still inside the timed region
        nxt = e2e_upload() if prefetch else None
        if args.from_head:
            o = run_step(imgs, head=(head_h[0].to(device, non_blocking=True), head_h[1].to(device, non_blocking=True)), uploads=ups)
        else:
            o = run_step(imgs, cand=(boxes_p.to(device, non_blocking=True), scores_p.to(device, non_blocking=True),
                                     seg_p.to(device, non_blocking=True)), uploads=ups)
        idx = o["topk_idx"].cpu()
        bx = o["boxes"].cpu()
        return idx, bx, nxt

    prefetch = bool(args.e2e_prefetch) and not args.e2e_serial
    for _ in range(max(1, min(args.warmup, 2))):
        idx_h, bx_h, _ = e2e_step()
    d2h = idx_h.numel() * 8 + bx_h.numel() * 4 + n_shelves * 4
    barrier(world)
    e0.record()
    up = None
    for s_i in range(args.steps):
        _, _, up = e2e_step(up, prefetch and s_i + 1 < args.steps)
    e1.record()
    barrier(world)
    ms_e2e_rank = e0.elapsed_time(e1)
    per_rank_e2e = gather_per_rank(ms_e2e_rank / args.steps, world)
    ms_e2e = max_over_ranks(ms_e2e_rank, world)
    e2e_value = total_crops * args.steps / (ms_e2e / 1e3)

    # -------- per-kernel rooflines from the events recorded inside the timed region
    kernels = {}
    for name, d in stages.items():
        ms = d["ms"] / d["launches"]
        ent = {"launches_per_step": d["launches"] / args.steps, "avg_ms": round(ms, 4)}
        if name == "sim_topk":
            tf = d["flops"] / d["launches"] / (ms / 1e3) / 1e12
            # timed inside a long, power-capped step: the sustained bf16 figure is the denominator (the burst one is for a
            # kernel timed alone, as tools/kbench.py does)
            ent.update({"bound": "tensor", "achieved_tflops": round(tf, 2), "frac": round(tf / peaks["tensor_sustained"], 4),
                        "peak_tflops": peaks["tensor_sustained"]})
        else:
            gbs = d["bytes"] / d["launches"] / (ms / 1e3) / 1e9
            ent.update({"bound": "hbm", "achieved_gbs": round(gbs, 1), "frac": round(gbs / peaks["hbm"], 4),
                        "algorithmic_bytes_per_launch": d["bytes"] / d["launches"]})
        ent["share_of_step"] = round(d["ms"] / ms_total, 4)
        if name == "crop_resize" and pad4_crops:
            # SURVEY's algorithmic bytes count 12 B per output pixel; the 4-channel layout the trunk is fed WRITES 16 B per pixel
            moved = d["bytes"] + 4.0 * 256 * 256 * K_crops * args.steps
            ent["bytes_moved_per_launch"] = moved / d["launches"]
            ent["frac_of_bytes_moved"] = round(moved / d["launches"] / (ms / 1e3) / 1e9 / peaks["hbm"], 4)
        if name == "maxpool2x2":
            ent["note"] = "a7: drives the torch trunk (not a SURVEY 8(a) kernel of the path itself)"
        kernels[name] = ent
    ours_ms = sum(d["ms"] for d in stages.values())
    # `roofline` = the kernel of the PATH with the largest share of the step, summed over its launches; the
    # 2x2 max-pools that drive the torch trunk (a7) take more of the step and are listed under `kernels` with their own fraction
    path_stages = {n: d for n, d in stages.items() if n != "maxpool2x2"} or stages
    dom = max(path_stages, key=lambda n: path_stages[n]["ms"])
    dd = stages[dom]
    traffic = ncu_traffic(dom, dd["bytes"] / dd["launches"])
    if dom == "sim_topk":
        dom_tf = dd["flops"] / dd["launches"] / (dd["ms"] / dd["launches"] / 1e3) / 1e12
        roofline = {"kernel": dom, "bound": "tensor", "achieved": round(dom_tf, 1), "peak": peaks["tensor_sustained"], "unit": "TFLOP/s",
                    "frac": round(dom_tf / peaks["tensor_sustained"], 4), "traffic": traffic,
                    "peak_source": peaks["source"] + " (bf16_tflops_sustained: timed inside a long step)"}
    else:
        dom_gbs = dd["bytes"] / dd["launches"] / (dd["ms"] / dd["launches"] / 1e3) / 1e9
        roofline = {"kernel": dom, "bound": "hbm", "achieved": round(dom_gbs, 1), "peak": peaks["hbm"], "unit": "GB/s",
                    "frac": round(dom_gbs / peaks["hbm"], 4), "traffic": traffic,
                    "launches_per_step": dd["launches"] / args.steps, "avg_launch_ms": round(dd["ms"] / dd["launches"], 4),
                    "algorithmic_bytes_per_launch": dd["bytes"] / dd["launches"],
                    "peak_source": peaks["source"] + " (hbm_gbs, measured copy bandwidth)"}
        for extra in ("bytes_moved_per_launch", "frac_of_bytes_moved"):
            if extra in kernels[dom]:
                roofline[extra] = kernels[dom][extra]

    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_total / args.steps, 3),
        "ms_per_step_per_rank": {"min": round(min(per_rank), 3), "max": round(max(per_rank), 3)}, "higher_is_better": True,
        "scaling": "strong" if args.workload == "c4" and not args.shelves else "weak", "vs_baseline": None,
        # the arithmetic type of the path: the library's own kernels compute in fp32 in every mode; the torch trunk (not rebuilt)
        # runs cuDNN TF32 convolutions unless --strict-fp32 (reported next to the headline under other_configs.strict_fp32)
        "dtype": ({"f32": "tf32 trunk convolutions (torch/cuDNN default on GPU), f32 kernels" if torch.backends.cudnn.allow_tf32
                   else "f32 (strict: fp32 convolutions and kernels)", "f16": "f16 trunk, f32 kernels", "bf16": "bf16 trunk, f32 kernels"}[args.trunk]),
        "data": "synthetic",
        "config": {"workload": ("%s: %d shelves 2048x1536 x ~%d products per GPU -> " + ("GLN head outputs (800x1066, 160956 anchors) -> "
                                "sigmoid/top-1000/decode -> " if args.from_head else "") + "NMS(0.5, top200) -> 256x256 crops -> "
                                "VGG16(random init)+%s -> cosine top-%d vs %d x 1024 gallery") % (args.workload, n_shelves, args.products,
                                ("R-MAC(%d levels)" % args.rmac_levels) if args.rmac_levels else "MAC", args.k, G),
                   "crops_per_step_per_gpu": K_crops, "trunk": "torch cuDNN VGG16 convolutions, not rebuilt (%s%s%s, tf32 convs %s)%s" % (
                       args.trunk, ", channels_last" if args.channels_last else "",
                       "" if args.plain_trunk else ", torch.cudnn_convolution_relu", torch.backends.cudnn.allow_tf32,
                       "" if args.plain_trunk else "; the 2x2 max-pools between them run in the library (cvpce_maxpool2x2_nhwc)" +
                       ("" if args.no_pad_channel else "; crops carry a zero 4th channel (conv1_1 against a zero-padded weight, same output)")),
                   "trunk_batch": args.trunk_batch,
                   "similarity": "tcgen05 bf16 + exact fp32 re-score" if args.sim == "tc" else "cuda-core fp32",
                   "l2": ("inputs (%.1f GB of shelf images per step) larger than L2" % (n_shelves * 3 * 1536 * 2048 * 4 / 1e9))
                         if n_shelves * 3 * 1536 * 2048 * 4 > 126e6 else
                         ("inputs (%.0f MB) fit L2, but every step streams %.1f GB of crops and trunk activations through it before "
                          "the inputs are read again" % (n_shelves * 3 * 1536 * 2048 * 4 / 1e6, K_crops * (0.79e6 + 60e6) / 1e9)),
                   "parallelism": "dp%d (independent images, no data-path collective)" % world},
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": round(ms_e2e / args.steps, 3),
                "ms_per_step_per_rank": {"min": round(min(per_rank_e2e), 3), "max": round(max(per_rank_e2e), 3)},
                "upload": "all images before the step" if args.e2e_serial else
                          ("copy stream, groups of %d shelves, overlapped with compute (HotPath.upload_images); " % args.upload_group) +
                          ("two staging buffers: step s+1's images go up behind step s" if args.e2e_prefetch else "each step uploads its own images first")},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "kernels": kernels,
        "own_kernels_share_of_step": round(ours_ms / ms_total, 4),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # BASELINE.md section 3 protocol: 1 warm-up + best of 3 on all host threads with per-stage seconds, plus a 1-thread figure
        line["cpu_baseline"] = cpu_reference_rate(args, G, sample_crops=args.cpu_sample_crops, repeats=3, warmup=1, one_thread_crops=8)
    return line


# ------------------------------------------------------------------------------------------- reference / cpu baseline
NN_NATIVE_BUDGET = 8e9  # bytes of the two gathered [Q,G,D] fp32 tensors the reference form of nearest_neighbors may allocate


def cpu_reference_rate(args, G, sample_crops, repeats, warmup=0, one_thread_crops=0):
    """Times the CPU port of the reference path (oracle/torch_port.py: the reference's torch / torchvision CPU ops in the
    reference's order) on a bounded sample: one shelf's NMS of all candidates, then the first `sample_crops` surviving boxes
    through crop loop -> VGG16 fp32 -> MAC -> nearest_neighbors, on all host threads; best of `repeats` after `warmup`, with the
    seconds of each stage of the best run.  `one_thread_crops` > 0 adds a single-thread figure."""
    from oracle import torch_port
    from cvpce_b200 import synth
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    H, W = synth.SHELF_HW
    torch.manual_seed(synth.WEIGHT_SEED)
    enc = torch_port.MACVGGPort().eval().requires_grad_(False)
    gen = torch.Generator().manual_seed(3)
    gallery = torch.randn((G, 1024), generator=gen).abs_()
    gallery = gallery / gallery.norm(dim=1, keepdim=True)
    image = torch.rand((3, H, W), generator=torch.Generator().manual_seed(100))
    b, s = synth.shelf_candidates(100, H, W, args.products)
    b, s = torch.from_numpy(b), torch.from_numpy(s)
    batch = 32  # Classifier's default batch_size (production.py:23)
    # nearest_neighbors in the reference's own form gathers two [Q,G,D] fp32 tensors: run it natively on the reference's
    # batch of 32 queries when that fits the budget, else in query chunks (a restatement of how it would have to be called)
    nn_chunk = None if 2.0 * batch * G * 1024 * 4 <= NN_NATIVE_BUDGET else max(1, int(NN_NATIVE_BUDGET // (2.0 * G * 1024 * 4)))
    times, stage_runs = [], []
    n = 0
    for it in range(warmup + repeats):
        stages = {}
        t0 = time.perf_counter()
        boxes, idx = torch_port.hot_path(enc, gallery, image, b, s, k=args.k, detections_per_img=200, max_crops=sample_crops,
                                         batch_size=batch, nn_chunk=nn_chunk, stages=stages)
        dt = time.perf_counter() - t0
        n = len(boxes)
        if it >= warmup:
            times.append(dt)
            stage_runs.append(stages)
    best = int(np.argmin(times))
    out = {"value": round(n / times[best], 3), "unit": UNIT, "cores": threads, "kind": "port",
           "sample": "1 shelf: NMS of %d candidates, first %d kept boxes -> crop loop -> VGG16 fp32 -> MAC -> nearest_neighbors vs %d x 1024 "
                     "(%s); torch CPU ops of the reference, %d warm-up + best of %d"
                     % (len(s), n, G, "reference-native form, batches of %d queries" % batch if nn_chunk is None
                        else "reference form in chunks of %d queries (nn_chunk: the native batch of %d would gather %.1f GB)" % (nn_chunk, batch, 2.0 * batch * G * 4096 / 1e9),
                        warmup, repeats),
           "seconds_per_step": [round(t, 3) for t in times],
           "stage_seconds": {k: round(v, 4) for k, v in stage_runs[best].items()},
           "stage_note": "NMS covers all %d candidates of the shelf; the other stages cover the %d sampled crops" % (len(s), n)}
    if one_thread_crops > 0:
        torch.set_num_threads(1)
        try:
            st1 = {}
            t0 = time.perf_counter()
            boxes1, _ = torch_port.hot_path(enc, gallery, image, b, s, k=args.k, detections_per_img=200, max_crops=one_thread_crops,
                                            batch_size=batch, nn_chunk=nn_chunk, stages=st1)
            dt1 = time.perf_counter() - t0
            out["one_thread"] = {"value": round(len(boxes1) / dt1, 3), "unit": UNIT, "cores": 1, "crops": len(boxes1),
                                 "seconds": round(dt1, 3), "stage_seconds": {k: round(v, 4) for k, v in st1.items()}, "runs": 1}
        finally:
            torch.set_num_threads(threads)
    return out


def run_reference(args, world, rank):
    if rank != 0:
        return None
    G = args.gallery or (8192 if args.workload in ("c3", "c4") else 180)
    base = cpu_reference_rate(args, G, sample_crops=args.cpu_sample_crops, repeats=args.steps, warmup=args.warmup)
    secs = base.pop("seconds_per_step")
    n = args.cpu_sample_crops
    value = n * len(secs) / sum(secs)
    base["value"] = round(value, 3)
    return {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(1e3 * sum(secs) / len(secs), 2), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "%s (bounded sample per step: %s)" % (args.workload, base["sample"]),
                       "processes": "ONE CPU process on rank 0 with all host threads, whatever --gpus says (the other ranks exit 0): "
                                    "the reference's CPU path is not data-parallel over GPUs",
                       "value_is": "crops of the sample / seconds, mean over the timed steps; per-crop extrapolation to the repo arm's "
                                   "%s crops per step is linear (encoder-bound)" % ("~6400" if args.workload == "c3" else "larger")},
            "cpu_baseline": base,
            "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


# ------------------------------------------------------------------------------------------------------- workload c1
def run_c1(args, device, with_cpu):
    """BASELINE.json configs[0]: DIHE random-init VGG16 embed + MAC + cosine top-1 for 64 synthetic crops vs a 180-item gallery,
    through the Classifier mirror (production.py:57-74): S = 224 (as BASELINE states) and S = 256 (the reference's native
    CLASSIFICATION_IMAGE_SIZE).  `value`: crops already on the device; `e2e`: crops in pinned host memory, indices back on the host
    (classify() returns python lists of annotation strings, like the reference).  With `with_cpu` the reference's CPU path
    (oracle/torch_port.py) is timed on exactly this configuration (1 warm-up + best of 2) and top-1 agreement is reported."""
    from cvpce_b200 import production
    from oracle import torch_port
    a1 = argparse.Namespace(**vars(args))
    a1.trunk = "f32"
    enc = build_encoder(a1, device)
    res = {}
    gen = torch.Generator().manual_seed(1000)
    gal_imgs = torch.rand((180, 3, 256, 256), generator=gen) * 2 - 1  # gallery images arrive in [-1,1] (datautils.py:412-414)
    samples = [(gal_imgs[i], gal_imgs[i], [], "sku%03d" % i) for i in range(180)]
    clf = production.Classifier(enc, samples, device=device, emb_device=device, batch_size=32, num_workers=0, k=1)
    cpu_enc = cpu_gal = None
    if with_cpu:
        torch.manual_seed(__import__("cvpce_b200.synth", fromlist=["x"]).WEIGHT_SEED)
        cpu_enc = torch_port.MACVGGPort().eval().requires_grad_(False)
        with torch.no_grad():
            cpu_gal = torch.cat([cpu_enc(gal_imgs[i:i + 32]) for i in range(0, 180, 32)])
    for S in (224, 256):
        crops_h = torch.rand((64, 3, S, S), generator=torch.Generator().manual_seed(7)).pin_memory()
        crops_d = crops_h.to(device)
        steps, warm = 20, 5
        for _ in range(warm):
            lab = clf.classify(crops_d)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            lab = clf.classify(crops_d)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        for _ in range(2):
            clf.classify(crops_h)
        e0.record()
        for _ in range(steps):
            lab_h = clf.classify(crops_h)
        e1.record()
        torch.cuda.synchronize()
        ms_e2e = e0.elapsed_time(e1) / steps
        ent = {"workload": "c1: 64 synthetic %dx%d crops -> VGG16(random init)+MAC -> cosine top-1 vs 180 x 1024 gallery (Classifier.classify)" % (S, S),
               "value": round(64 / ms * 1e3, 1), "unit": UNIT, "ms_per_step": round(ms, 3), "steps": steps,
               "e2e": round(64 / ms_e2e * 1e3, 1), "e2e_ms_per_step": round(ms_e2e, 3), "h2d_bytes_per_step": int(crops_h.numel() * 4),
               "d2h_bytes_per_step": 64 * 8, "same_labels_resident_and_e2e": lab == lab_h}
        if with_cpu:
            ts = []
            for it in range(3):
                st = {}
                t0 = time.perf_counter()
                idx = torch_port.classify(cpu_enc, cpu_gal, crops_h, k=1, batch_size=32, stages=st)
                ts.append((time.perf_counter() - t0, st))
            best = min(ts[1:], key=lambda t: t[0])
            ent["cpu_baseline"] = {"value": round(64 / best[0], 3), "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                                   "sample": "the whole configuration (64 crops, reference-native nearest_neighbors), 1 warm-up + best of 2",
                                   "seconds": round(best[0], 3), "stage_seconds": {k: round(v, 4) for k, v in best[1].items()}}
            ent["top1_agreement_with_cpu_fp32"] = round(float(np.mean([lab[i][0] == "sku%03d" % int(idx[i, 0]) for i in range(64)])), 4)
        res["S%d" % S] = ent
    del clf, enc
    torch.cuda.empty_cache()
    return res


# ----------------------------------------------------------------------------------------------------- workload c5
def gather_per_rank(ms, world):
    """[ms of rank 0, ..., ms of rank world-1] on every rank."""
    if world == 1:
        return [ms]
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [float(x.item()) for x in out]


def c5_parity_check(args, searcher, q, k, world, rank, device, n_check=512):
    """In-run check on a query subsample: the search through the configured exchange (peer scatter) == the NCCL all-gather
    form == a single-rank search of the WHOLE gallery (rank 0 regenerates all shards' rows: the synthetic generator does not
    depend on the sharding), bit for bit -- scores and indices.  Returns (ok, detail)."""
    from cvpce_b200 import dist as cdist
    import torch.distributed as dist
    G, D = searcher.global_rows, searcher.D
    sub = q[torch.linspace(0, q.shape[0] - 1, n_check, device=device).long()].contiguous()
    s_a, i_a = searcher.search(sub, k)
    ok, detail = True, {}
    if world > 1:
        other = cdist.ShardedGallery(searcher.g32, searcher.idx_base, world, rank, searcher.group, exchange="nccl", global_rows=G)
        s_b, i_b = other.search(sub, k)
        same = bool(torch.equal(s_a, s_b) and torch.equal(i_a, i_b))
        detail["p2p_equals_nccl" if searcher.exchange == "p2p" else "nccl_equals_nccl"] = same
        ok = ok and same
        # partitioned-queries form (what the e2e leg runs) against the resident-queries form
        lo, hi = cdist.shard_range(n_check, world, rank)
        s_c, i_c = searcher.search_partitioned(sub[lo:hi], lo, n_check, k)
        same = bool(torch.equal(s_a, s_c) and torch.equal(i_a, i_c))
        detail["partitioned_equals_resident"] = same
        ok = ok and same
        del other
    if rank == 0:
        whole = cdist.ShardedGallery.synthetic(G, D, 1, 0, device, seed=5)
        s_w, i_w = whole.search(sub, k)
        same = bool(torch.equal(s_a, s_w) and torch.equal(i_a, i_w))
        detail["sharded_equals_single_rank_whole_gallery"] = same
        ok = ok and same
        del whole
        torch.cuda.empty_cache()
    if world > 1:
        flag = torch.tensor([1 if ok else 0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(int(flag.item()))
    return ok, detail


def run_ours_c5(args, world, rank, local):
    """BASELINE.json configs[4]: 1M x 512 gallery sharded over the ranks, 100k queries, cosine top-5, all-gather + merge.
    Strong scaling: the gallery is fixed, each rank searches G/world rows for ALL queries.
    `value`: queries resident in HBM on every rank.  `e2e`: the queries start in PINNED HOST memory, partitioned over the ranks
    (each rank uploads 1/world of them; cvpce_rows_prepare_scatter replicates the block to every rank over NVLink), each rank
    downloads its block of the merged result; upload of query chunk c+1 overlaps the search of chunk c."""
    from cvpce_b200 import dist as cdist
    device = torch.device("cuda", local)
    peaks = measured_peaks()
    G = args.gallery if (args.gallery and args.workload == "c5") else 1_000_000
    Q, D, k = args.queries, 512, args.k
    steps, warmup = args.steps, max(args.warmup, 3)
    searcher = cdist.ShardedGallery.synthetic(G, D, world, rank, device, seed=5, exchange=args.exchange)
    q = cdist.synthetic_queries(Q, D, device, seed=6)
    # the timed region is tens of milliseconds: start sampling clocks before the warm-up so that the sampler (200 ms period)
    # has readings taken under the same load
    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(warmup):
        sim, idx = searcher.search(q, k)
    barrier(world)
    l0 = searcher.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sim, idx = searcher.search(q, k)
    e1.record()
    barrier(world)
    ms_rank = e0.elapsed_time(e1)
    per_rank = gather_per_rank(ms_rank / steps, world)
    ms = max_over_ranks(ms_rank, world)
    launches = searcher.launches - l0
    # ---- end to end: queries in pinned host memory, partitioned over the ranks and over `chunks` query chunks
    chunks = max(1, args.c5_chunks)
    parts = cdist.partition_query_chunks(Q, world, rank, chunks)
    blocks_h = [(q[g0: g0 + n].cpu().pin_memory(), lo, qc) for (g0, n, lo, qc) in parts]
    out_h = [(torch.empty((n, k), dtype=torch.float32).pin_memory(), torch.empty((n, k), dtype=torch.int64).pin_memory())
             for (_, n, _, _) in parts]
    for _ in range(2):
        res = searcher.search_host(blocks_h, k, out=out_h)
    # the e2e result (this rank's rows) equals the resident-queries result
    e2e_same = all(bool(torch.equal(sh, sim[g0: g0 + n].cpu()) and torch.equal(ih, idx[g0: g0 + n].cpu()))
                   for (sh, ih), (g0, n, _, _) in zip(res, parts))
    barrier(world)
    e0.record()
    for _ in range(steps):
        searcher.search_host(blocks_h, k, out=out_h)
    e1.record()
    barrier(world)
    ms_e2e_rank = e0.elapsed_time(e1)
    per_rank_e2e = gather_per_rank(ms_e2e_rank / steps, world)
    ms_e2e = max_over_ranks(ms_e2e_rank, world)
    clocks = sampler.stop() if sampler else None
    parity_ok, parity = c5_parity_check(args, searcher, q, k, world, rank, device)
    parity["e2e_rows_equal_resident"] = e2e_same
    parity_ok = parity_ok and e2e_same
    flops = 2.0 * Q * G * D * steps
    tf = flops / (ms / 1e3) / 1e12 / world
    h2d = sum(b[0].numel() * 4 for b in blocks_h)
    d2h = sum(sh.numel() * 4 + ih.numel() * 8 for sh, ih in out_h)
    line = {"metric": "gallery queries matched/sec (cosine top-%d)" % k, "value": round(Q * steps / (ms / 1e3), 1), "unit": "queries/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": round(ms / steps, 3),
            "ms_per_step_per_rank": {"min": round(min(per_rank), 3), "max": round(max(per_rank), 3)},
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16 operands, f32 accumulate + exact f32 re-score",
            "data": "synthetic", "config": {"workload": "c5: %d x %d gallery sharded over %d GPU(s), %d queries, top-%d, all-gather merge" % (G, D, world, Q, k),
                               "exchange": "none (1 GPU)" if world == 1 else ("selection-kernel stores into peer gather buffers over NVLink (symmetric memory) + barrier"
                                                                           if searcher.exchange == "p2p" else "NCCL all-gather"),
                               "l2": "operands (%.0f MB bf16 gallery shard + %.0f MB fp32 for the re-score per GPU) larger than L2" % (G / world * D * 2 / 1e6, G / world * D * 4 / 1e6)},
            "roofline": {"kernel": "sim_topk", "bound": "tensor", "achieved": round(tf, 1), "peak": peaks["tensor_sustained"], "unit": "TFLOP/s",
                         "frac": round(tf / peaks["tensor_sustained"], 4), "traffic": None, "per_gpu": True,
                         "peak_source": peaks["source"] + " (bf16_tflops_sustained); the time includes prepare + re-score + exchange + merge"},
            "e2e": {"value": round(Q * steps / (ms_e2e / 1e3), 1), "unit": "queries/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "h2d_bytes_per_step_all_ranks": int(Q * D * 4), "ms_per_step": round(ms_e2e / steps, 3),
                    "ms_per_step_per_rank": {"min": round(min(per_rank_e2e), 3), "max": round(max(per_rank_e2e), 3)},
                    "how": "each rank uploads 1/%d of the queries (%d chunk(s), upload of chunk c+1 overlapped with the search of chunk c), "
                           "%s, each rank downloads its rows of the merged top-k" % (world, chunks,
                           "replicates its block into every rank's query buffers with peer stores (cvpce_rows_prepare_scatter)"
                           if (world > 1 and searcher.exchange == "p2p") else ("all-gathers the blocks (NCCL)" if world > 1 else "no exchange"))},
            "c5_parity": parity_ok, "c5_parity_detail": parity,
            "gpu_launches": launches, "clocks": clocks}
    del searcher, q, blocks_h
    torch.cuda.empty_cache()
    return line


def _brief(l):
    """The fields of a full bench line that other_configs keeps."""
    out = {"workload": l["config"]["workload"], "value": l["value"], "unit": l["unit"], "ms_per_step": l["ms_per_step"],
           "steps": l["steps"], "warmup": l["warmup"], "dtype": l["dtype"], "e2e": l["e2e"]["value"],
           "e2e_ms_per_step": l["e2e"]["ms_per_step"], "h2d_bytes_per_step": l["e2e"]["h2d_bytes_per_step"],
           "d2h_bytes_per_step": l["e2e"]["d2h_bytes_per_step"]}
    if "ms_per_step_per_rank" in l:
        out["ms_per_step_per_rank"] = l["ms_per_step_per_rank"]
    return out


def main():
    args = parse()
    if args.impl == "reference":
        world = int(os.environ.get("WORLD_SIZE", "1"))
        rank = int(os.environ.get("RANK", "0"))
        line = run_reference(args, world, rank)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device; the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    world, rank, local = dist_setup(args)
    torch.backends.cudnn.benchmark = bool(args.cudnn_benchmark)
    if args.strict_fp32:
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
    import copy
    if args.workload == "c5":
        line = run_ours_c5(args, world, rank, local)
    else:
        line = run_ours_shelves(args, world, rank, local)
        headline_default = args.trunk == "f32" and not args.strict_fp32 and not args.plain_trunk
        if headline_default and args.workload == "c3" and not args.no_other_configs:
            other = {}
            # (1) BASELINE.json configs[4] on the driver's clock at EVERY N: the 1M x 512 gallery sharded over the ranks
            torch.cuda.empty_cache()
            a5 = copy.copy(args)
            a5.workload, a5.gallery, a5.steps, a5.warmup = "c5", None, max(3, min(args.steps, 5)), 3
            l5 = run_ours_c5(a5, world, rank, local)
            other["c5"] = {"workload": l5["config"]["workload"], "exchange": l5["config"]["exchange"], "value": l5["value"], "unit": l5["unit"],
                           "ms_per_step": l5["ms_per_step"], "ms_per_step_per_rank": l5["ms_per_step_per_rank"], "steps": l5["steps"],
                           "warmup": l5["warmup"], "scaling": "strong", "dtype": l5["dtype"],
                           "tflops_per_gpu": l5["roofline"]["achieved"], "frac_of_bf16_tflops_sustained": l5["roofline"]["frac"],
                           "e2e": l5["e2e"], "c5_parity": l5["c5_parity"], "c5_parity_detail": l5["c5_parity_detail"],
                           "gpu_launches": l5["gpu_launches"], "clocks": l5["clocks"]}
            if world == 1:
                # (2) the parity mode of the headline ("parity mode keeps fp32"): the same configuration with fp32
                # convolutions; ~10x slower (cuDNN fp32 convolutions run on the CUDA cores), so one warm-up and two steps
                torch.cuda.empty_cache()
                a32 = copy.copy(args)
                a32.no_cpu_baseline, a32.steps, a32.warmup = True, 2, 1
                tf32_state = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
                torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
                try:
                    other["strict_fp32"] = _brief(run_ours_shelves(a32, world, rank, local))
                finally:
                    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32_state
                # (3) BASELINE.json configs[1] (one shelf, ~200 boxes, 180 references): the same step at batch 1, the latency case
                torch.cuda.empty_cache()
                a2 = copy.copy(args)
                a2.workload, a2.shelves, a2.gallery, a2.no_cpu_baseline, a2.steps, a2.warmup = "c2", None, None, True, 20, 5
                other["c2"] = _brief(run_ours_shelves(a2, world, rank, local))
                # (4) BASELINE.json configs[0]: 64 given crops vs 180 references through Classifier.classify, CPU reference beside it
                torch.cuda.empty_cache()
                other["c1"] = run_c1(args, torch.device("cuda", local), with_cpu=not args.no_cpu_baseline)
            line["other_configs"] = other
        if world == 1 and headline_default and not args.no_extension_modes:
            # reported next to the headline, never instead of it: the same step with the half-precision trunk hand-off
            # (same 10-bit mantissa as the TF32 convolutions of the headline, same measured deviation)
            a16 = copy.copy(args)
            a16.trunk, a16.no_cpu_baseline, a16.steps = "f16", True, max(1, min(args.steps, 2))
            torch.cuda.empty_cache()
            l16 = run_ours_shelves(a16, world, rank, local)
            line["extension_modes"] = {"trunk_f16": {"value": l16["value"], "unit": l16["unit"], "ms_per_step": l16["ms_per_step"],
                                                     "e2e": l16["e2e"]["value"], "steps": a16.steps,
                                                     "note": "half weights/activations, fp32 accumulation; not the headline"}}
            # the north-star's R-MAC pooling (3 scales, per-region L2) instead of the reference's MAC: the same step, the pooling
            # kernel reads relu4_3 (32 x 32) and relu5_3 (16 x 16) channels-last through the persistent TMA-staged kernel
            ar = copy.copy(args)
            ar.rmac_levels, ar.no_cpu_baseline, ar.steps = 3, True, max(1, min(args.steps, 2))
            torch.cuda.empty_cache()
            lr = run_ours_shelves(ar, world, rank, local)
            line["extension_modes"]["rmac_l3"] = {"value": lr["value"], "unit": lr["unit"], "ms_per_step": lr["ms_per_step"],
                                                  "e2e": lr["e2e"]["value"], "steps": ar.steps,
                                                  "region_pool_embed": lr["kernels"].get("region_pool_embed"),
                                                  "note": "R-MAC (3 scales, 14 regions per map, per-region L2) over relu4_3 + relu5_3; not the headline "
                                                          "(the reference pools with MAC)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
