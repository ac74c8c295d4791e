#!/usr/bin/env python
"""Development aid: per-phase cycle counts of nms_batched_kernel (block 0).  Needs a library built with
CVPCE_NVCC_EXTRA=-DCVPCE_PHASE_TIMERS python -m cvpce_b200.build --force"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import _lib, ops, synth, detection

lib = _lib.load()
buf = (ctypes.c_longlong * 8)()


def phases(tag):
    lib.cvpce_debug_nms_phases(buf, 1)
    v = list(buf)
    names = ["sort", "gather+range", "A grid test", "B mask", "C resolve", "panels"]
    print(tag, {n: x for n, x in zip(names, v)}, "total cycles", sum(v[:5]))


dev = "cuda"
lib.cvpce_debug_nms_phases(buf, 1)
b, s = synth.random_boxes(50, 5000)
bd, sd = torch.from_numpy(b).to(dev), torch.from_numpy(s).to(dev)
seg = torch.tensor([0, 5000], dtype=torch.int32, device=dev)
ops.nms_batched(bd, sd, seg, 5000, 0.5, None, 1000)
phases("stress 5000 (max_keep 1000)")
ops.nms_batched(bd, sd, seg, 5000, 0.5, None, 0)
phases("stress 5000 (no cap)")
bb, ss = synth.shelf_candidates(100, *synth.SHELF_HW, 200)
seg = torch.tensor([0, len(ss)], dtype=torch.int32, device=dev)
ops.nms_batched(torch.from_numpy(bb).to(dev), torch.from_numpy(ss).to(dev), seg, len(ss), 0.5, None, 200)
phases("c3 shelf %d" % len(ss))
H, W = 800, 1066
anchors = [torch.from_numpy(a).to(dev) for a in synth.fpn_anchors(H, W)]
n_per = [a.shape[0] for a in anchors]
g = torch.Generator(device=dev).manual_seed(5)
lg = -4.6 + 1.5 * torch.randn((1, sum(n_per), 1), device=dev, generator=g)
rg = 0.3 * torch.randn((1, sum(n_per), 4), device=dev, generator=g)
head = {"cls_logits": list(lg.split(n_per, 1)), "bbox_regression": list(rg.split(n_per, 1))}
sizes = torch.tensor([[H, W]], dtype=torch.int32, device=dev)
detection.postprocess_detections_batch(head, anchors, sizes, 0.05, 1000, 0.5, 1000)
phases("gln trained-like")
