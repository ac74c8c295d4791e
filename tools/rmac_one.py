#!/usr/bin/env python
"""One R-MAC shape through ops.region_pool_embed a few times (for ncu captures of rmac_staged_kernel):
    python tools/rmac_one.py [B=256] [nhwc|nchw] [f32|bf16]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import _lib, ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cl = (sys.argv[2] if len(sys.argv) > 2 else "nhwc") == "nhwc"
dt = torch.bfloat16 if (len(sys.argv) > 3 and sys.argv[3] == "bf16") else torch.float32
f4 = torch.rand((B, 512, 32, 32), device="cuda").to(dt)
f5 = torch.rand((B, 512, 16, 16), device="cuda").to(dt)
if cl:
    f4, f5 = f4.contiguous(memory_format=torch.channels_last), f5.contiguous(memory_format=torch.channels_last)
regs = [ops.rmac_regions(32, 32, 3), ops.rmac_regions(16, 16, 3)]
out = torch.empty((B, 1024), device="cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
for i in range(5):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.region_pool_embed([f4, f5], regions=regs, per_region_l2=True, out=out)
    e1.record()
    torch.cuda.synchronize()
    print("call %d: %.4f ms" % (i, e0.elapsed_time(e1)), flush=True)
print("algorithmic bytes per launch:", f4.element_size() * (f4.numel() + f5.numel()) + 4 * 1024 * B)
