#!/usr/bin/env python
"""Development aid: where block 0 of rmac_staged_kernel spends its cycles (producer thread; lane 0 of consumer warp 0).
Needs a library variant built with -DCVPCE_RMAC_PROF:
    python tools/build_variant.py rmacprof rmac_staged.cu -DCVPCE_RMAC_PROF
    CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so python tools/rmac_prof.py"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import _lib, ops

lib = _lib.load()
buf = (ctypes.c_longlong * 16)()
NAMES = ["producer: waiting for a free stage", "producer: total", "consumer: waiting for a full stage", "consumer: region updates",
         "consumer: per-map finish (norms, descriptor)", "consumer: total", "consumer: barrier after the walk (waiting for the slowest warp)",
         "consumer: per-image finish (final norm, store)", "NCHW warp 0: column passes", "NCHW warp 0: cell maxima", "NCHW warp 0: region maxima",
         "NCHW warp 0: plane groups processed"]


def run(tag, feats, regs, **knobs):
    out = torch.empty((feats[0].shape[0], sum(f.shape[1] for f in feats)), device="cuda")
    with _lib.tuning(rmac_kernel=2, **knobs):
        ops.region_pool_embed(feats, regions=regs, per_region_l2=True, out=out)
        lib.cvpce_debug_rmac_prof(buf, 1)
        ops.region_pool_embed(feats, regions=regs, per_region_l2=True, out=out)
        lib.cvpce_debug_rmac_prof(buf, 1)
    v = list(buf)
    print(tag)
    for n, x in zip(NAMES, v):
        print("   %-70s %9d cycles" % (n, x))


for B in (1, 128):
    for cl in (True, False):
        for dt in (torch.float32, torch.bfloat16):
            f4 = torch.rand((B, 512, 32, 32), device="cuda").to(dt)
            f5 = torch.rand((B, 512, 16, 16), device="cuda").to(dt)
            if cl:
                f4, f5 = f4.contiguous(memory_format=torch.channels_last), f5.contiguous(memory_format=torch.channels_last)
            regs = [ops.rmac_regions(32, 32, 3), ops.rmac_regions(16, 16, 3)]
            run("B=%d %s %s" % (B, "nhwc" if cl else "nchw", dt), [f4, f5], regs)
