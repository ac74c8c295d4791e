#!/usr/bin/env python
"""Development aid: where block 0 of the staged R-MAC kernels spends its cycles (clock64 counters of a -DCVPCE_RMAC_PROF build):
    python tools/build_variant.py rmacprof rmac_staged.cu -DCVPCE_RMAC_PROF
    CVPCE_B200_SO=cvpce_b200/variants/librmacprof.so python tools/rmac_prof.py"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import _lib, ops

lib = _lib.load()
buf = (ctypes.c_longlong * 16)()
NAMES = ["producer: waiting for a free stage", "producer: total", "consumer thread 0: waiting for a full stage",
         "consumer thread 0: row-segment ends (plain kernel: incl. the region stage)", "plain kernel: per-map finish (norms, descriptor)",
         "plain kernel: consumers total", "plain kernel: barrier after a map", "plain kernel: per-image finish (final norm, store)",
         "-", "-", "-", "-", "ws kernel: prologue (parameter copy, barrier init)", "ws kernel: entry -> streaming warps done",
         "ws kernel: entry -> region warps done", "ws kernel: wait for the very first chunk"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def run(tag, feats, regs, kernel):
    out = torch.empty((feats[0].shape[0], sum(f.shape[1] for f in feats)), device="cuda")
    with _lib.tuning(rmac_kernel=kernel):
        for _ in range(2):
            ops.region_pool_embed(feats, regions=regs, per_region_l2=True, out=out)
        lib.cvpce_debug_rmac_prof(buf, 1)
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.region_pool_embed(feats, regions=regs, per_region_l2=True, out=out)
        e1.record()
        torch.cuda.synchronize()
        lib.cvpce_debug_rmac_prof(buf, 1)
    print("%s: %.1f us (events around the call, L2 flushed)" % (tag, 1e3 * e0.elapsed_time(e1)))
    for n, x in zip(NAMES, list(buf)):
        if n != "-" and x:
            print("   %-75s %9d cycles" % (n, x))


for B in (64, 128, 256, 512):
    for kernel in (2, 4):
        f4 = torch.rand((B, 512, 32, 32), device="cuda").contiguous(memory_format=torch.channels_last)
        f5 = torch.rand((B, 512, 16, 16), device="cuda").contiguous(memory_format=torch.channels_last)
        regs = [ops.rmac_regions(32, 32, 3), ops.rmac_regions(16, 16, 3)]
        run("B=%d channels-last f32, kernel %s" % (B, {2: "rmac_staged_kernel", 4: "rmac_staged_ws_kernel"}[kernel]), [f4, f5], regs, kernel)
        del f4, f5
