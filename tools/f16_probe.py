#!/usr/bin/env python
"""Probe (not a product path): how fast and how accurate is the torch VGG16 trunk in fp16 (10-bit mantissa like TF32, fp32
accumulation, half the activation bytes) next to the TF32 trunk bench.py times?  Prints crops/s of the trunk alone and the
embedding deviation of both against strict fp32."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import classification, synth

dev = torch.device("cuda")
torch.manual_seed(synth.WEIGHT_SEED)
enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(dev).enable_fast_trunk()
g = torch.Generator(device=dev).manual_seed(11)
x = (torch.rand((256, 3, 256, 256), device=dev, generator=g) * 2 - 1).contiguous(memory_format=torch.channels_last)


def run(e, xx):
    f4, f5 = e.features(xx, normalized=True)
    d = torch.cat([f4.float().amax(dim=(2, 3)), f5.float().amax(dim=(2, 3))], 1)
    return d / d.norm(dim=1, keepdim=True).clamp(min=1e-8)


def rate(e, xx, n=5):
    for _ in range(2):
        run(e, xx)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        run(e, xx)
    torch.cuda.synchronize()
    return xx.shape[0] * n / (time.perf_counter() - t0)


with torch.no_grad():
    torch.backends.cudnn.allow_tf32 = False
    ref = run(enc, x[:64])
    torch.backends.cudnn.allow_tf32 = True
    tf = run(enc, x[:64])
    r_tf = rate(enc, x)
    e16 = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    e16.load_state_dict(enc.state_dict())
    e16 = e16.to(dev).half().enable_fast_trunk()
    x16 = x.half()
    h = run(e16, x16[:64])
    r_h = rate(e16, x16)
for name, e in (("tf32", tf), ("fp16", h)):
    d = (e - ref)
    print("%s: ||d||/||e|| max %.3e, max abs %.3e, min cosine %.8f" % (
        name, (d.norm(dim=1) / ref.norm(dim=1)).max().item(), d.abs().max().item(),
        torch.nn.functional.cosine_similarity(e, ref).min().item()))
print("trunk alone (256-crop batches, torch max over the maps): tf32 %.0f crops/s, fp16 %.0f crops/s" % (r_tf, r_h))
