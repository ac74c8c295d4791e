#!/usr/bin/env python
"""How far do the embeddings move when the (torch) trunk's convolutions run in TF32 (torch's default on GPU, what bench.py
times) instead of strict fp32?  Prints max / mean relative deviation, cosine, and top-1 / top-5 agreement vs a gallery."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import classification, synth

dev = torch.device("cuda")
torch.manual_seed(synth.WEIGHT_SEED)
enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(dev).enable_fast_trunk()
g = torch.Generator(device=dev).manual_seed(11)
x = torch.rand((128, 3, 256, 256), device=dev, generator=g) * 2 - 1
gal = torch.rand((8192, 3, 64, 64), device=dev, generator=g)
out = {}
with torch.no_grad():
    for tf32 in (False, True):
        torch.backends.cudnn.allow_tf32 = tf32
        out[tf32] = enc(x)
        if not tf32:
            # gallery: embeddings of other random crops (strict fp32), so that neighbours are as close as random-init allows
            G = torch.cat([enc(torch.nn.functional.interpolate(gal[i:i + 256], size=(256, 256))) for i in range(0, 2048, 256)])
    # --trunk f16: half weights / activations (10-bit mantissa like TF32), fp32 accumulation, library kernels reading half
    e16 = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False)
    e16.load_state_dict(enc.state_dict())
    e16 = e16.to(dev).half().enable_fast_trunk()
    with torch.no_grad():
        out["f16"] = e16(x.half())
a, b = out[False], out[True]
rel = ((a - b).abs() / a.abs().clamp(min=1e-6))
print("embedding tf32 vs fp32: max rel %.3e  mean rel %.3e  max abs %.3e  min cosine %.8f" % (
    rel.max().item(), rel.mean().item(), (a - b).abs().max().item(), torch.nn.functional.cosine_similarity(a, b).min().item()))
sa, ia = classification.similarity_topk(G, a, 5)
for name, b in (("tf32", out[True]), ("f16", out["f16"])):
    rel = ((a - b).abs() / a.abs().clamp(min=1e-6))
    sb, ib = classification.similarity_topk(G, b, 5)
    print("%s vs fp32: ||d||/||e|| max %.3e, mean elementwise rel %.3e, max abs %.3e, min cosine %.8f | similarity max rel diff %.3e, "
          "top-1 agreement %.4f, top-5 list agreement %.4f" % (
              name, ((a - b).norm(dim=1) / a.norm(dim=1)).max().item(), rel.mean().item(), (a - b).abs().max().item(),
              torch.nn.functional.cosine_similarity(a, b).min().item(), ((sa - sb).abs() / sa.abs()).max().item(),
              (ia[:, 0] == ib[:, 0]).float().mean().item(), (ia == ib).all(dim=1).float().mean().item()))
