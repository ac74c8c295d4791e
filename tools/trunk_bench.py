#!/usr/bin/env python
"""How fast can the (not rebuilt) torch VGG16 trunk be driven? Development probe, not part of the product."""
import sys, os, time
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvpce_b200 import classification

dev = torch.device("cuda")
torch.manual_seed(0)
enc = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
x = torch.randn(B, 3, 256, 256, device=dev)


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n


def fused_trunk(layers, x):
    i = 0
    layers = list(layers)
    while i < len(layers):
        l = layers[i]
        if isinstance(l, torch.nn.Conv2d) and i + 1 < len(layers) and isinstance(layers[i + 1], torch.nn.ReLU):
            x = torch.cudnn_convolution_relu(x, l.weight, l.bias, l.stride, l.padding, l.dilation, l.groups)
            i += 2
        elif isinstance(l, torch.nn.MaxPool2d):
            x = F.max_pool2d(x, l.kernel_size, l.stride)
            i += 1
        else:
            x = l(x)
            i += 1
    return x


with torch.no_grad():
    ref4, ref5 = enc.features(x, normalized=True)
    for bench in (False, True):
        torch.backends.cudnn.benchmark = bench
        t = timeit(lambda: enc.features(x, normalized=True))
        print("nchw f32 benchmark=%s: %.1f crops/s" % (bench, B / t), flush=True)
    xc = x.contiguous(memory_format=torch.channels_last)
    encc = enc.to(memory_format=torch.channels_last)
    for bench in (False, True):
        torch.backends.cudnn.benchmark = bench
        t = timeit(lambda: encc.features(xc, normalized=True))
        print("nhwc f32 benchmark=%s: %.1f crops/s" % (bench, B / t), flush=True)
    f4, f5 = encc.features(xc, normalized=True)
    print("  nhwc vs nchw max rel diff f5: %.3e" % float((f5 - ref5).abs().max() / ref5.abs().max()))
    for bench in (False, True):
        torch.backends.cudnn.benchmark = bench
        try:
            def run():
                a = fused_trunk(encc.block1, xc)
                return a, fused_trunk(encc.block2, a)
            t = timeit(run)
            print("nhwc f32 fused conv+bias+relu benchmark=%s: %.1f crops/s" % (bench, B / t), flush=True)
            f4, f5 = run()
            print("  fused vs nchw max rel diff f5: %.3e  channels_last out: %s" % (
                float((f5 - ref5).abs().max() / ref5.abs().max()), f5.is_contiguous(memory_format=torch.channels_last)))
        except Exception as e:
            print("fused failed:", repr(e)[:300])
    torch.backends.cudnn.benchmark = False
    try:
        def run_nchw():
            a = fused_trunk(enc.block1, x)
            return a, fused_trunk(enc.block2, a)
        enc2 = classification.macvgg_embedder("vgg16", pretrained=False).eval().requires_grad_(False).to(dev)
        enc2.load_state_dict(enc.state_dict())
        t = timeit(lambda: (fused_trunk(enc2.block1, x)))
        print("nchw f32 fused block1 only: %.1f crops/s" % (B / t))
    except Exception as e:
        print("nchw fused failed:", repr(e)[:300])
    torch.backends.cudnn.allow_tf32 = False
    t = timeit(lambda: encc.features(xc, normalized=True))
    print("nhwc strict-fp32 (no tf32): %.1f crops/s" % (B / t), flush=True)
    torch.backends.cudnn.allow_tf32 = True
    e16 = encc.to(torch.bfloat16)
    x16 = xc.to(torch.bfloat16)
    torch.backends.cudnn.benchmark = True
    t = timeit(lambda: e16.features(x16, normalized=True))
    print("nhwc bf16 benchmark=True: %.1f crops/s" % (B / t), flush=True)
