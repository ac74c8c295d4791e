#!/usr/bin/env python
"""Probe: does cuDNN run VGG16's first convolution faster on 4-channel channels-last input (zero 4th channel, zero-padded
weight) than on the 3-channel input, which needs nhwcAddPaddingKernel?  Development probe, not part of the product."""
import torch

torch.manual_seed(0)
dev = "cuda"
B = 256
w = torch.randn(64, 3, 3, 3, device=dev) * 0.1
b = torch.randn(64, device=dev) * 0.1
x3 = torch.randn(B, 3, 256, 256, device=dev).contiguous(memory_format=torch.channels_last)
x4 = torch.zeros(B, 4, 256, 256, device=dev).contiguous(memory_format=torch.channels_last)
x4[:, :3] = x3
w3 = w.contiguous(memory_format=torch.channels_last)
w4 = torch.zeros(64, 4, 3, 3, device=dev)
w4[:, :3] = w
w4 = w4.contiguous(memory_format=torch.channels_last)


def run(x, ww):
    return torch.cudnn_convolution_relu(x, ww, b, (1, 1), (1, 1), (1, 1), 1)


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


y3, y4 = run(x3, w3), run(x4, w4)
print("max |y4 - y3| = %.3e (max |y| %.3f)" % (float((y4 - y3).abs().max()), float(y3.abs().max())))
print("conv1_1 on 3 channels: %.3f ms   on 4 channels: %.3f ms   (B = %d)" % (timeit(lambda: run(x3, w3)), timeit(lambda: run(x4, w4)), B))
