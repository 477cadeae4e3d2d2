#!/usr/bin/env python3
"""histogram256 on 64 MiB: per-call time in a Python loop vs inside a CUDA graph of 32 calls (4 rotating inputs)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

n = int(os.environ.get("HIST_MIB", "64")) << 20
bufs = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(4)]
for i, b in enumerate(bufs):
    prim.generate(b, "bytes", i * n, n, 2009)
h = torch.empty(256, dtype=torch.int32, device="cuda")
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for k in range(8):
        prim.histogram256(h, bufs[k & 3], n)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for k in range(32):
            prim.histogram256(h, bufs[k & 3], n)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    print("graph of 32 calls: %.2f us per call" % (a.elapsed_time(b) / 320 * 1e3))
    a.record()
    for k in range(320):
        prim.histogram256(h, bufs[k & 3], n)
    b.record()
    torch.cuda.synchronize()
    print("python loop      : %.2f us per call" % (a.elapsed_time(b) / 320 * 1e3))
