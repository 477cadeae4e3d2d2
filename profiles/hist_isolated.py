#!/usr/bin/env python3
"""histogram256 on 64 MiB as ONE call between device synchronisations (what a caller who histograms once gets), per
CTA geometry for short inputs (b200_tune_hist_launch: 1024 threads / one 512-thread CTA per SM / two per SM), against
the same geometry called back to back.  CUDA events around the single call: launch latency is inside."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

n = 64 << 20
bufs = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(4)]
for i, b in enumerate(bufs):
    prim.generate(b, "bytes", i * n, n, 2009)
h = torch.empty(256, dtype=torch.int32, device="cuda")
for small in (0, 1, 2):
    prim.lib.b200_tune_hist_launch(1, small)
    iso = []
    for i in range(44):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        prim.histogram256(h, bufs[i & 3], n)
        b.record()
        torch.cuda.synchronize()
        if i >= 4:
            iso.append(a.elapsed_time(b) * 1e3)
    for _ in range(8):
        prim.histogram256(h, bufs[0], n)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(40):
        prim.histogram256(h, bufs[i & 3], n)
    b.record()
    torch.cuda.synchronize()
    print("short-input geometry %d: isolated median %.2f us (min %.2f), back to back %.2f us per call" %
          (small, float(np.median(iso)), min(iso), a.elapsed_time(b) * 1e3 / 40), flush=True)
prim.lib.b200_tune_hist_launch(1, 1)
