#!/usr/bin/env python3
"""Parity (torch double cumsum) + timing of the integral image at the sample's size and at HBM-sized images.
B200_II_ROWS=8/16/32 forces the band height."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

for h, w in ((1080, 1920), (4096, 4096), (16384, 16384), (32768, 16384), (8190, 16380)):
    img = torch.randint(0, 256, (h, w), dtype=torch.uint8, device="cuda")
    out = torch.empty(h * w, dtype=torch.int32, device="cuda")
    prim.integral_image(out, img, w, h)
    want = img.to(torch.int64).cumsum(1).cumsum(0).to(torch.int32).reshape(-1) if h * w <= (1 << 28) else None
    ok = "unchecked" if want is None else ("OK" if bool((out == want).all()) else "MISMATCH")
    del want
    reps = 200 if h * w < (1 << 24) else 20
    for _ in range(3):
        prim.integral_image(out, img, w, h)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        prim.integral_image(out, img, w, h)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / reps * 1e3
    print("integral image %5d x %5d: %s  %9.1f us  %6.0f GB/s (5 B/pixel)  %.2f Gpixel/s" %
          (w, h, ok, us, 5.0 * h * w / us / 1e3, h * w / us / 1e3), flush=True)
    del img, out
