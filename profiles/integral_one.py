#!/usr/bin/env python3
"""One integral image call per size (for ncu launch lists): II_SHAPE=h,w"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

h, w = (int(v) for v in os.environ.get("II_SHAPE", "16384,16384").split(","))
img = torch.randint(0, 256, (h, w), dtype=torch.uint8, device="cuda")
out = torch.empty(h * w, dtype=torch.int32, device="cuda")
for _ in range(2):
    prim.integral_image(out, img, w, h)
torch.cuda.synchronize()
