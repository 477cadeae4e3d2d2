#!/usr/bin/env python3
"""Parity + timing of the stream compaction at 2^28 (keep 80 %: rand() % 5 like the sample)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

N = 1 << 28
OFF = int(os.environ.get("COMPACT_OFFSET", "0"))  # 1: 4-byte aligned only -> register-staged fallback kernel
src = torch.empty(N + 4, dtype=torch.int32, device="cuda")[OFF:OFF + N]
dst = torch.empty(N, dtype=torch.int32, device="cuda")
cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
for mod in (5, 2, 1):
    prim.generate(src, "key_mod", 0, N, 7, mod)
    prim.filter_arr(dst, cnt, src, N)
    kept = int(cnt.item())
    want = src[src > 0]
    ok = kept == want.numel() and bool((dst[:kept] == want).all())
    for _ in range(3):
        prim.filter_arr(dst, cnt, src, N)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        prim.filter_arr(dst, cnt, src, N)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / 20 * 1e3
    print("compaction 2^28 keep %.0f%%: %s  %.1f us  %.0f GB/s (4 B/elem + 4 B/kept)" %
          (100.0 * kept / N, "OK" if ok else "MISMATCH", us, (4 * N + 4 * kept) / us / 1e3))
