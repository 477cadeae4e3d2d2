#!/usr/bin/env python3
"""Fixed cost of one histogram256 / reduce<int> call (back-to-back launches, CUDA events) against the input size."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
from cuda_samples_b200._lib import lib  # noqa: E402


def timeit(fn, iters=200, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


n = 64 << 20
bufs = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(4)]
for i, b in enumerate(bufs):
    prim.generate(b, "bytes", i * n, n, 2009)
h = torch.empty(256, dtype=torch.int32, device="cuda")
red = torch.empty(4, dtype=torch.int32, device="cuda")
it = [0]
for cl in (1, 2):
    lib.b200_tune_hist_cluster(cl)
    for nb in (16 << 10, 1 << 20, 4 << 20, 16 << 20, 32 << 20, 64 << 20):
        def step():
            prim.histogram256(h, bufs[it[0] & 3], nb)
            it[0] += 1
        us = timeit(step)
        print("hist256 cluster=%d %6d KiB: %7.2f us  (stream time at 6.5 TB/s: %5.2f us)" % (cl, nb >> 10, us, nb / 6.5e6))
lib.b200_tune_hist_cluster(0)
for nb in (16 << 10, 1 << 20, 16 << 20, 64 << 20):
    def step_r():
        prim.reduce(nb // 4, 256, 4, 7, bufs[it[0] & 3].view(torch.int32), red)
        it[0] += 1
    us = timeit(step_r)
    print("reduce<int>       %6d KiB: %7.2f us" % (nb >> 10, us))
