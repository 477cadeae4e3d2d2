#!/usr/bin/env python3
"""Knob sweeps on one B200 (CUDA-event timing, device-resident 2^28-element inputs).
    python profiles/experiments.py [scan] [sort] [hist]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
from cuda_samples_b200._lib import lib  # noqa: E402

N = 1 << 28
what = sys.argv[1:] or ["scan", "sort", "hist"]


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3  # us


if "scan" in what:
    src = torch.empty(N, dtype=torch.int32, device="cuda")
    dst = torch.empty(N, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", 0, N, 2009)
    for tma, back in ((0, 0), (1, 0), (1, 100), (1, 200), (1, 400), (1, 800)):
        lib.b200_tune_scan_tma(tma)
        lib.b200_tune_scan_backoff(back)
        us = timeit(lambda: prim.scanExclusiveLarge(dst, src, 1, N), 20)
        print("scan 2^28  tma=%d backoff=%4d ns: %8.1f us  %7.1f GB/s" % (tma, back, us, 8 * N / us / 1e3))
    lib.b200_tune_scan_tma(1)
    lib.b200_tune_scan_backoff(0)
    for batch, L in ((1024, 1 << 18), (1 << 14, 1 << 14), (4, 1 << 26)):
        us = timeit(lambda: prim.scanExclusiveLarge(dst, src, batch, L), 20)
        print("scan batch=%d L=%d: %8.1f us  %7.1f GB/s" % (batch, L, us, 8 * N / us / 1e3))
    del src, dst

if "sort" in what:
    sk = torch.empty(N, dtype=torch.int32, device="cuda")
    sv = torch.arange(N, dtype=torch.int32, device="cuda")
    dk, dv, bk, bv = (torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(4))
    for kind, param in (("key_unique", 0), ("key_mod", 65536)):
        prim.generate(sk, kind, 0, N, 2009, param)
        for lb in (1, 4, 8):
            lib.b200_tune_sort_lookback(lb)
            us = timeit(lambda: prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1), 5, 2)
            print("sort 2^28 %-10s lookback=%d per round: %8.1f us  %6.2f Gkeys/s" % (kind, lb, us, N / us / 1e3))
    lib.b200_tune_sort_lookback(4)
    del sk, sv, dk, dv, bk, bv

if "hist" in what:
    n = 64 << 20
    bufs = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(4)]
    for i, b in enumerate(bufs):
        prim.generate(b, "bytes", i * n, n, 2009)
    h = torch.empty(256, dtype=torch.int32, device="cuda")
    it = [0]

    def step():
        prim.histogram256(h, bufs[it[0] & 3], n)
        it[0] += 1
    for cl in (1, 2, 4, 8):
        lib.b200_tune_hist_cluster(cl)
        us = timeit(step, 40, 8)
        print("hist256 64 MiB cluster=%d: %7.2f us  %7.1f GB/s" % (cl, us, n / us / 1e3))
    lib.b200_tune_hist_cluster(0)
    del bufs
    big = torch.empty(4 << 30, dtype=torch.uint8, device="cuda")
    prim.generate(big, "bytes", 0, 4 << 30, 2009)
    for nb in (256 << 20, 1 << 30, 4 << 30):
        us = timeit(lambda: prim.histogram256(h, big, nb), 10, 3)
        print("hist256 %5d MiB: %8.2f us  %7.1f GB/s" % (nb >> 20, us, nb / us / 1e3))
