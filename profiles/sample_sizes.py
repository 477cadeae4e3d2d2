#!/usr/bin/env python3
"""The primitives at the sizes the samples' own drivers use (reduction: 2^24 elements; scan: 6.8 M elements as arrays
of 4 .. 262144; mergeSort: 4 Mi pairs; histogram: 64 MiB), CUDA events, device-resident.   python profiles/sample_sizes.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402


def t(fn, it=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it * 1e3


# rotate over several buffers so that the inputs do not simply sit in the 126 MB L2
R = 8
n = 1 << 24
xi = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(R)]
for x in xi:
    prim.generate(x, "i32_byte", 0, n, 1)
oi = torch.empty(64, dtype=torch.int32, device="cuda")
k = [0]


def red():
    k[0] = (k[0] + 1) % R
    prim.reduce(n, 256, 64, 7, xi[k[0]], oi)


us = t(red)
print("reduce<int> 2^24 (rotating over %d inputs): %6.2f us  %6.0f GB/s" % (R, us, 4.0 * n / us / 1e3))
xf = [x.view(torch.float32) for x in xi]
of = torch.empty(64, dtype=torch.float32, device="cuda")


def redf():
    k[0] = (k[0] + 1) % R
    prim.reduce(n, 256, 64, 7, xf[k[0]], of)


us = t(redf)
print("reduce<float> 2^24: %6.2f us  %6.0f GB/s" % (us, 4.0 * n / us / 1e3))
N = 13 * 1048576 // 2
dst = [torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(R)]
for L in (4, 64, 1024, 2048, 65536, 262144):
    def sc():
        k[0] = (k[0] + 1) % R
        if L <= 1024:
            prim.scanExclusiveShort(dst[k[0]], xi[k[0]], N // L, L)
        else:
            prim.scanExclusiveLarge(dst[k[0]], xi[k[0]], N // L, L)
    us = t(sc)
    print("scan %d arrays of %6d: %6.2f us  %6.0f GB/s" % (N // L, L, us, 8.0 * N / us / 1e3))
m = 4 << 20
sk = [torch.empty(m, dtype=torch.int32, device="cuda") for _ in range(4)]
for s in sk:
    prim.generate(s, "key_mod", 0, m, 2009, 65536)
sv = torch.arange(m, dtype=torch.int32, device="cuda")
dk, dv, bk, bv = (torch.empty(m, dtype=torch.int32, device="cuda") for _ in range(4))
prim.initMergeSort()


def so():
    k[0] = (k[0] + 1) % 4
    prim.mergeSort(dk, dv, bk, bv, sk[k[0]], sv, m, 1)


us = t(so, 100)
print("mergeSort 4 Mi pairs, keys %% 65536: %6.2f us  %6.2f Gkeys/s" % (us, m / us / 1e3))
