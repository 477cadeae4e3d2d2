#!/usr/bin/env python3
"""Batched short-array key-value sort (the sortingNetworks API: bitonicSort / oddEvenMergeSort) -- parity against a
per-array stable sort and microseconds per call, for the shapes of the sample (N = 2^20 pairs, keys = rand() % 65536)
and larger batches, 32-bit random keys included.  16 B of traffic per pair (keys + values in, keys + values out)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

peak = 6458.1
for batch, L, kind, param in ((16384, 64, "key_mod", 65536), (4096, 256, "key_mod", 65536), (1024, 1024, "key_mod", 65536),
                              (256, 4096, "key_mod", 65536), (1 << 20, 64, "key_mod", 65536), (1 << 14, 4096, "key_mod", 65536),
                              (1 << 14, 4096, "u32", 0), (1 << 20, 64, "u32", 0), (30000, 1000, "key_mod", 65536),
                              (128, 8192, "key_mod", 65536)):
    n = batch * L
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    if kind == "u32":
        prim.generate(sk.view(torch.uint8), "bytes", 0, 4 * n, 5)
    else:
        prim.generate(sk, kind, 0, n, 5, param)
    sv = torch.arange(n, dtype=torch.int32, device="cuda")
    dk, dv = torch.empty_like(sk), torch.empty_like(sv)
    ok = True
    for direction in (1, 0):
        prim.bitonicSort(dk, dv, sk, sv, batch, L, direction)
        ku = (sk.to(torch.int64) & 0xffffffff).view(batch, L)
        rk, ri = torch.sort(ku, dim=1, stable=True, descending=(direction == 0))
        want_v = (ri + (torch.arange(batch, device="cuda") * L).view(batch, 1)).view(-1)
        ok &= bool((dk.to(torch.int64) & 0xffffffff).eq(rk.view(-1)).all()) and bool(dv.to(torch.int64).eq(want_v).all())
        del ku, rk, ri, want_v
    for _ in range(3):
        prim.bitonicSort(dk, dv, sk, sv, batch, L, 1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    a.record()
    for _ in range(reps):
        prim.bitonicSort(dk, dv, sk, sv, batch, L, 1)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / reps
    print("%8d x %5d  %-18s parity %-6s %9.1f us  %7.2f Gpairs/s  %7.1f GB/s at 16 B/pair = %.3f of measured peak"
          % (batch, L, kind + ("%%%d" % param if param else ""), "OK" if ok else "FAILED", us, n / us / 1e3,
             16 * n / us / 1e3, 16 * n / us / 1e3 / peak), flush=True)
    del sk, sv, dk, dv
