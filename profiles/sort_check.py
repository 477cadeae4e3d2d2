#!/usr/bin/env python3
"""Correctness + timing of the KV sort for one build of the library (B200PRIM_LIB selects a variant).
Checks stability bit-exactly against torch.sort(stable=True) on ragged and full sizes, both directions,
then times 2^28 pairs (CUDA events).   python profiles/sort_check.py [tag]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

if os.environ.get("SORT_LB"):  # look-back rows fetched per round trip (1, 4 or 8)
    prim.lib.b200_tune_sort_lookback(int(os.environ["SORT_LB"]))
if os.environ.get("SORT_PAIRS"):  # 0: intermediate results in the caller's buf arrays, 1: library pair buffers
    prim.lib.b200_tune_sort_pairs(int(os.environ["SORT_PAIRS"]))
tag = sys.argv[1] if len(sys.argv) > 1 else os.environ.get("B200PRIM_LIB", "default")
ok = True
for n in (() if os.environ.get("SORT_CHECK_SKIP_PARITY") else (1, 31, 4096, 4097, 12345, (1 << 20) + 77, 1 << 24)):
    for kind, param in (("key_mod", 300), ("key_unique", 0), ("u32", 0)):
        sk = torch.empty(n, dtype=torch.int32, device="cuda")
        if kind == "u32":  # uniform random 32-bit keys: four random bytes each
            prim.generate(sk.view(torch.uint8), "bytes", 0, 4 * n, 7 + n)
        else:
            prim.generate(sk, kind, 0, n, 7 + n, param)
        sv = torch.arange(n, dtype=torch.int32, device="cuda")
        dk, dv, bk, bv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4))
        for direction in (1, 0):
            prim.mergeSort(dk, dv, bk, bv, sk, sv, n, direction)
            torch.cuda.synchronize()
            ku = sk.to(torch.int64) & 0xffffffff
            rk, ri = torch.sort(ku, stable=True, descending=(direction == 0))
            good = bool((dk.to(torch.int64) & 0xffffffff).eq(rk).all()) and bool(dv.to(torch.int64).eq(ri).all())
            if not good:
                ok = False
                print("[%s] MISMATCH n=%d kind=%s dir=%d" % (tag, n, kind, direction))
print("[%s] parity %s" % (tag, "OK" if ok else "FAILED"))

N = 1 << 28
sk = torch.empty(N, dtype=torch.int32, device="cuda")
sv = torch.arange(N, dtype=torch.int32, device="cuda")
dk, dv, bk, bv = (torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(4))
for kind, param in (("key_unique", 0), ("u32", 0)):
    if kind == "u32":
        prim.generate(sk.view(torch.uint8), "bytes", 0, 4 * N, 2009)
    else:
        prim.generate(sk, kind, 0, N, 2009, param)
    for _ in range(2):
        prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) / 5 * 1e3
    print("[%s] sort 2^28 %-10s %8.1f us  %6.2f Gkeys/s" % (tag, kind, us, N / us / 1e3))
