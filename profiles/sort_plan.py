#!/usr/bin/env python3
"""Device-side pass plan of the KV sort (passes whose digit is the same for every key are skipped): stability
bit-exact against torch.sort(stable=True) on key sets whose varying bytes sit in every position, both directions,
with the plan on and off; then timing at 2^28 and at the sample's own 4 Mi pairs.   python profiles/sort_plan.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402


def keys(n, kind, seed):
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    if kind == "u32":
        prim.generate(sk.view(torch.uint8), "bytes", 0, 4 * n, seed)
        return sk
    prim.generate(sk, "key_mod", 0, n, seed, {"mod65536": 65536, "mod256": 256, "const": 1, "hi16": 65536,
                                              "byte1": 256, "byte3": 256, "bytes0+2": 65536}[kind])
    if kind == "hi16":
        sk <<= 16
    elif kind == "byte1":
        sk <<= 8
    elif kind == "byte3":
        sk <<= 24
    elif kind == "bytes0+2":
        sk.copy_((sk & 0xff) | ((sk & 0xff00) << 8))
    elif kind == "const":
        sk += 0x12345678
    return sk


KINDS = ("u32", "mod65536", "mod256", "const", "hi16", "byte1", "byte3", "bytes0+2")
for plan in (1, 0):
    prim.lib.b200_tune_sort_plan(plan)
    ok = True
    for n in (65536, (1 << 20) + 77, 1 << 22):
        for kind in KINDS:
            sk = keys(n, kind, 7 + n)
            sv = torch.arange(n, dtype=torch.int32, device="cuda")
            dk, dv, bk, bv = (torch.full((n,), -1, dtype=torch.int32, device="cuda") for _ in range(4))
            for direction in (1, 0):
                dk.fill_(-1), dv.fill_(-1)
                prim.mergeSort(dk, dv, bk, bv, sk, sv, n, direction)
                torch.cuda.synchronize()
                ku = sk.to(torch.int64) & 0xffffffff
                rk, ri = torch.sort(ku, stable=True, descending=(direction == 0))
                good = bool((dk.to(torch.int64) & 0xffffffff).eq(rk).all()) and bool(dv.to(torch.int64).eq(ri).all())
                if not good:
                    ok = False
                    print("plan=%d MISMATCH n=%d kind=%s dir=%d" % (plan, n, kind, direction))
    print("plan=%d parity %s" % (plan, "OK" if ok else "FAILED"), flush=True)

for N in (1 << 28, 1 << 22):
    sv = torch.arange(N, dtype=torch.int32, device="cuda")
    dk, dv, bk, bv = (torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(4))
    for kind in ("u32", "mod65536", "mod256", "const"):
        sk = keys(N, kind, 2009)
        for plan in (0, 1):
            prim.lib.b200_tune_sort_plan(plan)
            it = 5 if N > (1 << 24) else 50
            for _ in range(3):
                prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(it):
                prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) / it * 1e3
            print("plan=%d sort 2^%d %-9s %8.1f us  %6.2f Gkeys/s" % (plan, N.bit_length() - 1, kind, us, N / us / 1e3),
                  flush=True)
    del sv, dk, dv, bk, bv
prim.lib.b200_tune_sort_plan(1)
