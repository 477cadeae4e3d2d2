#!/usr/bin/env python3
"""Chained scan of long arrays: tiles that wait for their prefix in shared memory (round 1) against the reload
pipeline (stage freed after phase A, phase B re-reads the tile from L2).  Parity of both against torch.cumsum, then
timing (CUDA events).   python profiles/scan_reload.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

N = 1 << 28
src = torch.empty(N, dtype=torch.int32, device="cuda")
dst = torch.empty(N, dtype=torch.int32, device="cuda")
prim.generate(src, "u31", 0, N, 2009)


def want(batch, L):
    x = src[:batch * L].view(batch, L).to(torch.int64)
    return ((torch.cumsum(x, 1) - x) & 0xFFFFFFFF)


def t(fn, it=20):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it * 1e3


if os.environ.get("SCAN_RELOAD_ONE"):  # target of an ncu capture: three launches of the reload pipeline, nothing else
    prim.lib.b200_tune_scan_reload(1)
    for _ in range(3):
        prim.scanExclusiveLarge(dst, src, 1, N)
    torch.cuda.synchronize()
    sys.exit(0)
for reload in (0, 1):
    prim.lib.b200_tune_scan_reload(reload)
    ok = True
    for batch, L in ((1, (1 << 20) + 4), (3, (1 << 21) + 12), (1, 1 << 26), (5, 3 << 20), (1, (1 << 27) - 8)):
        dst.fill_(-1)
        prim.scanExclusiveLarge(dst, src, batch, L)
        torch.cuda.synchronize()
        got = dst[:batch * L].view(batch, L).to(torch.int64) & 0xFFFFFFFF
        good = bool((got == want(batch, L)).all()) and bool((dst[batch * L:batch * L + 16] == -1).all())
        ok &= good
        if not good:
            print("reload=%d MISMATCH batch=%d L=%d" % (reload, batch, L))
    # in place
    cp = src[:1 << 24].clone()
    prim.scanExclusiveLarge(cp, cp, 1, 1 << 24)
    ok &= bool(((cp.to(torch.int64) & 0xFFFFFFFF) == want(1, 1 << 24)[0]).all())
    print("reload=%d parity %s" % (reload, "OK" if ok else "FAILED"), flush=True)
    for batch, L in ((1, N), (16, 1 << 24), (256, 1 << 20), (1, 1 << 25)):
        us = t(lambda: prim.scanExclusiveLarge(dst, src, batch, L))
        print("reload=%d scan %5d x 2^%-2d  %8.1f us  %7.1f GB/s" % (reload, batch, L.bit_length() - 1, us,
                                                                    8.0 * batch * L / us / 1e3), flush=True)
prim.lib.b200_tune_scan_reload(0)
