#!/usr/bin/env python3
"""Chained scan of long arrays and the compaction on the same pipeline: the shipped pipeline against the variants
measured and REJECTED in round 2 (profiles/r02x_scan_*.log).  The variants are not in the tree: apply
profiles/r02x_scan_parked_two_ctas_rejected.patch (modes 2..4: tiles parked in register sets; 22..44: fewer stages;
100..113 with -DB200_SCAN_PARK_SWEEP: two CTAs of 8 scan warps per SM) or profiles/r02x_scan_aggregate_warps_rejected.patch
(mode 0: dedicated aggregate warps, mode 2: the shipped pipeline; -DB200_SCAN_LOOK_WARPS=n) to csrc/ first.
Parity of every mode against torch.cumsum / boolean indexing, then timing (CUDA events)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

N = 1 << 28
MODES = [int(m) for m in os.environ.get("SCAN_PARK_MODES", "0,2").split(",")]
src = torch.empty(N, dtype=torch.int32, device="cuda")
dst = torch.empty(N, dtype=torch.int32, device="cuda")
cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
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


for mode in MODES:
    prim.lib.b200_tune_scan_reload(mode)
    ok = True
    for batch, L in ((1, (1 << 20) + 4), (3, (1 << 21) + 12), (1, 1 << 26), (5, 3 << 20), (1, (1 << 27) - 8),
                     (150, 1 << 20)):
        dst.fill_(-1)
        prim.scanExclusiveLarge(dst, src, batch, L)
        torch.cuda.synchronize()
        got = dst[:batch * L].view(batch, L).to(torch.int64) & 0xFFFFFFFF
        good = bool((got == want(batch, L)).all()) and bool((dst[batch * L:batch * L + 16] == -1).all())
        ok &= good
        if not good:
            print("mode=%d MISMATCH batch=%d L=%d" % (mode, batch, L))
    cp = src[:1 << 24].clone()  # in place
    prim.scanExclusiveLarge(cp, cp, 1, 1 << 24)
    ok &= bool(((cp.to(torch.int64) & 0xFFFFFFFF) == want(1, 1 << 24)[0]).all())
    dst.fill_(-1)
    prim.scanInclusive(dst, src, 1, 1 << 25)
    x = src[:1 << 25].to(torch.int64)
    ok &= bool(((dst[:1 << 25].to(torch.int64) & 0xFFFFFFFF) == (torch.cumsum(x, 0) & 0xFFFFFFFF)).all())
    del x
    print("mode=%d scan parity %s" % (mode, "OK" if ok else "FAILED"), flush=True)
    for batch, L in ((1, N), (16, 1 << 24), (256, 1 << 20), (1, 1 << 25)):
        us = t(lambda: prim.scanExclusiveLarge(dst, src, batch, L))
        print("mode=%d scan %5d x 2^%-2d  %8.1f us  %7.1f GB/s" % (mode, batch, L.bit_length() - 1, us,
                                                                  8.0 * batch * L / us / 1e3), flush=True)
    if mode in (0, 2):
        csrc = torch.empty(N, dtype=torch.int32, device="cuda")
        for mod in (5, 2):
            prim.generate(csrc, "key_mod", 0, N, 7, mod)
            for n in (N, (1 << 24) + 4 * 12345):
                dst.fill_(-1)
                prim.filter_arr(dst, cnt, csrc, n)
                kept = int(cnt.item())
                w = csrc[:n][csrc[:n] > 0]
                good = kept == w.numel() and bool((dst[:kept] == w).all()) and bool((dst[kept:kept + 8] == -1).all())
                del w
                if not good:
                    print("mode=%d compaction MISMATCH mod=%d n=%d" % (mode, mod, n))
            us = t(lambda: prim.filter_arr(dst, cnt, csrc, N))
            kept = int(cnt.item())
            print("mode=%d compaction 2^28 keep %.0f%%  %8.1f us  %7.1f GB/s" %
                  (mode, 100.0 * kept / N, us, (4.0 * N + 4.0 * kept) / us / 1e3), flush=True)
        del csrc
prim.lib.b200_tune_scan_reload(0)
