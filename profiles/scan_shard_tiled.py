#!/usr/bin/env python3
"""One rank's share of the sharded scan on ONE GPU with a world-1 peer window: shard total + exchange + CHAINED scan
seeded with the carry (round-2 path, 5.4 TB/s second pass) against tile sums + exchange + tile prefixes + scan of
INDEPENDENT tiles (no look-back).  Shard sizes of the 2^28-element config at 8 / 4 / 2 GPUs.
    python profiles/scan_shard_tiled.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

prim.p2p_connect(0, 1, [prim.p2p_window_create()])
for n in (1 << 25, 1 << 26, 1 << 27):
    src = torch.empty(n, dtype=torch.int32, device="cuda")
    dst = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", 0, n, 2009)
    carry = torch.empty(1, dtype=torch.int32, device="cuda")
    work = torch.empty(2 + 16, dtype=torch.float64, device="cuda")
    ref = None
    for tiled, pdl in ((0, 1), (1, 0), (1, 1)):
        prim.lib.b200_tune_scan_shard_tiled(tiled)
        prim.lib.b200_tune_scan_pdl(pdl)

        def step():
            prim.scan_exclusive_shard(dst, src, n, work, carry)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            step()
        b.record()
        torch.cuda.synchronize()
        if ref is None:
            ref = dst.clone()
        same = bool((ref == dst).all())
        us = a.elapsed_time(b) * 1e3 / 20
        print("shard 2^%d u32 (%4d MiB): %s, PDL %d -> %6.1f us per call (%.0f GB/s at 12 B/element), same result: %s" %
              (n.bit_length() - 1, 4 * n >> 20, "tile sums + independent tiles" if tiled else "total + chained scan       ",
               pdl, us, 12.0 * n / us / 1e3, same), flush=True)
    del src, dst
prim.lib.b200_tune_scan_shard_tiled(1)
prim.lib.b200_tune_scan_pdl(1)
prim.p2p_disconnect()
