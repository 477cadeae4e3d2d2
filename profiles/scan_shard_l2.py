#!/usr/bin/env python3
"""One rank's share of the sharded scan (shard total + exchange, then the chained scan seeded with the carry) on ONE
GPU with a world-1 peer window: does reading the shard back to front in the total pass leave its head in the L2 for
the scan?  Shard sizes of the 2^28-element config at 8 / 4 / 2 GPUs.   python profiles/scan_shard_l2.py"""
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
    for rev, pdl in ((0, 0), (1, 0), (1, 1)):
        prim.lib.b200_tune_scan_total_reverse(rev)
        prim.lib.b200_tune_scan_pdl(pdl)

        def step():
            prim.scan_total_exchange(carry, src, n, work)
            prim.scanExclusiveLarge(dst, src, 1, n, d_carry=carry)
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
        print("shard 2^%d u32 (%4d MiB): total pass %s, scan %s -> %.1f us per (total + exchange + scan), same "
              "result: %s" % (n.bit_length() - 1, 4 * n >> 20, "back to front" if rev else "front to back",
                              "overlaps the exchange (PDL)" if pdl else "behind the total kernel",
                              a.elapsed_time(b) * 1e3 / 20, same), flush=True)
    del src, dst
prim.lib.b200_tune_scan_total_reverse(1)
prim.lib.b200_tune_scan_pdl(1)
prim.p2p_disconnect()
