#!/usr/bin/env python3
"""histogram256 / histogram64 on 64 MiB (BASELINE configs[0]): microseconds per call for every launch shape --
programmatic dependent launch on/off x CTA geometry for short inputs (1024 threads, 512 x 1 per SM, 512 x 2 per SM) --
from a Python/ctypes loop and from a CUDA graph of 40 calls, over 4 rotating inputs (256 MiB > L2? no: B200's L2 is
126 MB, so 4 x 64 MiB rotate through it) and over ONE input (L2-resident), each checked against torch.bincount."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

n = 64 << 20
bufs = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(4)]
for i, b in enumerate(bufs):
    prim.generate(b, "bytes", i * n, n, 2009)
want = [torch.bincount(b.to(torch.int64), minlength=256).to(torch.int32) for b in bufs]
want64 = [w.view(64, 4).sum(1).to(torch.int32) for w in want]
h = torch.empty(256, dtype=torch.int32, device="cuda")
h64 = torch.empty(64, dtype=torch.int32, device="cuda")


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


for pdl in (0, 1):
    for small in (0, 1, 2):
        prim.lib.b200_tune_hist_launch(pdl, small)
        ok = True
        for i in range(8):  # back-to-back calls on alternating inputs: the result of every call must be its own
            prim.histogram256(h, bufs[i & 3], n)
            if i in (3, 7):
                ok &= bool((h == want[i & 3]).all())
        for i in range(4):
            prim.histogram64(h64, bufs[i], n)
        ok &= bool((h64 == want64[3]).all())
        prim.histogram256(h, bufs[1], n)
        prim.histogram64(h64, bufs[2], n)
        prim.histogram256(h, bufs[0], n)
        torch.cuda.synchronize()
        ok &= bool((h == want[0]).all()) and bool((h64 == want64[2]).all())
        row = {}
        for label, nb in (("rotating4", 4), ("one_input_L2", 1)):
            it = [0]

            def step():
                prim.histogram256(h, bufs[it[0] % nb], n)
                it[0] += 1
            row[label + "_loop"] = timed(step, 200)
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(4):
                    step()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    for _ in range(40):
                        step()
                row[label + "_graph"] = timed(g.replay, 10) / 40
                del g
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
        prim.histogram256(h, bufs[2], n)
        torch.cuda.synchronize()
        ok &= bool((h == want[2]).all())
        print("pdl=%d small_ctas=%d  parity %s | us per call: loop %.2f  graph %.2f | one input (L2): loop %.2f  graph %.2f"
              % (pdl, small, "OK" if ok else "FAILED", row["rotating4_loop"], row["rotating4_graph"],
                 row["one_input_L2_loop"], row["one_input_L2_graph"]), flush=True)
prim.lib.b200_tune_hist_launch(1, 1)
