#!/usr/bin/env python3
"""Launch every hot-path kernel at its bench size (1 warm-up + 2 launches each) -- the target of the
`ncu --set full` capture whose summaries live in this directory.  Not a benchmark: numbers taken under
a profiler are never reported as throughput."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
N = 1 << 28
torch.cuda.set_device(0)
REPS = int(os.environ.get("PROFILE_REPS", "3"))

if which in ("all", "reduce"):
    xi = torch.empty(N, dtype=torch.int32, device="cuda")
    xf = torch.empty(N, dtype=torch.float32, device="cuda")
    prim.generate(xi, "i32_byte", 0, N, 1)
    prim.generate(xf, "f32_byte", 0, N, 1)
    oi, of = torch.empty(64, dtype=torch.int32, device="cuda"), torch.empty(64, dtype=torch.float32, device="cuda")
    for _ in range(REPS):
        prim.reduce(N, 256, 64, 7, xi, oi)
        prim.reduce(N, 256, 64, 7, xf, of)
    torch.cuda.synchronize()
    del xi, xf

if which in ("all", "hist"):
    big = 1 << 30
    d = torch.empty(big, dtype=torch.uint8, device="cuda")
    prim.generate(d, "bytes", 0, big, 2009)
    h256, h64 = torch.empty(256, dtype=torch.int32, device="cuda"), torch.empty(64, dtype=torch.int32, device="cuda")
    for _ in range(REPS):
        prim.histogram256(h256, d, big)
        prim.histogram64(h64, d, big)
        prim.histogram256(h256, d, 64 << 20)
    torch.cuda.synchronize()
    del d

if which in ("all", "scan"):
    src = torch.empty(N, dtype=torch.int32, device="cuda")
    dst = torch.empty(N, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", 0, N, 2009)
    for _ in range(REPS):
        prim.scanExclusiveLarge(dst, src, 1, N)
        prim.scanExclusiveShort(dst, src, N // 1024, 1024)
    prim.generate(src, "key_mod", 0, N, 7, 5)
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    for _ in range(REPS):
        prim.filter_arr(dst, cnt, src, N)
    torch.cuda.synchronize()
    del src, dst

if which in ("all", "image"):
    ih = iw = 16384
    img = torch.empty(ih * iw, dtype=torch.uint8, device="cuda")
    prim.generate(img, "bytes", 0, ih * iw, 11)
    sat = torch.empty(ih * iw, dtype=torch.int32, device="cuda")
    for _ in range(REPS):
        prim.integral_image(sat, img, iw, ih)
    torch.cuda.synchronize()
    del img, sat

if which in ("all", "sort"):
    sk = torch.empty(N, dtype=torch.int32, device="cuda")
    sv = torch.arange(N, dtype=torch.int32, device="cuda")
    prim.generate(sk, "key_unique", 0, N, 2009)
    dk, dv, bk, bv = (torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(4))
    for _ in range(2):
        prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
    torch.cuda.synchronize()

if which in ("all", "seg"):
    for batch, L in ((16384, 64), (16384, 4096)):
        n = batch * L
        sk = torch.empty(n, dtype=torch.int32, device="cuda")
        prim.generate(sk, "key_mod", 0, n, 2009, 65536)
        sv = torch.arange(n, dtype=torch.int32, device="cuda")
        dk, dv = torch.empty(n, dtype=torch.int32, device="cuda"), torch.empty(n, dtype=torch.int32, device="cuda")
        for _ in range(2):
            prim.bitonicSort(dk, dv, sk, sv, batch, L, 1)
        torch.cuda.synchronize()
print("done", which)
