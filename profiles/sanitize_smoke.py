#!/usr/bin/env python3
"""Small shapes of every primitive (ragged sizes, unaligned views) for compute-sanitizer runs:
    compute-sanitizer --tool memcheck  python profiles/sanitize_smoke.py
    compute-sanitizer --tool racecheck python profiles/sanitize_smoke.py      (one tool per gpurun call)
(Round 1: compute-sanitizer is closed on this GPU pool, so the script only runs as a plain small-shape parity check.)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
import oracle  # noqa: E402

rng = np.random.default_rng(0)
ok = True
for nbytes in (5, 4099, 70001):
    h = rng.integers(0, 256, nbytes + 3, dtype=np.uint8)
    d = torch.from_numpy(h).cuda()[1:1 + nbytes]
    o256, o64 = torch.empty(256, dtype=torch.int32).cuda(), torch.empty(64, dtype=torch.int32).cuda()
    prim.histogram_fused(o256, o64, d, nbytes)
    ok &= bool((o256.cpu().numpy().view(np.uint32) == oracle.histogram256(h[1:1 + nbytes])).all())
for n in (3, 1025, 50001):
    x = rng.random(n + 1, dtype=np.float32)
    out = torch.empty(4, dtype=torch.float32).cuda()
    prim.reduce(n, 256, 4, 7, torch.from_numpy(x).cuda()[1:], out)
    ok &= abs(float(out[0]) - oracle.reduce(x[1:])) <= 1e-5 * abs(oracle.reduce(x[1:]))
for batch, L in ((3, 5), (4, 256), (2, 8192 * 3 + 4), (1, 8192 * 40), (3, 20000)):
    s = rng.integers(0, 2**32, batch * L, dtype=np.uint32)
    dst = torch.empty(batch * L, dtype=torch.int32).cuda()
    prim.scanExclusiveLarge(dst, torch.from_numpy(s.view(np.int32)).cuda(), batch, L)
    ok &= bool((dst.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(s, batch, L)).all())
for n in (7, 4097, 30011):
    k = rng.integers(0, 1000, n, dtype=np.uint32)
    v = np.arange(n, dtype=np.uint32)
    t = [torch.from_numpy(k.view(np.int32)).cuda(), torch.from_numpy(v.view(np.int32)).cuda()]
    t += [torch.empty(n, dtype=torch.int32).cuda() for _ in range(4)]
    prim.mergeSort(t[2], t[3], t[4], t[5], t[0], t[1], n, 0)
    wk, wv = oracle.stable_sort_kv(k, v, 0)
    ok &= bool((t[2].cpu().numpy().view(np.uint32) == wk).all() and (t[3].cpu().numpy().view(np.uint32) == wv).all())
a, b = rng.random(5 * 1003, dtype=np.float32), rng.random(5 * 1003, dtype=np.float32)
c = torch.empty(5, dtype=torch.float32).cuda()
prim.scalarProd(c, torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), 5, 1003)
torch.cuda.synchronize()
print("sanitize_smoke results correct:", ok)
sys.exit(0 if ok else 1)
