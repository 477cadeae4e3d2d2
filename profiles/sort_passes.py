#!/usr/bin/env python3
"""One sort of 2^28 pairs with keys < 65536: passes 0,1 see random digits, passes 2,3 a single digit.
Run under `ncu --metrics gpu__time_duration.sum` to compare the per-pass kernel durations."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

N = 1 << 28
sk = torch.empty(N, dtype=torch.int32, device="cuda")
sv = torch.arange(N, dtype=torch.int32, device="cuda")
dk, dv, bk, bv = (torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(4))
prim.generate(sk, "key_mod", 0, N, 2009, 65536)
for _ in range(2):
    prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
torch.cuda.synchronize()
print("done")
