#!/usr/bin/env python3
"""Per-tile timeline of one radix pass (debug build: python cuda_samples_b200/build.py --trace).
    B200PRIM_LIB=cuda_samples_b200/libb200prim_trace.so python profiles/sort_trace.py
Slots (globaltimer ns, thread 0 of each tile), r02 order of the pass (early counts): 0 ticket drawn, 1 keys in
registers, 2 per-warp digit counts done, 3 tile counts published, 6 ranked + keys staged, 4 look-back done (digit 0),
5 all digits resolved and values staged (barrier).  SORT_LB=1|4|8 sets the look-back rows per round trip."""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
from cuda_samples_b200._lib import lib  # noqa: E402

if os.environ.get("SORT_LB"):
    lib.b200_tune_sort_lookback(int(os.environ["SORT_LB"]))
N = 1 << 28
TILE = int(os.environ.get("B200_SORT_TILE", "8192"))
tiles = (N + TILE - 1) // TILE
sk = torch.empty(N, dtype=torch.int32, device="cuda")
sv = torch.arange(N, dtype=torch.int32, device="cuda")
dk, dv, bk, bv = (torch.empty(N, dtype=torch.int32, device="cuda") for _ in range(4))
prim.generate(sk, "key_unique", 0, N, 2009)
for _ in range(2):
    prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)
trace = torch.zeros(tiles * 8, dtype=torch.int64, device="cuda")
lib.b200_debug_set_sort_trace.argtypes = [ctypes.c_void_p]
lib.b200_debug_set_sort_trace(ctypes.c_void_p(trace.data_ptr()))
torch.cuda.synchronize()
prim.mergeSort(dk, dv, bk, bv, sk, sv, N, 1)  # the 4 passes overwrite each other: the trace shows the last pass
torch.cuda.synchronize()
lib.b200_debug_set_sort_trace(ctypes.c_void_p(0))
t = trace.cpu().numpy().reshape(tiles, 8).astype(np.float64)
mid = slice(tiles // 4, 3 * tiles // 4)


def stat(name, x):
    x = x[mid]
    print("%-40s mean %7.2f us  p10 %7.2f  p50 %7.2f  p90 %7.2f" % (name, x.mean() / 1e3, np.percentile(x, 10) / 1e3,
                                                                  np.percentile(x, 50) / 1e3, np.percentile(x, 90) / 1e3))


print("pass span %.1f us for %d tiles" % ((t[:, 5].max() - t[:, 0].min()) / 1e3, tiles))
stat("ticket -> keys in registers", t[:, 1] - t[:, 0])
stat("early digit counts (16 REDs)", t[:, 2] - t[:, 1])
stat("barrier + warp order + publish", t[:, 3] - t[:, 2])
stat("publish -> ranked, keys staged", t[:, 6] - t[:, 3])
stat("ranked -> look-back done (digit 0)", t[:, 4] - t[:, 6])
stat("values staged + wait for all digits", t[:, 5] - t[:, 4])
stat("TOTAL ticket -> ready to store", t[:, 5] - t[:, 0])
raw = trace.cpu().numpy().reshape(tiles, 8)[:, 7]
polls, stalls = (raw & 0xFFFFFFFF)[mid], (raw >> 32)[mid]
print("look-back of digit 0: polls per tile mean %.1f p50 %.0f p90 %.0f; of which found an unpublished row: mean %.1f"
      % (polls.mean(), np.percentile(polls, 50), np.percentile(polls, 90), stalls.mean()))
