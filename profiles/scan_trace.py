#!/usr/bin/env python3
"""Per-tile timeline of the TMA chained scan (debug build: python cuda_samples_b200/build.py --trace).
    B200PRIM_LIB=cuda_samples_b200/libb200prim_trace.so python profiles/scan_trace.py
Slots (globaltimer ns): 0 load issued, 1 landed (look-back warp saw `full`), 2 aggregate ready, 3 look-back done,
4 = number of polls, 5 scan warp 0 got the prefix, 6 scan warp 0 finished, 7 store issued."""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
from cuda_samples_b200._lib import lib  # noqa: E402

N = 1 << 28
lib.b200_tune_scan_backoff(int(os.environ.get('SCAN_BACKOFF', '0')))
tiles = -(-N // 10240)  # 40 KiB tiles (the geometry of arrays >= 2^20 elements)
src = torch.empty(N, dtype=torch.int32, device="cuda")
dst = torch.empty(N, dtype=torch.int32, device="cuda")
prim.generate(src, "u31", 0, N, 2009)
MODE = os.environ.get("SCAN_TRACE_MODE", "scan")  # "compact0": the compaction pipeline with nothing kept (no stores)
cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
if MODE == "compact0":
    src.neg_()
    pass


def run():
    if MODE == "compact0":
        prim.filter_arr(dst, cnt, src, N)
    else:
        prim.scanExclusiveLarge(dst, src, 1, N)


for _ in range(3):
    run()
trace = torch.zeros(tiles * 8, dtype=torch.int64, device="cuda")
lib.b200_debug_set_scan_trace.argtypes = [ctypes.c_void_p]
lib.b200_debug_set_scan_trace(ctypes.c_void_p(trace.data_ptr()))
torch.cuda.synchronize()
run()
torch.cuda.synchronize()
lib.b200_debug_set_scan_trace(ctypes.c_void_p(0))
t = trace.cpu().numpy().reshape(tiles, 8).astype(np.float64)
t0 = t[:, 0].min()
mid = slice(tiles // 4, 3 * tiles // 4)  # steady state


def stat(name, x):
    x = x[mid]
    print("%-34s mean %8.2f us   p10 %8.2f   p50 %8.2f   p90 %8.2f" % (name, x.mean() / 1e3, np.percentile(x, 10) / 1e3,
                                                                    np.percentile(x, 50) / 1e3, np.percentile(x, 90) / 1e3))


print("kernel span %.1f us for %d tiles" % ((t[:, 6].max() - t0) / 1e3, tiles))
stat("load issued -> landed", t[:, 1] - t[:, 0])
stat("landed -> aggregate ready", t[:, 2] - t[:, 1])
stat("aggregate ready -> look-back done", t[:, 3] - t[:, 2])
stat("look-back done -> scan warp has it", t[:, 5] - t[:, 3])
stat("scan warp: prefix -> stage released", t[:, 6] - t[:, 5])
pass
stat("TOTAL load issued -> stage released", t[:, 6] - t[:, 0])
print("polls per look-back: mean %.2f  p50 %.0f  p90 %.0f  max %.0f" % (t[mid, 4].mean(), np.percentile(t[mid, 4], 50),
                                                                      np.percentile(t[mid, 4], 90), t[mid, 4].max()))
# how far ahead of the resolved prefix front does a tile's aggregate appear?
agg, done = t[:, 2], t[:, 3]
order = np.arange(tiles)
print("look-back done time is monotone in tile index for %.1f %% of neighbours" % (100.0 * (np.diff(done) >= 0).mean()))
front = np.maximum.accumulate(done)
stat("aggregate ready -> prefix front passes", front - agg)
