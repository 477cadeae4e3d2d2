#!/usr/bin/env python3
"""Per-tile timeline of the TMA chained scan (debug build: python cuda_samples_b200/build.py --trace).
    B200PRIM_LIB=cuda_samples_b200/libb200prim_trace.so python profiles/scan_trace.py
Slots (globaltimer ns): 0 load issued, 1 scan warp 0 starts phase A (tile landed and its turn has come), 2 aggregate
ready (look-back warp woke up), 3 look-back done, 4 = number of polls, 5 scan warp 0 got the prefix, 6 scan warp 0 has
scanned the tile (results in registers, stores being issued), 7 scan warp 0 starts waiting for the prefix.
SCAN_MODE = b200_tune_scan_reload mode (0 tiles wait in shared memory, 1 reload, 2..4 parked register sets)."""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
from cuda_samples_b200._lib import lib  # noqa: E402

N = 1 << 28
lib.b200_tune_scan_backoff(int(os.environ.get('SCAN_BACKOFF', '100')))
lib.b200_tune_scan_reload(int(os.environ.get('SCAN_MODE', '0')))
print("SCAN_MODE", os.environ.get('SCAN_MODE', '0'))
tiles = -(-N // 10240)  # 40 KiB tiles (the geometry of arrays >= 2^20 elements)
src = torch.empty(N, dtype=torch.int32, device="cuda")
dst = torch.empty(N, dtype=torch.int32, device="cuda")
prim.generate(src, "u31", 0, N, 2009)
MODE = os.environ.get("SCAN_TRACE_MODE", "scan")  # "compact0": the compaction pipeline with nothing kept (no stores)
cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
if MODE == "compact0":
    src.neg_()
if MODE == "compact":  # keep 80 % like the sample's rand() % 5
    prim.generate(src, "key_mod", 0, N, 7, 5)


def run():
    if MODE in ("compact0", "compact"):
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
stat("load issued -> phase A starts", t[:, 1] - t[:, 0])
stat("phase A starts -> aggregate ready", t[:, 2] - t[:, 1])
stat("aggregate ready -> warp 0 waits for prefix", t[:, 7] - t[:, 2])
stat("warp 0 waits for prefix (blocked)", t[:, 5] - t[:, 7])
stat("aggregate ready -> look-back done", t[:, 3] - t[:, 2])
stat("look-back done -> scan warp has it", t[:, 5] - t[:, 3])
stat("scan warp: prefix -> tile scanned", t[:, 6] - t[:, 5])
per_sm = np.diff(np.sort(t[mid, 6]))
print("tile completion rate (whole GPU): %.1f ns per tile -> %.2f us per tile per SM" % (per_sm.mean(), per_sm.mean() * 148 / 1e3))
pass
stat("TOTAL load issued -> stage released", t[:, 6] - t[:, 0])
cta = (trace.cpu().numpy().reshape(tiles, 8)[:, 4] >> 32).astype(np.int64)
t[:, 4] = (trace.cpu().numpy().reshape(tiles, 8)[:, 4] & 0xffffffff).astype(np.float64)
# per-CTA sequences (tiles of one CTA in ticket order = the order its scan warps take them)
gaps = {"scanned(b) -> phase A starts (next tile in the stage queue): stores + wait for landing": [],
        "phase A starts -> same warp waits for a prefix: phase A": []}
ahead = int(os.environ.get("SCAN_AHEAD", "3"))
for c in range(148):
    idx = np.nonzero(cta == c)[0]
    idx = idx[(idx > tiles // 4) & (idx < 3 * tiles // 4)]
    if len(idx) < ahead + 4:
        continue
    # phase B of idx[i] is followed by phase A of idx[i + ahead + 1]; phase A of idx[i + ahead] precedes phase B of idx[i]
    a = t[idx[ahead + 1:], 1] - t[idx[:-(ahead + 1)], 6]
    b = t[idx[:-ahead], 7] - t[idx[ahead:], 1]
    gaps[list(gaps)[0]].extend(a.tolist())
    gaps[list(gaps)[1]].extend(b.tolist())
for k, v in gaps.items():
    v = np.array(v)
    if len(v):
        print("%-90s mean %6.2f us  p10 %6.2f  p50 %6.2f  p90 %6.2f" % (k, v.mean() / 1e3, np.percentile(v, 10) / 1e3,
                                                                      np.percentile(v, 50) / 1e3, np.percentile(v, 90) / 1e3))
print("polls per look-back: mean %.2f  p50 %.0f  p90 %.0f  max %.0f" % (t[mid, 4].mean(), np.percentile(t[mid, 4], 50),
                                                                      np.percentile(t[mid, 4], 90), t[mid, 4].max()))
# how far ahead of the resolved prefix front does a tile's aggregate appear?
agg, done = t[:, 2], t[:, 3]
order = np.arange(tiles)
print("look-back done time is monotone in tile index for %.1f %% of neighbours" % (100.0 * (np.diff(done) >= 0).mean()))
front = np.maximum.accumulate(done)
stat("aggregate ready -> prefix front passes", front - agg)
