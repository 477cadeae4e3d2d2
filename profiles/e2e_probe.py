#!/usr/bin/env python3
"""Where does the end-to-end (host buffer in, result out) reduce lose time?  Prints, for GPU 0 (or $LOCAL_RANK):
  * the box: CPUs allowed, NUMA nodes, the GPU's node, where torch-pinned and b200_host_alloc pages land;
  * plain pinned cudaMemcpyAsync H2D of 1 GiB (the PCIe roofline) from both kinds of buffer;
  * b200_reduce_{i32,f32}_host over 2^28 elements: zero-copy vs staged (8/32/128 MiB chunks).
    python profiles/e2e_probe.py"""
import glob
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402

dev = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(dev)
N = 1 << 28
print("allowed cpus:", len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:4], "...",
      "| nodes:", [os.path.basename(p) for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*"))])
for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
    try:
        print(" ", os.path.basename(p), "cpulist", open(p + "/cpulist").read().strip())
    except OSError as e:
        print(" ", p, e)

xi = torch.empty(N, dtype=torch.int32, device="cuda")
prim.generate(xi, "i32_byte", 0, N, 1)
want = int(xi.sum(dtype=torch.int64).item())
want = ((want + 2**31) % 2**32) - 2**31

tp = torch.empty(N, dtype=torch.int32, pin_memory=True)
tp.copy_(xi)
hb = prim.HostBuffer(4 * N)
print("b200_host_alloc:", hb.info())
ha = hb.array(np.int32)
torch.from_numpy(ha).copy_(xi)
torch.cuda.synchronize()
print("pcie probe, torch pinned   : %.1f GB/s" % prim.host_probe_h2d(tp, 4 * N, 5))
print("pcie probe, b200_host_alloc: %.1f GB/s" % prim.host_probe_h2d(ha, 4 * N, 5))


def run(label, buf, mode, chunk):
    prim.tune_host_path(mode, chunk)
    r = prim.reduce_host(buf)
    assert r == want, (label, r, want)
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        prim.reduce_host(buf)
        best = min(best, time.perf_counter() - t0)
    print("%-44s %7.2f ms  %6.1f GB/s" % (label, best * 1e3, 4 * N / best / 1e9), flush=True)


for name, buf in (("torch pinned", tp), ("b200_host_alloc", ha)):
    run(name + " zero-copy", buf, 1, 0)
    for mb in (8, 32, 128):
        run(name + " staged %d MiB" % mb, buf, 0, mb << 20)
pg = np.empty(N // 8, np.int32)  # pageable input still works (staged), 128 MiB is enough to see the rate
pg[:] = ha[:N // 8]
prim.tune_host_path(1, 32 << 20)
t0 = time.perf_counter()
r = prim.reduce_host(pg)
dt = time.perf_counter() - t0
print("pageable numpy input (staged, 128 MiB): %.1f GB/s, result ok: %s" % (pg.nbytes / dt / 1e9,
                                                                          r == ((int(pg.astype(np.int64).sum()) + 2**31) % 2**32) - 2**31))
