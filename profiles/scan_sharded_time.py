#!/usr/bin/env python3
"""The sharded long scan (one 2^28-element array over all ranks) with the chained second pass (round-2 path) and with
the tile-prefix second pass, timed like bench.py (barrier, CUDA events, max over ranks), parity of both against each
other and against the CPU gold of rank 0's shard.  Launch under torchrun:
    python -m torch.distributed.run --nproc-per-node G profiles/scan_sharded_time.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
import oracle  # noqa: E402
from cuda_samples_b200 import multi_gpu as mg  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
mg.enable_p2p()
n = (1 << 28) // world
src = torch.empty(n, dtype=torch.int32, device="cuda")
dst = torch.empty(n, dtype=torch.int32, device="cuda")
prim.generate(src, "u31", rank * n, n, 2009)
ref = None
for tiled in (0, 1):
    prim.lib.b200_tune_scan_shard_tiled(tiled)
    for _ in range(5):
        mg.scan_sharded(dst, src, n)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50):
        mg.scan_sharded(dst, src, n)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / 50 * 1e3], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if ref is None:
        ref = dst.clone()
    same = torch.tensor([int(bool((ref == dst).all()))], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    if rank == 0:
        gold = oracle.scan_exclusive(oracle.generate("u31", 0, n, 2009), 1, n)
        ok = bool((dst.cpu().numpy().view(np.uint32) == gold).all())
        print("world=%d  scan 2^28: %s second pass  %7.1f us (max over ranks)  %.0f GB/s at 8 B/element; both paths equal on "
              "every rank: %s; rank 0 == scanExclusiveHost: %s" %
              (world, "tile-prefix" if tiled else "chained    ", float(t), 8.0 * (1 << 28) / float(t) / 1e3, bool(same.item()), ok),
              flush=True)
prim.lib.b200_tune_scan_shard_tiled(1)
mg.check_p2p() if hasattr(mg, "check_p2p") else None
mg.disable_p2p()
dist.destroy_process_group()
