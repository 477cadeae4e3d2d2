#!/usr/bin/env python3
"""Where the time of the multi-GPU sort goes (CUDA events per phase, rank 0 prints).  Launch under torchrun:
    python -m torch.distributed.run --nproc-per-node G profiles/sort_sharded_phases.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim  # noqa: E402
from cuda_samples_b200 import api, multi_gpu as mg  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
mg.enable_p2p()
if os.environ.get("SORT_PEER_BULK"):  # 0: 8-byte lane stores into the peers' pair buffers, 1: cp.async.bulk from shared memory
    prim.lib.b200_tune_sort_peer_bulk(int(os.environ["SORT_PEER_BULK"]))
n = (1 << 28) // world
sk = torch.empty(n, dtype=torch.int32, device="cuda")
sv = torch.arange(rank * n, rank * n + n, dtype=torch.int64, device="cuda").to(torch.int32)
prim.generate(sk, "key_unique", rank * n, n, 2009)
for _ in range(3):
    mg.sort_sharded(sk, sv, n, 1)
b = mg._recv_buffers(n, 1.25, rank, world, None)
cap = b["capacity"]
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
acc = [0.0] * 5
REPS = 5
for _ in range(REPS):
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    ev[0].record()
    spl = b["work"]["spl"]
    api.sort_splitters(sk, n, 1, spl)  # jittered samples -> peer windows -> bitonic sort -> the same splitters everywhere
    ev[1].record()
    meta = torch.empty(world * world + 2, dtype=torch.int32, device="cuda")
    api.partition_exchange_kv(sk, sv, n, spl, 1, b["pairs"], None, cap, meta[:world * world], meta[world * world:])
    ev[2].record()
    w = b["work"]
    dk, dv = w["dk"], w["dv"]
    ev[3].record()
    api.sort_kv_dyn(dk, dv, None, None, b["pairs"][rank], None, cap, meta[world * world:], 1)
    ev[4].record()
    host = meta.cpu()
    ev[5].record()
    torch.cuda.synchronize()
    for i in range(5):
        acc[i] += ev[i].elapsed_time(ev[i + 1]) / REPS
if rank == 0:
    for name, t in zip(("splitters (one kernel, peer windows)", "partition + exchange (2 kernels)", "(nothing)",
                        "local sort (dyn count)", "meta.cpu()"), acc):
        print("world=%d  %-38s %8.1f us" % (world, name, t * 1e3))
    print("world=%d  total %.1f us; received %d of capacity %d" % (world, sum(acc) * 1e3, int(host[-2]), cap))
mg.disable_p2p()
dist.destroy_process_group()
