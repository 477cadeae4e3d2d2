#!/usr/bin/env python3
"""Parity of the sharded paths on real GPUs (one process per GPU, NCCL).  Launch with
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/multi_gpu_check.py
Every rank regenerates the global input with the counter-based generator on the CPU (oracle) and checks
its own slice of the result against the oracle run on the whole input."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_samples_b200 as prim  # noqa: E402
import oracle  # noqa: E402
from cuda_samples_b200 import multi_gpu as mg  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    failures = []
    sort_gold = {}

    def check(name, ok):
        if not ok:
            failures.append(name)

    def small_exchanges(tag):
        # histogram: 16 MiB per rank
        n = 16 << 20
        d = torch.empty(n, dtype=torch.uint8, device="cuda")
        prim.generate(d, "bytes", rank * n, n, 2009)
        h256, h64 = mg.histogram_sharded(d, n)
        whole = oracle.generate("bytes", 0, world * n, 2009)
        check(tag + "hist256", (h256.cpu().numpy().view(np.uint32) == oracle.histogram256(whole)).all())
        check(tag + "hist64", (h64.cpu().numpy().view(np.uint32) == oracle.histogram64(whole)).all())

        # reduce: 2^24 per rank
        n = 1 << 24
        xi = torch.empty(n, dtype=torch.int32, device="cuda")
        xf = torch.empty(n, dtype=torch.float32, device="cuda")
        prim.generate(xi, "i32_byte", rank * n, n, 1)
        prim.generate(xf, "f32_byte", rank * n, n, 1)
        gi, gf = mg.reduce_sharded(xi, n), mg.reduce_sharded(xf, n)
        check(tag + "reduce_int", int(gi.item()) == oracle.reduce(oracle.generate("i32_byte", 0, world * n, 1)))
        want = oracle.reduce(oracle.generate("f32_byte", 0, world * n, 1))
        check(tag + "reduce_float", abs(float(gf.item()) - want) <= 1e-5 * abs(want))
        ti, tf = mg.reduce_many_sharded([xi, xf], [n, n]).typed()
        check(tag + "reduce_many", ti == int(gi.item()) and abs(tf - want) <= 1e-5 * abs(want))

        # scan: one array of world * 2^24 elements, contiguous shards
        src = torch.empty(n, dtype=torch.int32, device="cuda")
        prim.generate(src, "u31", rank * n, n, 2009)
        dst = torch.empty_like(src)
        mg.scan_sharded(dst, src, n)
        want = oracle.scan_exclusive(oracle.generate("u31", 0, world * n, 2009), 1, world * n)[rank * n:(rank + 1) * n]
        check(tag + "scan", (dst.cpu().numpy().view(np.uint32) == want).all())

    small_exchanges("nccl:")
    # the same three paths with the exchange fused into the producing kernel (NVLink peer windows, csrc/p2p.cuh)
    check("enable_p2p", mg.enable_p2p())
    pay = torch.arange(rank * 100, rank * 100 + 24, dtype=torch.int32, device="cuda")
    got = torch.empty(world * 24, dtype=torch.int32, device="cuda")
    for _ in range(5):  # repeated exchanges exercise both slot parities
        prim.p2p_allgather(pay, 96, got)
    want_all = torch.cat([torch.arange(r * 100, r * 100 + 24, dtype=torch.int32) for r in range(world)])
    check("p2p_allgather", bool((got.cpu() == want_all).all()))
    for _ in range(3):
        small_exchanges("p2p:")
    torch.cuda.synchronize()
    check("p2p_errors", prim.p2p_errors() == 0)
    # Back to back with no host sync between the calls and the ranks deliberately out of step: every exchange must
    # make every rank wait for every rank, or a rank that needs nobody's words (rank 0 of a scan) runs ahead and
    # overwrites window slots its peers have not read yet (seen on 8 GPUs as lost words, never as wrong results).
    # All four sharded primitives alternate (histogram -> scan -> reduce -> sort), so every kind of exchange follows
    # every other kind through the same window slots.
    n = 1 << 24
    src = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", rank * n, n, 2009)
    dst = torch.empty_like(src)
    want = oracle.scan_exclusive(oracle.generate("u31", 0, world * n, 2009), 1, world * n)[rank * n:(rank + 1) * n]
    xi = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(xi, "i32_byte", rank * n, n, 1)
    want_sum = oracle.reduce(oracle.generate("i32_byte", 0, world * n, 1))
    nb = 16 << 20
    hb = torch.empty(nb, dtype=torch.uint8, device="cuda")
    prim.generate(hb, "bytes", rank * nb, nb, 2009)
    whole = oracle.generate("bytes", 0, world * nb, 2009)
    want_h256, want_h64 = oracle.histogram256(whole), oracle.histogram64(whole)
    del whole
    ns = 1 << 22
    skeys = oracle.generate("key_unique", 0, world * ns, 7, 0)
    svals = np.arange(world * ns, dtype=np.uint32)
    sk = torch.from_numpy(skeys[rank * ns:(rank + 1) * ns].view(np.int32).copy()).cuda()
    sv = torch.from_numpy(svals[rank * ns:(rank + 1) * ns].view(np.int32).copy()).cuda()
    want_sk, want_sv = oracle.stable_sort_kv(skeys, svals, 1)
    sort_gold[("key_unique", 0, 1)] = (want_sk, want_sv)
    mg.sort_sharded(sk, sv, ns, 1)  # first call sizes the receive buffers (it synchronises); the loop below does not
    prim.p2p_set_timeout_ms(2000)
    for late in range(min(world, 3)):
        torch.cuda.synchronize()
        dist.barrier()
        if rank == late:
            time.sleep(0.05)
        for _ in range(24):
            h256, h64 = mg.histogram_sharded(hb, nb)
            mg.scan_sharded(dst, src, n)
            gi = mg.reduce_sharded(xi, n)
            dk, dv, meta, _ = mg.sort_sharded(sk, sv, ns, 1, sync=False)
        torch.cuda.synchronize()
        tag = " (rank %d late)" % late
        check("p2p back-to-back hist" + tag, (h256.cpu().numpy().view(np.uint32) == want_h256).all() and
              (h64.cpu().numpy().view(np.uint32) == want_h64).all())
        check("p2p back-to-back scan" + tag, (dst.cpu().numpy().view(np.uint32) == want).all())
        check("p2p back-to-back reduce" + tag, int(gi.item()) == want_sum)
        host = meta.cpu().numpy().view(np.uint32)
        m = host[:world * world].reshape(world, world)
        n_out, off = int(host[-2]), int(m[:, :rank].sum())
        check("p2p back-to-back sort" + tag, int(host[-1]) == 0 and int(m.sum()) == world * ns and
              (dk[:n_out].cpu().numpy().view(np.uint32) == want_sk[off:off + n_out]).all() and
              (dv[:n_out].cpu().numpy().view(np.uint32) == want_sv[off:off + n_out]).all())
        check("p2p back-to-back errors" + tag, prim.p2p_errors() == 0)
    del hb, sk, sv, skeys, svals
    prim.p2p_set_timeout_ms(30000)
    del src, dst, xi
    # latency of the exchange alone: NCCL all_gather vs the peer-window kernel (CUDA events, 200 calls each)
    small = torch.zeros(2, dtype=torch.float64, device="cuda")
    out = torch.empty(2 * world, dtype=torch.float64, device="cuda")
    for name, fn in (("nccl all_gather_into_tensor", lambda: dist.all_gather_into_tensor(out, small)),
                     ("p2p_allgather kernel", lambda: prim.p2p_allgather(small, 16, out))):
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(200):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if rank == 0:
            print("exchange of 16 B per rank, world=%d: %-28s %.2f us per call" % (world, name, e0.elapsed_time(e1) * 5),
                  flush=True)
    mg.disable_p2p()

    # sort: world * 2^22 pairs, three datasets, both directions
    n = 1 << 22

    def sort_checks(tag, **kw):
        for kind, param in (("key_unique", 0), ("key_tile_unique", 1 << 12), ("key_mod", 1000)):
            keys = oracle.generate(kind, 0, world * n, 7, param)
            vals = np.arange(world * n, dtype=np.uint32)
            for direction in (1, 0):
                sk = torch.from_numpy(keys[rank * n:(rank + 1) * n].view(np.int32).copy()).cuda()
                sv = torch.from_numpy(vals[rank * n:(rank + 1) * n].view(np.int32).copy()).cuda()
                dk, dv, n_out, m = mg.sort_sharded(sk, sv, n, direction, **kw)
                if (kind, param, direction) not in sort_gold:
                    sort_gold[(kind, param, direction)] = oracle.stable_sort_kv(keys, vals, direction)
                wk, wv = sort_gold[(kind, param, direction)]
                off = int(m[:, :rank].sum())  # rows = source, cols = destination: everything sent to lower ranks
                gk, gv = dk[:n_out].cpu().numpy().view(np.uint32), dv[:n_out].cpu().numpy().view(np.uint32)
                check("%s sort %s dir=%d" % (tag, kind, direction), int(m.sum()) == world * n and
                      (gk == wk[off:off + n_out]).all() and (gv == wv[off:off + n_out]).all())

    sort_checks("nccl:")
    # partition pass storing straight into the peers' receive buffers (twice: the buffers are reused), and with
    # receive buffers that are too small: every rank must fall back to the NCCL path with the same result
    mg.enable_p2p()
    sort_checks("p2p:")
    sort_checks("p2p again:")
    sort_checks("p2p overflow->fallback:", capacity_factor=0.5)
    torch.cuda.synchronize()
    check("p2p_errors after sort", prim.p2p_errors() == 0)
    mg.disable_p2p()

    flag = torch.tensor([len(failures)], dtype=torch.int32, device="cuda")
    dist.all_reduce(flag)
    if failures:
        print("rank %d FAILED: %s" % (rank, failures), flush=True)
    if rank == 0:
        print("multi_gpu_check world=%d: %s" % (world, "OK" if int(flag.item()) == 0 else "FAILED"), flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
