"""Host logic of the sharded paths (cuda_samples_b200/multi_gpu.py) on CPU: gloo, world_size 2.
The local kernels are replaced by the oracle-backed test double in tests/_cpu_ops.py; what is under
test is the exchange logic: bins all_reduce, rank-ordered float sum, shard carry from all_gathered totals,
splitter choice, bucket-count exchange, all-to-all layout in source-rank order, global stability."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, case, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from _cpu_ops import OracleOps
        from cuda_samples_b200 import multi_gpu as mg
        ops = OracleOps()
        rng = np.random.default_rng(1234)  # same stream on every rank: the global input
        if case == "histogram":
            n = 1 << 18
            data = rng.integers(0, 256, world * n, dtype=np.uint8)
            shard = torch.from_numpy(data[rank * n:(rank + 1) * n].copy())
            h256, h64 = mg.histogram_sharded(shard, n, ops=ops)
            ok = (h256.numpy().view(np.uint32) == oracle.histogram256(data)).all() and \
                 (h64.numpy().view(np.uint32) == oracle.histogram64(data)).all()
        elif case == "reduce":
            n = 1 << 18
            xi = rng.integers(-2**31, 2**31, world * n, dtype=np.int64).astype(np.int32)
            xf = rng.random(world * n, dtype=np.float32)
            gi = mg.reduce_sharded(torch.from_numpy(xi[rank * n:(rank + 1) * n].copy()), n, ops=ops)
            gf = mg.reduce_sharded(torch.from_numpy(xf[rank * n:(rank + 1) * n].copy()), n, ops=ops)
            want = oracle.reduce(xf)
            ok = int(gi.item()) == oracle.reduce(xi) and abs(float(gf.item()) - want) <= 1e-5 * abs(want)
            # the one-collective form (what bench.py times)
            pend = mg.reduce_many_sharded([torch.from_numpy(xi[rank * n:(rank + 1) * n].copy()),
                                           torch.from_numpy(xf[rank * n:(rank + 1) * n].copy())], [n, n], ops=ops)
            ti, tf = pend.typed()
            ok = ok and ti == oracle.reduce(xi) and abs(tf - want) <= 1e-5 * abs(want)
        elif case == "scan":
            n = (1 << 16) + 12
            src = rng.integers(0, 2**32, world * n, dtype=np.uint32)
            s = torch.from_numpy(src[rank * n:(rank + 1) * n].view(np.int32).copy())
            d = torch.empty_like(s)
            mg.scan_sharded(d, s, n, ops=ops)
            want = oracle.scan_exclusive(src, 1, world * n)[rank * n:(rank + 1) * n]
            ok = (d.numpy().view(np.uint32) == want).all()
        else:  # sort_asc / sort_desc / sort_dups
            direction = 0 if case == "sort_desc" else 1
            n = 50000 + 17 * rank  # unequal shards
            sizes = [50000 + 17 * r for r in range(world)]
            total = sum(sizes)
            hi = 40 if case == "sort_dups" else 2**32
            keys = rng.integers(0, hi, total, dtype=np.uint32)
            vals = np.arange(total, dtype=np.uint32)
            off = sum(sizes[:rank])
            k = torch.from_numpy(keys[off:off + n].view(np.int32).copy())
            v = torch.from_numpy(vals[off:off + n].view(np.int32).copy())
            dk, dv, n_out, m = mg.sort_sharded(k, v, n, direction, ops=ops, samples_per_rank=256)
            # gather every rank's output range and compare with the global stable sort
            outs = [None] * world
            dist.all_gather_object(outs, (dk[:n_out].numpy().view(np.uint32).copy(),
                                          dv[:n_out].numpy().view(np.uint32).copy()))
            gk = np.concatenate([o[0] for o in outs])
            gv = np.concatenate([o[1] for o in outs])
            wk, wv = oracle.stable_sort_kv(keys, vals, direction)
            ok = gk.size == total and (gk == wk).all() and (gv == wv).all() and int(m.sum()) == total
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case", ["histogram", "reduce", "scan", "sort_asc", "sort_desc", "sort_dups"])
def test_sharded_paths_world2_gloo(case):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0, "rank crashed or hung"
    res = dict(q.get(timeout=5) for _ in range(world))
    assert res == {0: True, 1: True}


def test_shard_carry_wraps():
    from cuda_samples_b200 import multi_gpu as mg
    totals = torch.from_numpy(np.array([0xFFFFFFF0, 0x20, 5], np.uint32).view(np.int32).copy())
    assert mg.shard_carry(totals, 0).numpy().view(np.uint32)[0] == 0
    assert mg.shard_carry(totals, 1).numpy().view(np.uint32)[0] == 0xFFFFFFF0
    assert mg.shard_carry(totals, 2).numpy().view(np.uint32)[0] == 0x10


def test_choose_splitters_is_deterministic_and_ordered():
    from cuda_samples_b200 import multi_gpu as mg
    s = torch.tensor([9, 1, 5, 3, 7, 2, 8, 4], dtype=torch.int64)
    spl = mg.choose_splitters(s, 4)
    assert spl.tolist() == [3, 5, 8]
