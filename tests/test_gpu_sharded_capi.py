"""The sharded primitives through the C ABI only, ONE process driving all devices (csrc/sharded.cu: b200_ctx_create +
b200_histogram_sharded / b200_reduce_sharded_* / b200_scan_exclusive_sharded / b200_sort_kv_sharded) -- the
simpleMultiGPU-style caller a C program would write, checked against the CPU gold of the concatenated input.
The one-device context runs everywhere; the all-devices test needs >= 2 GPUs (the driver's 1-GPU box skips it; the
same check runs inside `bench.py --gpus N`'s parity block through the multi-process path)."""
import numpy as np
import pytest

import oracle

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prim():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import cuda_samples_b200 as p
    return p


def on(dev, arr):
    return torch.from_numpy(arr).to("cuda:%d" % dev)


def sync_all(devs):
    for d in devs:
        torch.cuda.synchronize(d)


def run_all(prim, devs, seed):
    rng = np.random.default_rng(seed)
    G = len(devs)
    ctx = prim.ShardedContext(devs)
    try:
        # ---- histogram: ragged shard sizes, every device gets the global bins
        sizes = [(3 << 20) + 16 * i + (5 if i == G - 1 else 0) for i in range(G)]
        shards = [rng.integers(0, 256, s, dtype=np.uint8) for s in sizes]
        d_data = [on(d, s) for d, s in zip(devs, shards)]
        h256 = [torch.empty(256, dtype=torch.int32, device="cuda:%d" % d) for d in devs]
        h64 = [torch.empty(64, dtype=torch.int32, device="cuda:%d" % d) for d in devs]
        sync_all(devs)
        for _ in range(3):  # repeated calls exercise both slot parities of the exchange windows
            ctx.histogram(h256, h64, d_data, sizes)
        ctx.synchronize()
        whole = np.concatenate(shards)
        for a, b in zip(h256, h64):
            assert (a.cpu().numpy().view(np.uint32) == oracle.histogram256(whole)).all()
            assert (b.cpu().numpy().view(np.uint32) == oracle.histogram64(whole)).all()

        # ---- reduce: int wraps like reduceCPU<int>, float / double within the stated tolerance of Kahan
        ns = [(1 << 20) + 3 * i for i in range(G)]
        xi = [rng.integers(-2**31, 2**31, n, dtype=np.int64).astype(np.int32) for n in ns]
        xf = [rng.random(n, dtype=np.float32) for n in ns]
        xd = [rng.random(n) for n in ns]
        assert ctx.reduce([on(d, x) for d, x in zip(devs, xi)], ns) == oracle.reduce(np.concatenate(xi))
        want = oracle.reduce(np.concatenate(xf))
        assert abs(ctx.reduce([on(d, x) for d, x in zip(devs, xf)], ns) - want) <= 1e-5 * abs(want)
        want = oracle.reduce(np.concatenate(xd))
        assert abs(ctx.reduce([on(d, x) for d, x in zip(devs, xd)], ns) - want) <= 1e-12 * abs(want)

        # ---- scan: one array in contiguous shards of different lengths (one of them short)
        ns = [(1 << 20) + 8192 * i for i in range(G)]
        ns[-1] = 4099
        src = [rng.integers(0, 2**32, n, dtype=np.uint32) for n in ns]
        d_src = [on(d, s.view(np.int32)) for d, s in zip(devs, src)]
        d_dst = [torch.empty_like(t) for t in d_src]
        sync_all(devs)
        for _ in range(2):
            ctx.scan_exclusive(d_dst, d_src, ns)
        ctx.synchronize()
        whole = np.concatenate(src)
        want = oracle.scan_exclusive(whole, 1, whole.size)
        off = 0
        for t, n in zip(d_dst, ns):
            assert (t.cpu().numpy().view(np.uint32) == want[off:off + n]).all()
            off += n

        # ---- the sharded scan 24 times back to back, no host synchronisation in between, short peer timeout: a device
        #      that ran ahead of its peers' reads of the exchange windows would lose a word, and synchronize() raises
        if G > 1:
            ns2 = [(1 << 20) + 8192 * i for i in range(G)]  # every shard takes the tile-prefix path
            src2 = [rng.integers(0, 2**32, n, dtype=np.uint32) for n in ns2]
            d_src2 = [on(d, s.view(np.int32)) for d, s in zip(devs, src2)]
            d_dst2 = [torch.empty_like(t) for t in d_src2]
            sync_all(devs)
            torch.cuda.set_device(devs[0])
            prim.p2p_set_timeout_ms(2000)  # (applies to every device of the context)
            try:
                for _ in range(24):
                    ctx.scan_exclusive(d_dst2, d_src2, ns2)
                ctx.synchronize()
            finally:
                prim.p2p_set_timeout_ms(30000)
            whole = np.concatenate(src2)
            want = oracle.scan_exclusive(whole, 1, whole.size)
            off = 0
            for t, n in zip(d_dst2, ns2):
                assert (t.cpu().numpy().view(np.uint32) == want[off:off + n]).all()
                off += n

        # ---- sort: three key distributions, both directions; too small a capacity is a status, not a crash
        n = 1 << 20
        cap = int(1.5 * n) if G > 1 else n
        for kind, param in (("key_unique", 0), ("key_tile_unique", 1 << 12), ("key_mod", 1000)):
            keys = oracle.generate(kind, 0, G * n, 7, param)
            vals = np.arange(G * n, dtype=np.uint32)
            sk = [on(d, keys[i * n:(i + 1) * n].view(np.int32).copy()) for i, d in enumerate(devs)]
            sv = [on(d, vals[i * n:(i + 1) * n].view(np.int32).copy()) for i, d in enumerate(devs)]
            dk = [torch.empty(cap, dtype=torch.int32, device="cuda:%d" % d) for d in devs]
            dv = [torch.empty(cap, dtype=torch.int32, device="cuda:%d" % d) for d in devs]
            sync_all(devs)
            for direction in (1, 0):
                n_out = ctx.sort_kv(dk, dv, cap, sk, sv, [n] * G, direction)
                assert sum(n_out) == G * n
                wk, wv = oracle.stable_sort_kv(keys, vals, direction)
                gk = np.concatenate([t[:m].cpu().numpy().view(np.uint32) for t, m in zip(dk, n_out)])
                gv = np.concatenate([t[:m].cpu().numpy().view(np.uint32) for t, m in zip(dv, n_out)])
                assert (gk == wk).all() and (gv == wv).all(), (kind, direction)
        if G > 1:
            for t in dk + dv:
                t.fill_(0x5A5A5A5A)
            sync_all(devs)
            with pytest.raises(prim.B200Error):
                ctx.sort_kv(dk, dv, n // 4, sk, sv, [n] * G, 1)
            sync_all(devs)
            for t in dk + dv:  # "nothing was moved": the destination arrays are as the caller left them
                assert bool((t == 0x5A5A5A5A).all())
            n_out = ctx.sort_kv(dk, dv, cap, sk, sv, [n] * G, 1)  # and the context is still usable afterwards
            assert sum(n_out) == G * n
    finally:
        ctx.destroy()


def test_sharded_c_abi_one_device(prim):
    run_all(prim, [0], 1)


def test_sharded_c_abi_one_process_all_devices(prim):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    run_all(prim, list(range(min(n, 8))), 2)
    run_all(prim, list(range(min(n, 8))), 3)  # a second context over the same devices (create / destroy is clean)


def test_large_smem_kernels_launch_on_a_second_device(prim):
    """cudaFuncSetAttribute is per device: the >48 KiB-smem scan and sort kernels must launch on every
    device of a process, not only on the first one that used them."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    rng = np.random.default_rng(4)
    for dev in (0, 1):
        with torch.cuda.device(dev):
            src = rng.integers(0, 2**32, 1 << 20, dtype=np.uint32)
            s = on(dev, src.view(np.int32))
            d = torch.empty_like(s)
            prim.scanExclusiveLarge(d, s, 1, src.size)
            assert (d.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(src, 1, src.size)).all()
            keys = rng.integers(0, 2**32, 1 << 18, dtype=np.uint32)
            vals = np.arange(keys.size, dtype=np.uint32)
            t = [on(dev, keys.view(np.int32)), on(dev, vals.view(np.int32))]
            t += [torch.empty(keys.size, dtype=torch.int32, device="cuda:%d" % dev) for _ in range(4)]
            prim.mergeSort(t[2], t[3], t[4], t[5], t[0], t[1], keys.size, 1)
            wk, wv = oracle.stable_sort_kv(keys, vals, 1)
            assert (t[2].cpu().numpy().view(np.uint32) == wk).all() and (t[3].cpu().numpy().view(np.uint32) == wv).all()
