"""GPU parity tests (run on a B200: pytest -m gpu).

Every test drives the CUDA path through the C ABI of libb200prim.so (ctypes; torch only provides
device memory and the stream) and compares it with the oracle (oracle/: CPU restatement of the
samples' gold code) or with the committed golden vectors produced by the reference's own gold code.
Bar: bit-exact for histogram / scan / sort / integer reduce; fp32 <= 1e-5, fp64 <= 1e-12 relative
to the Kahan-summed reduceCPU for floating-point reductions.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import oracle  # noqa: E402

FP32_RTOL = 1e-5   # BASELINE.json north_star
FP64_RTOL = 1e-12


@pytest.fixture(scope="module")
def prim():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import cuda_samples_b200 as p
    info = p.device_info()
    assert info["cc"][0] == 10, "libb200prim.so is sm_100a only"
    return p


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def hx(a):
    return "%016x" % oracle.fnv1a64(a)


# ================================================================================== histogram
def run_hist(prim, data_t, nbytes, which):
    if which == "256":
        out = torch.full((256,), -1, dtype=torch.int32).cuda()
        prim.histogram256(out, data_t, nbytes)
        return out.cpu().numpy().view(np.uint32)
    if which == "64":
        out = torch.full((64,), 12345, dtype=torch.int32).cuda()
        prim.histogram64(out, data_t, nbytes)
        return out.cpu().numpy().view(np.uint32)
    o256 = torch.full((256,), 7, dtype=torch.int32).cuda()
    o64 = torch.full((64,), 7, dtype=torch.int32).cuda()
    prim.histogram_fused(o256, o64, data_t, nbytes)
    return o256.cpu().numpy().view(np.uint32), o64.cpu().numpy().view(np.uint32)


def test_histogram_small_golden(prim, golden_small):
    d = dev(golden_small["hist_in"])
    assert (run_hist(prim, d, d.numel(), "256") == golden_small["hist256_out"]).all()
    assert (run_hist(prim, d, d.numel(), "64") == golden_small["hist64_out"]).all()
    f256, f64 = run_hist(prim, d, d.numel(), "fused")
    assert (f256 == golden_small["hist256_out"]).all() and (f64 == golden_small["hist64_out"]).all()


@pytest.mark.parametrize("nbytes", [0, 1, 3, 4, 15, 16, 17, 4096, 65536 + 5, (1 << 20) + 13, 3 * (1 << 22) + 1])
def test_histogram_sizes_and_ragged(prim, nbytes):
    rng = np.random.default_rng(nbytes)
    h = rng.integers(0, 256, nbytes + 3, dtype=np.uint8)
    d = dev(h)
    for offset in (0, 1, 3):  # unaligned starts
        view = d[offset:offset + nbytes]
        want256, want64 = oracle.histogram256(h[offset:offset + nbytes]), oracle.histogram64(h[offset:offset + nbytes])
        assert (run_hist(prim, view, nbytes, "256") == want256).all()
        assert (run_hist(prim, view, nbytes, "64") == want64).all()
        f256, f64 = run_hist(prim, view, nbytes, "fused")
        assert (f256 == want256).all() and (f64 == want64).all()


@pytest.mark.parametrize("fill", ["zeros", "ff", "two_values", "ramp"])
def test_histogram_skewed(prim, fill):
    n = (8 << 20) + 48
    if fill == "zeros":
        h = np.zeros(n, np.uint8)
    elif fill == "ff":
        h = np.full(n, 255, np.uint8)
    elif fill == "two_values":
        h = np.where(np.arange(n) % 3 == 0, 17, 200).astype(np.uint8)
    else:
        h = (np.arange(n) & 255).astype(np.uint8)
    d = dev(h)
    assert (run_hist(prim, d, n, "256") == oracle.histogram256(h)).all()
    assert (run_hist(prim, d, n, "64") == oracle.histogram64(h)).all()


def test_histogram_kat_sample_input(prim, kat):
    """The sample's own input (histogram/main.cpp:92-96): 64 MiB of srand(2009); rand()%256."""
    n = kat["histogram256"]["bytes"]
    h = oracle.sample_input("hist_bytes", n, kat["histogram256"]["seed"])
    d = dev(h)
    prim.initHistogram256()
    h256 = run_hist(prim, d, n, "256")
    assert (int(h256[0]), int(h256[255]), hx(h256)) == (kat["histogram256"]["h0"], kat["histogram256"]["h255"],
                                                        kat["histogram256"]["fnv"])
    prim.initHistogram64()
    h64 = run_hist(prim, d, n, "64")
    assert (int(h64[0]), int(h64[63]), hx(h64)) == (kat["histogram64"]["h0"], kat["histogram64"]["h63"],
                                                    kat["histogram64"]["fnv"])
    # repeated calls overwrite (not accumulate) and stay identical
    for _ in range(3):
        assert (run_hist(prim, d, n, "256") == h256).all()


@pytest.mark.parametrize("cluster", [1, 2, 4, 8])
def test_histogram_cluster_sizes(prim, cluster):
    from cuda_samples_b200._lib import check, lib
    rng = np.random.default_rng(cluster)
    h = rng.integers(0, 256, (32 << 20) + 100, dtype=np.uint8)
    d = dev(h)
    check(lib.b200_tune_hist_cluster(cluster), "tune")
    try:
        assert (run_hist(prim, d, h.size, "256") == oracle.histogram256(h)).all()
        assert (run_hist(prim, d, 5000, "64") == oracle.histogram64(h[:5000])).all()
    finally:
        check(lib.b200_tune_hist_cluster(0), "tune")


def test_histogram_dropin_symbols(prim):
    """The samples' own C entry points (legacy default stream, uint byteCount)."""
    from cuda_samples_b200._lib import lib
    rng = np.random.default_rng(3)
    h = rng.integers(0, 256, 1 << 22, dtype=np.uint8)
    d = dev(h)
    out = torch.zeros(256, dtype=torch.int32).cuda()
    torch.cuda.synchronize()
    lib.initHistogram256()
    lib.histogram256(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(d.data_ptr()), h.size)
    torch.cuda.synchronize()
    assert (out.cpu().numpy().view(np.uint32) == oracle.histogram256(h)).all()
    lib.initHistogram64()
    lib.histogram64(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(d.data_ptr()), h.size)
    torch.cuda.synchronize()
    assert (out[:64].cpu().numpy().view(np.uint32) == oracle.histogram64(h)).all()


def test_histogram_large_generated(prim):
    """1 GiB of generator bytes: the device generator equals the CPU twin and the bins match."""
    n = 1 << 30
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    prim.generate(d, "bytes", 0, n, 2009)
    f256, f64 = run_hist(prim, d, n, "fused")
    assert int(f256.astype(np.uint64).sum()) == n and int(f64.astype(np.uint64).sum()) == n
    want = np.zeros(256, np.uint64)
    chunk = 1 << 26
    for off in range(0, n, chunk):  # CPU regenerates chunk by chunk (counter-based)
        want += oracle.histogram256(oracle.generate("bytes", off, chunk, 2009))
    assert (f256 == want.astype(np.uint32)).all()
    assert (f64 == want.reshape(64, 4).sum(1).astype(np.uint32)).all()


# ===================================================================================== reduce
def run_reduce(prim, x_t, n, blocks=1):
    out = torch.full((max(blocks, 1),), 77, dtype=x_t.dtype, device="cuda")
    prim.reduce(n, 256, blocks, 7, x_t, out)
    return out.cpu().numpy()


@pytest.mark.parametrize("n", [1, 2, 3, 5, 31, 1023, 4096, (1 << 20) + 3, (1 << 24)])
def test_reduce_sizes(prim, n):
    rng = np.random.default_rng(n)
    xi = rng.integers(-2**31, 2**31, n + 3, dtype=np.int64).astype(np.int32)
    xf = rng.random(n + 3, dtype=np.float32)
    xd = rng.random(n + 3, dtype=np.float64)
    for off in (0, 1, 3):
        got = run_reduce(prim, dev(xi)[off:off + n], n, 64)
        assert int(got[0]) == oracle.reduce(xi[off:off + n]) and not got[1:].any()
        got = run_reduce(prim, dev(xf)[off:off + n], n, 64)
        want = oracle.reduce(xf[off:off + n])
        assert abs(float(got[0]) - want) <= FP32_RTOL * abs(want) and not got[1:].any()
        got = run_reduce(prim, dev(xd)[off:off + n], n, 3)
        want = oracle.reduce(xd[off:off + n])
        assert abs(float(got[0]) - want) <= FP64_RTOL * abs(want) and not got[1:].any()


def test_reduce_kat_sample_input(prim, kat):
    """reduction.cpp:483-491 inputs (glibc seed 1): int exact; float/double within the stated tolerance."""
    for case in kat["reduce"]["cases"]:
        n = case["n"]
        got = run_reduce(prim, dev(oracle.sample_input("reduce_i32", n, 1)), n, 64)
        assert int(got[0]) == case["i32"]
        got = run_reduce(prim, dev(oracle.sample_input("reduce_f32", n, 1)), n, 64)
        want = float.fromhex(case["f32_hex"])
        assert abs(float(got[0]) - want) <= FP32_RTOL * abs(want)
        got = run_reduce(prim, dev(oracle.sample_input("reduce_f64", n, 1)), n, 64)
        want = float.fromhex(case["f64_hex"])
        assert abs(float(got[0]) - want) <= FP64_RTOL * abs(want)


def test_reduce_int_wraps_like_reduceCPU(prim):
    n = 1 << 26
    x = np.full(n, 0xFF, np.int32)  # 255 * 2^26 > 2^31: wraps negative
    got = run_reduce(prim, dev(x), n)
    assert int(got[0]) == oracle.reduce(x) == ((255 << 26) + 2**31) % 2**32 - 2**31


def test_reduce_full_size_and_deterministic(prim):
    """BASELINE config 2: 2^28 elements, int bit-exact, float within tolerance, run-to-run identical."""
    n = 1 << 28
    xi = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(xi, "i32_byte", 0, n, 1)
    xf = torch.empty(n, dtype=torch.float32, device="cuda")
    prim.generate(xf, "f32_byte", 0, n, 1)
    hi = oracle.generate("i32_byte", 0, n, 1)
    assert (xi[:1 << 20].cpu().numpy() == hi[:1 << 20]).all()
    got_i = run_reduce(prim, xi, n, 64)
    assert int(got_i[0]) == oracle.reduce(hi)
    hf = oracle.generate("f32_byte", 0, n, 1)
    assert (xf[-(1 << 20):].cpu().numpy() == hf[-(1 << 20):]).all()
    want = oracle.reduce(hf)
    got_f = [float(run_reduce(prim, xf, n, 64)[0]) for _ in range(3)]
    assert abs(got_f[0] - want) <= FP32_RTOL * abs(want)
    assert got_f[0] == got_f[1] == got_f[2]


def test_reduce_dropin_symbols(prim):
    from cuda_samples_b200._lib import lib
    n = 1 << 22
    x = oracle.sample_input("reduce_i32", n, 1)
    d, out = dev(x), torch.full((64,), 5, dtype=torch.int32).cuda()
    torch.cuda.synchronize()
    lib._Z6reduceIiEviiiiPT_S1_(n, 256, 64, 7, ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(out.data_ptr()))
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    assert int(o.sum()) == oracle.reduce(x) and int(o[0]) == oracle.reduce(x)


# ======================================================================================= scan
def run_scan(prim, src_t, batch, L, short=None, carry=None):
    dst = torch.full_like(src_t, 0x5A5A5A5A)
    if short is None:
        short = L <= prim.MAX_SHORT_ARRAY_SIZE
    if short and carry is None:
        prim.scanExclusiveShort(dst, src_t, batch, L)
    else:
        prim.scanExclusiveLarge(dst, src_t, batch, L, d_carry=carry)
    return dst.cpu().numpy().view(np.uint32)


def as_i32(a):
    return np.ascontiguousarray(a).view(np.int32)


@pytest.mark.parametrize("L", [4, 32, 1024, 2048, 8192])
def test_scan_small_golden(prim, golden_small, L):
    src = golden_small["scan_in"]
    got = run_scan(prim, dev(as_i32(src)), src.size // L, L)
    assert (got == golden_small["scan_out_L%d" % L]).all()


def test_scan_kat_sample_input(prim, kat):
    """scan/main.cpp:45-56 input, every array length the sample's two loops visit (hash per length)."""
    n = kat["scan"]["n"]
    a = oracle.sample_input("scan_u32", n, kat["scan"]["seed"])
    d = dev(as_i32(a))
    prim.initScan()
    golden = {c["L"]: c for c in kat["scan"]["cases"]}
    L = prim.MIN_SHORT_ARRAY_SIZE
    while L <= prim.MAX_LARGE_ARRAY_SIZE:
        got = run_scan(prim, d, n // L, L)
        if L in golden:
            assert int(got[-1]) == golden[L]["last"] and hx(got) == golden[L]["fnv"], L
        else:
            assert (got == oracle.scan_exclusive(a, n // L, L)).all(), L
        L *= 2


@pytest.mark.parametrize("batch,L", [(1, 1), (5, 1), (3, 2), (7, 3), (1, 5), (9, 100), (4, 1000), (3, 8191),
                                     (2, 8193), (3, 20000), (1, 8192 * 33 + 7), (2, 3 * 8192), (1, 1 << 22),
                                     (16, 1 << 16), (1, (1 << 24) + 4), (0, 16)])
def test_scan_arbitrary_shapes(prim, batch, L):
    rng = np.random.default_rng(batch * 1000003 + L)
    src = rng.integers(0, 2**32, max(batch * L, 1), dtype=np.uint32)[:batch * L]
    if batch * L == 0:
        prim.scanExclusiveLarge(torch.empty(4, dtype=torch.int32).cuda(), torch.empty(4, dtype=torch.int32).cuda(),
                                batch, L)
        return
    got = run_scan(prim, dev(as_i32(src)), batch, L, short=False)
    assert (got == oracle.scan_exclusive(src, batch, L)).all()


def test_scan_unaligned_and_in_place(prim):
    rng = np.random.default_rng(9)
    src = rng.integers(0, 2**32, (1 << 20) + 8, dtype=np.uint32)
    d = dev(as_i32(src))
    for off in (1, 2, 3):
        n = 1 << 20
        got = run_scan(prim, d[off:off + n], 1, n, short=False)
        assert (got == oracle.scan_exclusive(src[off:off + n], 1, n)).all()
    inplace = d.clone()
    prim.scanExclusiveLarge(inplace, inplace, 4, 1 << 18)
    assert (inplace.cpu().numpy().view(np.uint32)[:1 << 20] == oracle.scan_exclusive(src[:1 << 20], 4, 1 << 18)).all()


def test_scan_carry_and_total(prim):
    rng = np.random.default_rng(10)
    n = (1 << 21) + 12
    src = rng.integers(0, 2**32, n, dtype=np.uint32)
    carry = torch.tensor([0x89ABCDEF - 2**32], dtype=torch.int32).cuda()
    got = run_scan(prim, dev(as_i32(src)), 1, n, carry=carry)
    assert (got == oracle.scan_exclusive(src, 1, n, 0x89ABCDEF)).all()
    tot = torch.zeros(1, dtype=torch.int32).cuda()
    prim.scan_total(tot, dev(as_i32(src)), n)
    assert int(tot.cpu().numpy().view(np.uint32)[0]) == int(src.astype(np.uint64).sum() % 2**32)


def test_scan_full_size(prim):
    """BASELINE config 3 on one GPU: batch 1, 2^28 u32 (64x the sample's cap), compared element-wise."""
    n = 1 << 28
    d = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(d, "u31", 0, n, 2009)
    dst = torch.empty_like(d)
    for _ in range(2):  # second run: descriptors from the first must not leak (epoch)
        dst.fill_(-1)
        prim.scanExclusiveLarge(dst, d, 1, n)
    src = oracle.generate("u31", 0, n, 2009)
    want = oracle.scan_exclusive(src, 1, n)
    got = dst.cpu().numpy().view(np.uint32)
    assert (got == want).all()
    # the sample's largest legal batched shape as well: 1024 arrays of 262144
    dst.fill_(-1)
    prim.scanExclusiveLarge(dst, d, 1024, 262144)
    assert (dst.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(src, 1024, 262144)).all()


def test_scan_register_staged_path(prim):
    """The chained scan has two stagings (TMA bulk copies by default, plain 16-byte loads otherwise):
    both must give the gold result."""
    from cuda_samples_b200._lib import lib
    rng = np.random.default_rng(21)
    try:
        for enable in (0, 1):
            lib.b200_tune_scan_tma(enable)
            for batch, L in ((1, 1 << 22), (3, 8192 * 5 + 4), (2, 3 * 8192), (5, 1 << 16), (1, 8192 * 700 + 12)):
                src = rng.integers(0, 2**32, batch * L, dtype=np.uint32)
                got = run_scan(prim, dev(as_i32(src)), batch, L, short=False)
                assert (got == oracle.scan_exclusive(src, batch, L)).all(), (enable, batch, L)
            # many launches back to back: descriptors of earlier launches must never be taken for current ones
            n = 1 << 20
            src = rng.integers(0, 2**32, n, dtype=np.uint32)
            d, dst = dev(as_i32(src)), torch.empty(n, dtype=torch.int32).cuda()
            for _ in range(50):
                prim.scanExclusiveLarge(dst, d, 1, n)
            assert (dst.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(src, 1, n)).all()
    finally:
        lib.b200_tune_scan_tma(1)


def test_scan_dropin_symbols(prim):
    from cuda_samples_b200._lib import lib
    rng = np.random.default_rng(12)
    src = rng.integers(0, 2**32, 1 << 20, dtype=np.uint32)
    d, out = dev(as_i32(src)), torch.zeros(1 << 20, dtype=torch.int32).cuda()
    torch.cuda.synchronize()
    lib.initScan()
    assert lib.scanExclusiveShort(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(d.data_ptr()), 1 << 12, 256) > 0
    torch.cuda.synchronize()
    assert (out.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(src, 1 << 12, 256)).all()
    lib.scanExclusiveLarge(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(d.data_ptr()), 4, 262144)
    torch.cuda.synchronize()
    assert (out.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(src, 4, 262144)).all()


# ======================================================================================= sort
def run_sort(prim, keys, vals, direction):
    n = keys.size
    sk, sv = dev(as_i32(keys)), dev(as_i32(vals))
    m = max(n, 1)
    dk, dv = torch.full((m,), -3, dtype=torch.int32).cuda(), torch.full((m,), -3, dtype=torch.int32).cuda()
    bk, bv = torch.empty(m, dtype=torch.int32).cuda(), torch.empty(m, dtype=torch.int32).cuda()
    prim.mergeSort(dk, dv, bk, bv, sk, sv, n, direction)
    # src untouched (mergeSort.cu:502 never writes src)
    assert (sk.cpu().numpy().view(np.uint32) == keys).all() and (sv.cpu().numpy().view(np.uint32) == vals).all()
    return dk.cpu().numpy().view(np.uint32)[:n], dv.cpu().numpy().view(np.uint32)[:n]


@pytest.mark.parametrize("name", ["d1", "d2", "d3"])
@pytest.mark.parametrize("direction", [1, 0])
def test_sort_small_golden(prim, golden_small, name, direction):
    keys = golden_small["sort_%s_keys" % name]
    vals = np.arange(keys.size, dtype=np.uint32)
    gk, gv = run_sort(prim, keys, vals, direction)
    assert (gk == golden_small["sort_%s_dir%d_msh_keys" % (name, direction)]).all()
    if name in ("d1", "d2"):  # bit-exact vs the reference's mergeSortHost where it is stable
        assert (gv == golden_small["sort_%s_dir%d_msh_vals" % (name, direction)]).all()
    sk, sv = oracle.stable_sort_kv(keys, vals, direction)
    assert (gk == sk).all() and (gv == sv).all()


@pytest.mark.parametrize("n", [0, 1, 2, 33, 1000, 1024, 8191, 8192, 8193, 100000, (1 << 20) + 77])
@pytest.mark.parametrize("direction", [1, 0])
def test_sort_sizes(prim, n, direction):
    rng = np.random.default_rng(n + direction)
    for keys in (rng.integers(0, 2**32, n, dtype=np.uint32), rng.integers(0, 7, n, dtype=np.uint32),
                 np.full(n, 0xFFFFFFFF, np.uint32), np.zeros(n, np.uint32)):
        vals = rng.integers(0, 2**32, n, dtype=np.uint32)
        gk, gv = run_sort(prim, keys, vals, direction)
        sk, sv = oracle.stable_sort_kv(keys, vals, direction)
        assert (gk == sk).all() and (gv == sv).all()


def test_sort_kat_sample_input(prim, kat):
    """mergeSort/main.cpp:46-71: N = 4 Mi, srand(2009); rand()%65536, val = i, ascending."""
    m = kat["mergeSort"]
    keys = oracle.sample_input("sort_keys", m["n"], m["seed"])
    vals = oracle.fill_values(m["n"])
    prim.initMergeSort()
    gk, gv = run_sort(prim, keys, vals, m["dir"])
    assert hx(gk) == m["stable"]["fnv_keys"] and hx(gv) == m["stable"]["fnv_vals"]
    # the sample's own acceptance test
    assert oracle.validate_sorted_keys(gk, keys, 65536, 1) == 1
    assert oracle.validate_sorted_values(gk, gv, keys) == (1, 1)
    if oracle.ref is not None:
        assert oracle.ref.validate_sorted_keys(gk, keys, 65536, 1) == 1
        assert oracle.ref.validate_sorted_values(gk, gv, keys) == 1


def test_sort_full_size_matches_mergeSortHost_hashes(prim):
    """BASELINE.json configs[4] at its full 2^28 pairs against the reference's own mergeSortHost
    (mergeSort_host.cpp:288-363), run once in the build container on the same counter-generated input (all keys
    distinct, where mergeSortHost is the stable order; 321 s on one core) and committed as FNV-1a-64 hashes of its
    output arrays: tests/golden/make_golden_sort_2p28.py -> tests/golden/kat_sort_2p28.json."""
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat_sort_2p28.json")) as f:
        m = json.load(f)
    n = m["n"]
    assert n == 1 << 28 and m["generated_by"] == "reference mergeSortHost"
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(sk, m["kind"], 0, n, m["seed"])
    sv = torch.arange(n, dtype=torch.int64, device="cuda").to(torch.int32)
    dk, dv, bk, bv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4))
    prim.mergeSort(dk, dv, bk, bv, sk, sv, n, m["dir"])
    gk, gv = dk.cpu().numpy().view(np.uint32), dv.cpu().numpy().view(np.uint32)
    assert [int(gk[0]), int(gv[0])] == m["first"] and [int(gk[-1]), int(gv[-1])] == m["last"]
    assert hx(gk) == m["fnv_keys"]
    assert hx(gv) == m["fnv_vals"]


@pytest.mark.parametrize("dataset", ["d1", "d2"])
def test_sort_matches_mergeSortHost_where_it_is_stable(prim, dataset):
    n = 1 << 20
    if dataset == "d1":
        keys = oracle.generate("key_unique", 0, n, 5)
    else:
        keys = oracle.generate("key_tile_unique", 0, n, 0, 1 << 12)  # each key 256 times, unique per tile
    vals = oracle.fill_values(n)
    for direction in (1, 0):
        gk, gv = run_sort(prim, keys, vals, direction)
        mk, mv = oracle.merge_sort_host(keys, vals, direction)
        assert (gk == mk).all() and (gv == mv).all()


def test_sort_full_size_properties(prim):
    """BASELINE config 5 on one GPU: 2^28 pairs; checked through size-independent properties on the GPU
    (sortedness, key/value pairing, stability, multiset through the value permutation)."""
    n = 1 << 28
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    sv = torch.arange(n, dtype=torch.int32, device="cuda")
    dk, dv, bk, bv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4))
    for kind, param, direction in (("key_unique", 0, 1), ("key_tile_unique", 1 << 20, 1), ("key_mod", 65536, 0)):
        prim.generate(sk, kind, 0, n, 2009, param)
        prim.mergeSort(dk, dv, bk, bv, sk, sv, n, direction)
        ku = dk.to(torch.int64) & 0xFFFFFFFF
        ordered = (ku[1:] >= ku[:-1]) if direction else (ku[1:] <= ku[:-1])
        assert bool(ordered.all())
        ties = ku[1:] == ku[:-1]
        assert bool((dv[1:][ties] > dv[:-1][ties]).all())          # stable: ties keep source order
        assert bool((sk[dv.to(torch.int64)] == dk).all())           # pairing: resKey[j] == srcKey[resVal[j]]
        seen = torch.zeros(n, dtype=torch.bool, device="cuda")
        seen[dv.to(torch.int64)] = True
        assert bool(seen.all())                                    # values are a permutation of 0..n-1
        del ku, ordered, ties, seen


def test_sort_dropin_symbols(prim):
    from cuda_samples_b200._lib import lib
    n = 1 << 18
    keys = oracle.sample_input("sort_keys", n, 2009)
    vals = oracle.fill_values(n)
    t = [dev(as_i32(keys)), dev(as_i32(vals))] + [torch.zeros(n, dtype=torch.int32).cuda() for _ in range(4)]
    sk, sv, bk, bv, dk, dv = [ctypes.c_void_p(x.data_ptr()) for x in t]
    torch.cuda.synchronize()
    lib.initMergeSort()
    lib.mergeSort(dk, dv, bk, bv, sk, sv, n, 1)
    torch.cuda.synchronize()
    wk, wv = oracle.stable_sort_kv(keys, vals, 1)
    assert (t[4].cpu().numpy().view(np.uint32) == wk).all() and (t[5].cpu().numpy().view(np.uint32) == wv).all()


@pytest.mark.parametrize("buckets", [1, 2, 8, 256])
@pytest.mark.parametrize("direction", [1, 0])
def test_partition_kv(prim, buckets, direction):
    rng = np.random.default_rng(buckets)
    n = 300000
    keys = rng.integers(0, 2**32, n, dtype=np.uint32)
    keys[::7] = keys[0]
    vals = np.arange(n, dtype=np.uint32)
    flip = np.uint32(0 if direction else 0xFFFFFFFF)
    spl = np.sort(rng.integers(0, 2**32, buckets - 1, dtype=np.uint32))
    if buckets > 2:
        spl[1] = spl[0]  # an empty bucket
    dk, dv = torch.empty(n, dtype=torch.int32).cuda(), torch.empty(n, dtype=torch.int32).cuda()
    counts = torch.full((buckets,), -1, dtype=torch.int32).cuda()
    d_spl = dev(as_i32(spl)) if buckets > 1 else torch.zeros(1, dtype=torch.int32).cuda()
    prim.partition_kv(dk, dv, dev(as_i32(keys)), dev(as_i32(vals)), n, d_spl, buckets, direction, counts)
    b = np.searchsorted(spl, keys ^ flip, side="right")
    order = np.argsort(b, kind="stable")
    assert (counts.cpu().numpy() == np.bincount(b, minlength=buckets)).all()
    assert (dk.cpu().numpy().view(np.uint32) == keys[order]).all()
    assert (dv.cpu().numpy().view(np.uint32) == vals[order]).all()


# ================================================================================== generator
@pytest.mark.parametrize("kind,param,dtype", [("bytes", 0, torch.uint8), ("u31", 0, torch.int32),
                                              ("i32_byte", 0, torch.int32), ("f32_byte", 0, torch.float32),
                                              ("key_unique", 0, torch.int32), ("key_tile_unique", 4096, torch.int32),
                                              ("key_mod", 65536, torch.int32), ("f64_byte", 0, torch.float64)])
def test_generator_matches_cpu_twin(prim, kind, param, dtype):
    n, first = 100003, 12345678901
    d = torch.empty(n, dtype=dtype, device="cuda")
    prim.generate(d, kind, first, n, 2009, param)
    want = oracle.generate(kind, first, n, 2009, param)
    assert d.cpu().numpy().tobytes() == want.tobytes()


# ====================================================================== host-buffer (e2e) variants
def test_host_variants(prim):
    rng = np.random.default_rng(77)
    n = (80 << 20) + 1234  # > 2 staging chunks
    hb = rng.integers(0, 256, n, dtype=np.uint8)
    assert (prim.histogram256_host(hb) == oracle.histogram256(hb)).all()
    assert (prim.histogram64_host(hb) == oracle.histogram64(hb)).all()
    xi = rng.integers(-2**31, 2**31, n // 4, dtype=np.int64).astype(np.int32)
    assert prim.reduce_host(xi) == oracle.reduce(xi)
    xf = rng.random(n // 4, dtype=np.float32)
    want = oracle.reduce(xf)
    assert abs(prim.reduce_host(xf) - want) <= FP32_RTOL * abs(want)
    xd = rng.random(n // 8, dtype=np.float64)
    want = oracle.reduce(xd)
    assert abs(prim.reduce_host(xd) - want) <= FP64_RTOL * abs(want)
    src = rng.integers(0, 2**32, n // 4, dtype=np.uint32)
    dst = np.empty_like(src)
    for batch, L in ((1, src.size), (src.size // 1024, 1024), (2, src.size // 2)):
        m = batch * L
        prim.scan_exclusive_host(dst, src, batch, L)
        assert (dst[:m] == oracle.scan_exclusive(src[:m], batch, L)).all(), (batch, L)
    keys = rng.integers(0, 1000, 1 << 21, dtype=np.uint32)
    vals = np.arange(keys.size, dtype=np.uint32)
    ok, ov = np.empty_like(keys), np.empty_like(vals)
    for direction in (1, 0):
        prim.sort_kv_host(ok, ov, keys, vals, keys.size, direction)
        wk, wv = oracle.stable_sort_kv(keys, vals, direction)
        assert (ok == wk).all() and (ov == wv).all()


def test_host_variants_pinned_zero_copy_and_staged(prim):
    """Pinned input (b200_host_alloc) runs zero-copy -- one kernel reads the host buffer over PCIe; the same buffer
    forced through the staged path (several chunk sizes, ragged tail) gives the same bits."""
    rng = np.random.default_rng(78)
    n = (9 << 20) + 77
    hb = prim.HostBuffer(4 * n)
    info = hb.info()
    assert info["placement"] is not None
    xi = hb.array(np.int32)[:n]
    xi[:] = rng.integers(-2**31, 2**31, n, dtype=np.int64).astype(np.int32)
    want_i = oracle.reduce(xi)
    xf_src = rng.random(n, dtype=np.float32)
    try:
        for mode, chunk in ((1, 0), (0, 1 << 20), (0, 8 << 20), (0, 64 << 20)):
            prim.tune_host_path(mode, chunk)
            assert prim.reduce_host(xi) == want_i, (mode, chunk)
            assert (prim.histogram256_host(xi) == oracle.histogram256(xi)).all(), (mode, chunk)
            assert (prim.histogram64_host(xi) == oracle.histogram64(xi)).all(), (mode, chunk)
        hf = hb.array(np.float32)[:n]
        hf[:] = xf_src
        want_f = oracle.reduce(hf)
        for mode, chunk in ((1, 0), (0, 1 << 20), (0, 32 << 20)):
            prim.tune_host_path(mode, chunk)
            assert abs(prim.reduce_host(hf) - want_f) <= FP32_RTOL * abs(want_f), (mode, chunk)
        assert prim.host_probe_h2d(hf, 4 * n, 2) > 1.0  # GB/s: the probe ran
    finally:
        prim.tune_host_path(0, 64 << 20)
        hb.free()


def test_reduce_in_place_like_threadFenceReduction(prim):
    """threadFenceReduction.cu:247 calls reduce(s, threads, blocks, d_odata, d_odata): output aliases the input."""
    rng = np.random.default_rng(79)
    for n, blocks in ((64, 64), (1000, 1), (1 << 16, 64)):
        x = rng.random(n, dtype=np.float32)
        d = torch.from_numpy(x.copy()).cuda()
        prim.reduce(n, 256, blocks, 7, d, d)
        got = d.cpu().numpy()
        want = oracle.reduce(x)
        assert abs(float(got[0]) - want) <= FP32_RTOL * abs(want)
        assert (got[1:blocks] == 0).all() and (got[blocks:] == x[blocks:]).all()
        xi = rng.integers(-1000, 1000, n, dtype=np.int32)
        di = torch.from_numpy(xi.copy()).cuda()
        prim.reduce(n, 256, blocks, 7, di, di)
        assert int(di[0]) == oracle.reduce(xi)


# ============================================================== unmodified sample drivers (drop-in)
@pytest.mark.parametrize("name,args", [("histogram", []), ("scan", []), ("mergeSort", []), ("sortingNetworks", []),
                                       ("reduction", ["--type=int"]), ("reduction", ["--type=float"]),
                                       ("reduction", ["--type=double"])])
def test_unmodified_sample_driver_links_and_passes(prim, name, args):
    """baseline/_ref/<sample>_b200 = the reference's own main.cpp + gold files, linked against
    libb200prim.so instead of the sample's .cu (baseline/build.py).  The driver makes its own input,
    calls the entry points, runs its own gold and exits 0 only when they agree."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref", name + "_b200")
    if not os.path.exists(exe):
        pytest.skip("baseline/_ref not built (needs /root/reference at build time)")
    r = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    if name == "reduction":
        assert "Test passed" in r.stdout
    if name == "mergeSort":
        assert "stable!" in r.stdout


# ============================================= reduction siblings: same kernels
@pytest.mark.parametrize("symbol", ["reduce", "reduceSinglePass", "reduceSinglePassMultiBlockCG"])
def test_single_pass_reduction_siblings(prim, symbol):
    """threadFenceReduction / reductionMultiBlockCG entry points: float, tolerance of the samples
    (|gpu - cpu| < 1e-8 * size, threadFenceReduction.cu:479-481) AND the north-star relative 1e-5."""
    from cuda_samples_b200._lib import lib
    n = 1 << 20
    x = oracle.sample_input("reduce_f32", n, 1)
    d, out = dev(x), torch.full((64,), 3.0, dtype=torch.float32).cuda()
    torch.cuda.synchronize()
    getattr(lib, symbol)(n, 128, 64, ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(out.data_ptr()))
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    want = oracle.reduce(x)
    assert abs(float(o[0]) - want) < 1e-8 * n and abs(float(o[0]) - want) <= FP32_RTOL * abs(want)
    assert not o[1:].any()
    # the samples' multi-pass callers reduce d_odata in place afterwards (threadFenceReduction.cu:247)
    getattr(lib, symbol)(64, 32, 1, ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(out.data_ptr()))
    torch.cuda.synchronize()
    assert abs(float(out[0]) - want) <= FP32_RTOL * abs(want)


@pytest.mark.parametrize("vn,en", [(256, 4096), (1, 1), (3, 7), (5, 100003), (7, 4100), (2, 1 << 22), (1, (1 << 24) + 4),
                                   (4096, 64), (0, 16)])
def test_scalar_prod(prim, vn, en):
    """scalarProd sample (256 x 4096 is its own config): compared with scalarProdCPU's restatement.
    Products are rounded to float and summed in double on both sides, so the results agree to the last
    float bit except where the double sums differ in their last bits (different summation order)."""
    rng = np.random.default_rng(vn * 131 + en)
    a = rng.random(max(vn * en, 1), dtype=np.float32)[:vn * en]
    b = (rng.random(max(vn * en, 1), dtype=np.float32)[:vn * en] - 0.25).astype(np.float32)
    c = torch.full((max(vn, 1),), -7.0, dtype=torch.float32).cuda()
    prim.scalarProd(c, dev(a) if vn else c, dev(b) if vn else c, vn, en)
    if vn == 0:
        return
    got, want = c.cpu().numpy()[:vn], oracle.scalar_prod(a, b, vn, en)
    scale = np.maximum(np.abs(want), 1e-30)
    assert (np.abs(got - want) <= 2.5e-7 * np.maximum(scale, np.sqrt(en) * 1e-3)).all()
    # the sample's own acceptance test: L1 error < 1e-6 (scalarProd.cu:143-167)
    if np.abs(want).sum() > 0:
        assert np.abs(got - want).sum() / np.abs(want).sum() < 1e-6
    assert (got == want).mean() > 0.9 or vn < 8


# ================================================================== CUDA graph capture / replay
def test_a_kernel_that_overlaps_its_predecessor_never_reads_what_the_predecessor_is_still_writing(prim):
    """reduce, histogram and the tile-sums pass release the next kernel on the stream at their very top, and the
    histogram / the scan with a carry read their input before they wait for the kernel in front.  When that input IS
    what the kernel in front writes (a histogram of the buffer a reduce zero-fills, a scan of fresh histogram bins),
    the launcher must fall back to an ordinary, fully ordered launch (internal.h: programmatic-launch guard)."""
    n_out = 8 << 20  # reduce leaves [sum, 0, 0, ...] in 32 MiB: the zero-fill takes the last CTA tens of microseconds
    x = torch.ones(1 << 20, dtype=torch.int32, device="cuda")
    for _ in range(5):
        buf = torch.full((n_out,), -1, dtype=torch.int32, device="cuda")  # bytes 0xff everywhere
        h = torch.empty(256, dtype=torch.int32, device="cuda")
        prim.reduce(x.numel(), 256, n_out, 6, x, buf)
        prim.histogram256(h, buf, 4 * n_out)  # histogram of the reduce's output buffer, same stream
        got = h.cpu().numpy().view(np.uint32)
        want = np.zeros(256, np.uint32)
        for b in np.array([1 << 20], np.uint32).view(np.uint8):  # bytes of the sum
            want[b] += 1
        want[0] += 4 * n_out - 4
        assert (got == want).all(), (int(got[0]), int(got[255]))
    # scan (with a carry, the variant that is launched programmatically) of bins a histogram has just written
    data = torch.from_numpy(np.random.default_rng(3).integers(0, 256, 1 << 26, dtype=np.uint8)).cuda()
    carry = torch.tensor([5], dtype=torch.int32, device="cuda")
    reps = 1 << 14  # 256 bins repeated: one array of 2^22 elements, long enough for the TMA-staged chained scan
    for _ in range(5):
        bins = torch.full((256 * reps,), 77, dtype=torch.int32, device="cuda")
        out = torch.empty_like(bins)
        prim.histogram256(bins[-256:], data, data.numel())  # overwrites the LAST 256 words: the scan's tail depends on them
        prim.scanExclusiveLarge(out, bins, 1, bins.numel(), d_carry=carry)
        hb = bins.cpu().numpy().view(np.uint32)
        want = (np.concatenate(([0], np.cumsum(hb[:-1], dtype=np.uint32))) + np.uint32(5)).astype(np.uint32)
        assert (out.cpu().numpy().view(np.uint32) == want).all()


def test_entry_points_capture_into_a_cuda_graph_and_replay(prim):
    """Every *_async entry point only enqueues work on the given stream, and the kernels re-arm their own
    tickets / epochs on the device, so a captured sequence can be replayed (the inputs change between replays)."""
    rng = np.random.default_rng(99)
    n = 1 << 20
    data = torch.empty(n, dtype=torch.uint8, device="cuda")
    xi = torch.empty(n, dtype=torch.int32, device="cuda")
    keys = torch.empty(n, dtype=torch.int32, device="cuda")
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    h256, h64 = torch.empty(256, dtype=torch.int32, device="cuda"), torch.empty(64, dtype=torch.int32, device="cuda")
    red = torch.empty(4, dtype=torch.int32, device="cuda")
    scan_out = torch.empty(n, dtype=torch.int32, device="cuda")
    dk, dv, bk, bv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4))
    ii = torch.empty(n, dtype=torch.int32, device="cuda")
    comp, ncomp = torch.empty(n, dtype=torch.int32, device="cuda"), torch.empty(1, dtype=torch.int32, device="cuda")

    def enqueue():
        prim.integral_image(ii, data, 2048, n // 2048)   # 4 launches incl. the library's own batched scan
        prim.filter_arr(comp, ncomp, xi, n)
        prim.histogram256(h256, data, n)
        prim.histogram64(h64, data, n)
        prim.reduce(n, 256, 4, 7, xi, red)
        prim.scanExclusiveLarge(scan_out, xi, 1, n)   # chained scan (tickets + epoch)
        prim.mergeSort(dk, dv, bk, bv, keys, vals, n, 1)

    def fill(seed):
        r = np.random.default_rng(seed)
        hb = r.integers(0, 256, n, dtype=np.uint8)
        hx_ = (r.integers(0, 2**31, n, dtype=np.int64) - (1 << 29)).astype(np.int32)
        hk = r.integers(0, 2**32, n, dtype=np.uint32)
        data.copy_(torch.from_numpy(hb))
        xi.copy_(torch.from_numpy(hx_))
        keys.copy_(torch.from_numpy(hk.view(np.int32)))
        return hb, hx_, hk

    def check(hb, hx_, hk):
        torch.cuda.synchronize()
        assert (h256.cpu().numpy().view(np.uint32) == oracle.histogram256(hb)).all()
        assert (h64.cpu().numpy().view(np.uint32) == oracle.histogram64(hb)).all()
        r = red.cpu().numpy()
        assert int(r[0]) == oracle.reduce(hx_) and not r[1:].any()
        assert (scan_out.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(hx_.view(np.uint32), 1, n)).all()
        wk, wv = oracle.stable_sort_kv(hk, np.arange(n, dtype=np.uint32), 1)
        assert (dk.cpu().numpy().view(np.uint32) == wk).all() and (dv.cpu().numpy().view(np.uint32) == wv).all()
        assert (ii.cpu().numpy().view(np.uint32).reshape(-1, 2048) == oracle.integral_image(hb.reshape(-1, 2048))).all()
        want = oracle.filter_positive(hx_)
        assert int(ncomp.item()) == want.size and (comp[:want.size].cpu().numpy() == want).all()

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        inputs = fill(1)
        enqueue()  # warm-up outside capture: scratch is sized, kernel attributes are set
        check(*inputs)
        g = torch.cuda.CUDAGraph()
        inputs = fill(2)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=side):
            enqueue()
        for seed in (3, 4, 5):
            inputs = fill(seed)
            g.replay()
            check(*inputs)


# ================================================= sortingNetworks API: segmented sort
@pytest.mark.parametrize("batch,L", [(1, 1 << 16), (2, 1 << 15), (256, 256), (257, 64), (16384, 64), (1 << 15, 2),
                                     (7, 1000), (300, 333), (1, 5)])
@pytest.mark.parametrize("direction", [1, 0])
def test_batched_sort_sorting_networks_api(prim, batch, L, direction):
    """bitonicSort / oddEvenMergeSort contract: every array of the batch sorted on its own; checked against the
    stable oracle per array and with the sample's own acceptance rules (keys ordered, resKey[j] == srcKey[resVal[j]])."""
    rng = np.random.default_rng(batch * 7 + L + direction)
    n = batch * L
    keys = rng.integers(0, 65536, n, dtype=np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    dk, dv = torch.full((n,), -1, dtype=torch.int32).cuda(), torch.full((n,), -1, dtype=torch.int32).cuda()
    sk, sv = dev(as_i32(keys)), dev(as_i32(vals))
    assert prim.bitonicSort(dk, dv, sk, sv, batch, L, direction) > 0
    gk, gv = dk.cpu().numpy().view(np.uint32), dv.cpu().numpy().view(np.uint32)
    assert (sk.cpu().numpy().view(np.uint32) == keys).all()
    for a in list(range(min(batch, 3))) + [batch - 1, batch // 2]:
        wk, wv = oracle.stable_sort_kv(keys[a * L:(a + 1) * L], vals[a * L:(a + 1) * L], direction)
        assert (gk[a * L:(a + 1) * L] == wk).all() and (gv[a * L:(a + 1) * L] == wv).all()
    # whole batch: ordered inside every array, pairing intact (validateValues: vals are global indices)
    g2 = gk.reshape(batch, L).astype(np.int64)
    if L > 1:
        d = np.diff(g2, axis=1)
        assert (d >= 0).all() if direction else (d <= 0).all()
    assert (keys[gv] == gk).all()
    assert (np.sort(gv.reshape(batch, L), axis=1) == vals.reshape(batch, L)).all()


@pytest.mark.parametrize("batch,L", [(3, 4096), (5, 3000), (1000, 1), (700, 3), (7, 4097), (300, 17), (1, 4096),
                                     (255, 16), (256, 16), (9, 2048), (2, 2049)])
@pytest.mark.parametrize("keys_kind", ["u16", "u32", "constant", "high_byte"])
def test_batched_sort_in_shared_memory_shapes(prim, batch, L, keys_kind):
    """Arrays of <= 4096 pairs are sorted inside one CTA's shared memory (seg_sort_kernel): ragged lengths (padding),
    tiles with 1 .. 255 arrays, keys whose bytes do or do not vary (passes are skipped per tile), full 32-bit keys,
    both directions; 4097 / 8192 take the composition path -- same contract either side of the threshold."""
    rng = np.random.default_rng(batch * 131 + L)
    n = batch * L
    keys = {"u16": lambda: rng.integers(0, 65536, n, dtype=np.uint32),
            "u32": lambda: rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32),
            "constant": lambda: np.full(n, 0xDEADBEEF, np.uint32),
            "high_byte": lambda: (rng.integers(0, 7, n, dtype=np.uint32) << 24) | 0x00ABCD12}[keys_kind]()
    vals = np.arange(n, dtype=np.uint32)
    sk, sv = dev(as_i32(keys)), dev(as_i32(vals))
    for direction in (1, 0):
        dk, dv = torch.full((n,), -1, dtype=torch.int32).cuda(), torch.full((n,), -1, dtype=torch.int32).cuda()
        prim.bitonicSort(dk, dv, sk, sv, batch, L, direction)
        gk, gv = dk.cpu().numpy().view(np.uint32), dv.cpu().numpy().view(np.uint32)
        order = np.argsort(keys.reshape(batch, L).astype(np.int64) * (1 if direction else -1), axis=1, kind="stable")
        wk = np.take_along_axis(keys.reshape(batch, L), order, axis=1).reshape(-1)
        wv = np.take_along_axis(vals.reshape(batch, L), order, axis=1).reshape(-1)
        assert (gk == wk).all() and (gv == wv).all(), (batch, L, keys_kind, direction)


# ============================================================ peer-window exchange (world = 1) and dyn-count sort
def test_sort_with_device_side_count(prim):
    """b200_sort_kv_dyn_async: the pair count is read on the device; capacity only sizes the launch."""
    rng = np.random.default_rng(11)
    cap = (1 << 18) + 4096
    keys = rng.integers(0, 1 << 32, cap, dtype=np.uint64).astype(np.uint32)
    vals = np.arange(cap, dtype=np.uint32)
    sk, sv = dev(keys.view(np.int32)), dev(vals.view(np.int32))
    for n in (0, 1, 8191, 8192, 100000, cap):
        for direction in (1, 0):
            dk, dv, bk, bv = (torch.full((cap,), -7, dtype=torch.int32, device="cuda") for _ in range(4))
            d_n = torch.tensor([n, 0], dtype=torch.int32, device="cuda")
            prim.sort_kv_dyn(dk, dv, bk, bv, sk, sv, cap, d_n, direction)
            wk, wv = oracle.stable_sort_kv(keys[:n], vals[:n], direction)
            assert (dk[:n].cpu().numpy().view(np.uint32) == wk).all(), (n, direction)
            assert (dv[:n].cpu().numpy().view(np.uint32) == wv).all(), (n, direction)
            assert bool((dk[n:] == -7).all()), "nothing beyond the counted pairs may be written"


@pytest.mark.parametrize("kind", ["mod65536", "mod256", "const", "hi16", "byte1", "byte3", "bytes0+2", "u32"])
def test_sort_skips_single_digit_passes(prim, kind):
    """The pass plan made on the device (radix_sort.cu: make_sort_plan): a pass whose digit is the same for every key
    is not run, the remaining passes are re-chained over the pair buffers.  The sample's own keys (rand() % 65536,
    mergeSort/main.cpp:48,68) have two such passes.  Stable order bit-exact for every position of the varying bytes,
    both directions, split and pair-buffer input, host and device-side count, and with the plan switched off."""
    rng = np.random.default_rng(41)
    n = (1 << 17) + 8192 + 5  # (the planned pipeline starts at 65536 pairs)
    r = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    keys = {"mod65536": r & 0xFFFF, "mod256": r & 0xFF, "const": np.full(n, 0x12345678, np.uint32),
            "hi16": r & 0xFFFF0000, "byte1": r & 0xFF00, "byte3": r & 0xFF000000, "bytes0+2": r & 0x00FF00FF,
            "u32": r}[kind].astype(np.uint32)
    vals = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    pairs = np.empty(2 * n, np.uint32)
    pairs[0::2], pairs[1::2] = keys, vals
    sp = dev(pairs.view(np.int32))
    try:
        for plan in (1, 0):
            prim.lib.b200_tune_sort_plan(plan)
            for direction in (1, 0):
                wk, wv = oracle.stable_sort_kv(keys, vals, direction)
                gk, gv = run_sort(prim, keys, vals, direction)
                assert (gk == wk).all() and (gv == wv).all(), (kind, plan, direction, "split input")
                m = n - 4097  # device-side count below the capacity, pair-buffer input
                dk, dv = (torch.full((n,), -7, dtype=torch.int32, device="cuda") for _ in range(2))
                d_n = torch.tensor([m, 0], dtype=torch.int32, device="cuda")
                prim.sort_kv_dyn(dk, dv, None, None, sp, None, n, d_n, direction)
                wk, wv = oracle.stable_sort_kv(keys[:m], vals[:m], direction)
                assert (dk[:m].cpu().numpy().view(np.uint32) == wk).all(), (kind, plan, direction, "pairs, dyn count")
                assert (dv[:m].cpu().numpy().view(np.uint32) == wv).all(), (kind, plan, direction, "pairs, dyn count")
                assert bool((dk[m:] == -7).all()) and bool((dv[m:] == -7).all())
    finally:
        prim.lib.b200_tune_sort_plan(1)
    assert (sp.cpu().numpy().view(np.uint32) == pairs).all()


def test_sort_from_a_pair_buffer(prim):
    """d_srcVal == NULL: the input is an array of interleaved (key, value) pairs -- the format of the sharded sort's
    receive buffers (one 8-byte peer store per pair).  Every pass then reads pairs; the result is the usual split
    dst arrays, the count may come from the device, and the pair buffer itself stays untouched."""
    rng = np.random.default_rng(23)
    cap = (1 << 17) + 8192 + 6
    keys = rng.integers(0, 1 << 32, cap, dtype=np.uint64).astype(np.uint32)
    keys[: cap // 2] &= 0xFFFF  # ties: stability matters
    vals = rng.integers(0, 1 << 32, cap, dtype=np.uint64).astype(np.uint32)
    pairs = np.empty(2 * cap, np.uint32)
    pairs[0::2], pairs[1::2] = keys, vals
    sp = dev(pairs.view(np.int32))
    for n in (0, 1, 2, 3, 8191, 8193, 70001, cap):
        for direction in (1, 0):
            dk, dv = (torch.full((cap,), -7, dtype=torch.int32, device="cuda") for _ in range(2))
            d_n = torch.tensor([n, 0], dtype=torch.int32, device="cuda")
            prim.sort_kv_dyn(dk, dv, None, None, sp, None, cap, d_n, direction)
            wk, wv = oracle.stable_sort_kv(keys[:n], vals[:n], direction)
            assert (dk[:n].cpu().numpy().view(np.uint32) == wk).all(), (n, direction)
            assert (dv[:n].cpu().numpy().view(np.uint32) == wv).all(), (n, direction)
            assert bool((dk[n:] == -7).all()) and bool((dv[n:] == -7).all())
    assert (sp.cpu().numpy().view(np.uint32) == pairs).all()
    # host-known count through the plain entry point
    n = 100003
    dk, dv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(2))
    prim.mergeSort(dk, dv, None, None, sp, None, n, 1)
    wk, wv = oracle.stable_sort_kv(keys[:n], vals[:n], 1)
    assert (dk.cpu().numpy().view(np.uint32) == wk).all() and (dv.cpu().numpy().view(np.uint32) == wv).all()


def test_peer_window_entry_points_world_1(prim):
    """The fused compute + exchange kernels with a one-rank window set: same results as the plain kernels
    (the 2/4/8-GPU case is tests/multi_gpu_check.py)."""
    h = prim.p2p_window_create()
    prim.p2p_connect(0, 1, [h])
    try:
        assert prim.p2p_world() == 1
        rng = np.random.default_rng(5)
        pay = dev(rng.integers(0, 1 << 31, 24, dtype=np.int64).astype(np.int32))
        got = torch.empty(24, dtype=torch.int32, device="cuda")
        for _ in range(3):
            prim.p2p_allgather(pay, 96, got)
        assert bool((got == pay).all())
        # histogram -> all-reduce
        data = rng.integers(0, 256, (1 << 22) + 48, dtype=np.uint8)
        o256 = torch.empty(256, dtype=torch.int32, device="cuda")
        o64 = torch.empty(64, dtype=torch.int32, device="cuda")
        for _ in range(2):
            prim.histogram_fused_allreduce(o256, o64, dev(data), data.size)
            assert (o256.cpu().numpy().view(np.uint32) == oracle.histogram256(data)).all()
            assert (o64.cpu().numpy().view(np.uint32) == oracle.histogram64(data)).all()
        # reduce -> exchange -> rank-order sum (two slots: int then float)
        xi = rng.integers(-2**31, 2**31, 1 << 20, dtype=np.int64).astype(np.int32)
        xf = rng.random(1 << 20, dtype=np.float32)
        slots = torch.zeros(2, dtype=torch.float64, device="cuda")
        gathered = torch.zeros(2, dtype=torch.float64, device="cuda")
        sums = torch.zeros(2, dtype=torch.float64, device="cuda")
        prim.reduce_partial(dev(xi), xi.size, slots[0:1])
        prim.reduce_partial_exchange(dev(xf), xf.size, slots, 1, 2, gathered, sums)
        s = sums.cpu().numpy()
        assert int(s[0]) == oracle.reduce(xi)
        want = oracle.reduce(xf)
        assert abs(float(s[1]) - want) <= FP32_RTOL * abs(want)
        assert (gathered.cpu().numpy() == s).all()
        # scan total -> carry (rank 0: no lower ranks)
        src = rng.integers(0, 1 << 32, 1 << 20, dtype=np.uint64).astype(np.uint32)
        carry = torch.full((1,), 99, dtype=torch.int32, device="cuda")
        work = torch.zeros(18, dtype=torch.float64, device="cuda")
        prim.scan_total_exchange(carry, dev(src.view(np.int32)), src.size, work)
        assert int(carry.item()) == 0
        total = int(work[0].item()) & 0xFFFFFFFF
        assert total == int(src.sum(dtype=np.uint64) & 0xFFFFFFFF)
        # one shard of a sharded scan, both steps: tile sums + exchange + tile prefixes, then independent tiles;
        # ragged / unaligned / short shards take the chained scan behind the total exchange; in place as well
        big = rng.integers(0, 1 << 32, (1 << 22) + 8192 * 3 + 4 * 77, dtype=np.uint64).astype(np.uint32)
        dbig = dev(big.view(np.int32))
        for tiled in (1, 0):
            prim.lib.b200_tune_scan_shard_tiled(tiled)
            for off, n in ((0, big.size), (0, 1 << 20), (0, 8192 * 4), (0, 8192 * 4 + 4), (0, 8192 * 3), (4, 1 << 21),
                           (1, (1 << 20) + 3), (0, 0), (0, 5)):
                x = big[off:off + n]
                out = torch.full((n + 8,), -5, dtype=torch.int32, device="cuda")
                prim.scan_exclusive_shard(out, dbig[off:off + n], n, work, carry)
                want = oracle.scan_exclusive(x, 1, n) if n else x
                assert (out[:n].cpu().numpy().view(np.uint32) == want).all(), (tiled, off, n)
                assert bool((out[n:] == -5).all()) and int(carry.item()) == 0
            inplace = dbig[:1 << 21].clone()
            prim.scan_exclusive_shard(inplace, inplace, 1 << 21, work, carry)
            assert (inplace.cpu().numpy().view(np.uint32) == oracle.scan_exclusive(big[:1 << 21], 1, 1 << 21)).all()
        prim.lib.b200_tune_scan_shard_tiled(1)
        torch.cuda.synchronize()
        assert prim.p2p_errors() == 0
    finally:
        prim.p2p_disconnect()
    assert prim.p2p_world() == 0


# ===================================================================== scan siblings
@pytest.mark.parametrize("n", [1, 5, 1024, 8191, 8192, 65536, (1 << 20) + 3, 1 << 24])
def test_scan_inclusive_matches_shfl_scan_gold(prim, n):
    rng = np.random.default_rng(n)
    src = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    s = dev(src.view(np.int32))
    d = torch.empty_like(s)
    prim.scanInclusive(d, s, 1, n)
    assert (d.cpu().numpy().view(np.uint32) == oracle.scan_inclusive(src)).all()
    if n == 65536:  # the sample's own input: all ones -> 1..n (shfl_scan.cu:246-248)
        ones = torch.ones(n, dtype=torch.int32, device="cuda")
        prim.scanInclusive(ones, ones, 1, n)  # in place like the sample
        assert (ones.cpu().numpy() == np.arange(1, n + 1)).all()


def test_scan_inclusive_batched(prim):
    rng = np.random.default_rng(3)
    for batch, L in ((4096, 256), (64, 16384), (3, 100000)):
        src = rng.integers(0, 1 << 32, batch * L, dtype=np.uint64).astype(np.uint32)
        d = torch.empty(batch * L, dtype=torch.int32, device="cuda")
        prim.scanInclusive(d, dev(src.view(np.int32)), batch, L)
        want = np.concatenate([oracle.scan_inclusive(src[a * L:(a + 1) * L]) for a in range(batch)])
        assert (d.cpu().numpy().view(np.uint32) == want).all(), (batch, L)


@pytest.mark.parametrize("n", [0, 1, 31, 8192, 8193, 100001, 1 << 22, (1 << 24) + 5])
@pytest.mark.parametrize("kind", ["sample", "signed", "none", "all"])
def test_compaction_matches_host_filter(prim, n, kind):
    rng = np.random.default_rng(n + len(kind))
    if kind == "sample":      # warpAggregatedAtomicsCG.cu:255-257: rand() % 5
        src = rng.integers(0, 5, n, dtype=np.int64).astype(np.int32)
    elif kind == "signed":
        src = rng.integers(-2**31, 2**31, n, dtype=np.int64).astype(np.int32)
    elif kind == "none":
        src = -rng.integers(0, 100, n, dtype=np.int64).astype(np.int32)
    else:
        src = rng.integers(1, 100, n, dtype=np.int64).astype(np.int32)
    s = dev(src) if n else torch.empty(0, dtype=torch.int32, device="cuda")
    dst = torch.full((max(n, 1),), -99, dtype=torch.int32, device="cuda")
    cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    prim.filter_arr(dst, cnt, s, n)
    want = oracle.filter_positive(src)
    assert int(cnt.item()) == want.size
    assert (dst[:want.size].cpu().numpy() == want).all()
    assert bool((dst[want.size:] == -99).all())


@pytest.mark.parametrize("shape", [(1080, 1920), (1, 1), (1, 4), (7, 128), (33, 132), (64, 4096), (100, 1000),
                                   (257, 515), (31, 3), (1000, 130), (2049, 2052), (4096, 8192)])
def test_integral_image_matches_oracle(prim, shape):
    h, w = shape
    rng = np.random.default_rng(h * 7919 + w)
    img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    if shape == (1080, 1920):  # the sample's input: every byte 1 (shfl_scan.cu:333), corner = w*h (:372-381)
        img[:] = 1
    d_img = torch.from_numpy(img).cuda()
    out = torch.full((h * w,), -1, dtype=torch.int32, device="cuda")
    prim.integral_image(out, d_img, w, h)
    got = out.cpu().numpy().view(np.uint32).reshape(h, w)
    assert (got == oracle.integral_image(img)).all()
    if shape == (1080, 1920):
        assert int(got[-1, -1]) == w * h and (got[0] == np.arange(1, w + 1)).all()


@pytest.mark.parametrize("rows", [8, 16, 32])
def test_integral_image_band_heights_and_wrap(prim, rows, monkeypatch):
    """All band heights give the same table; 255-valued 8192 x 4100 image wraps u32 exactly like the oracle."""
    import ctypes
    from cuda_samples_b200 import _lib
    knob = ctypes.c_int.in_dll(_lib.lib, "g_ii_rows")
    knob.value = rows
    try:
        h, w = 4100, 8192
        img = np.full((h, w), 255, np.uint8)
        img[::3, ::5] = 17
        out = torch.empty(h * w, dtype=torch.int32, device="cuda")
        prim.integral_image(out, torch.from_numpy(img).cuda(), w, h)
        assert (out.cpu().numpy().view(np.uint32).reshape(h, w) == oracle.integral_image(img)).all()
        img = np.random.default_rng(rows).integers(0, 256, (45, 77), dtype=np.uint8)   # unaligned variant
        out = torch.empty(45 * 77, dtype=torch.int32, device="cuda")
        prim.integral_image(out, torch.from_numpy(img).cuda(), 77, 45)
        assert (out.cpu().numpy().view(np.uint32).reshape(45, 77) == oracle.integral_image(img)).all()
    finally:
        knob.value = 0


def test_size_limits_are_status_codes_not_aborts(prim):
    """The extended ABI reports what it cannot do (documented limits in include/b200prim.h) as a status; nothing is
    launched and the output stays untouched.  Empty inputs are a no-op."""
    from cuda_samples_b200._lib import B200Error
    out = torch.full((16,), 7, dtype=torch.int32, device="cuda")
    img = torch.zeros(16, dtype=torch.uint8, device="cuda")
    with pytest.raises(B200Error):
        prim.integral_image(out, img, 1 << 16, 1 << 16)      # width * height = 2^32
    with pytest.raises(B200Error):
        prim.filter_arr(out, out, out, 1 << 30)              # n < 2^30
    with pytest.raises(B200Error):
        prim.mergeSort(out, out, out, out, out, out, 1 << 30, 1)
    prim.integral_image(out, img, 0, 5)
    prim.integral_image(out, img, 5, 0)
    torch.cuda.synchronize()
    assert bool((out == 7).all())
