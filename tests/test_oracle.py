"""Pin the oracle (CPU only).

The restatement in oracle/oracle.c must agree with
  * the committed golden vectors (tests/golden/, produced by the reference's own gold code), and
  * the reference gold compiled unmodified (oracle/_ref/libref_gold.so) whenever that file exists.
"""
import os

import numpy as np
import pytest

import oracle

HAVE_REF = oracle.ref is not None
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref/libref_gold.so not built")


def hx(a):
    return "%016x" % oracle.fnv1a64(a)


def test_fnv_known_answer():
    # FNV-1a 64 test vectors (draft-eastlake-fnv): "" and "a"
    assert oracle.fnv1a64(np.zeros(0, np.uint8)) == 0xcbf29ce484222325
    assert oracle.fnv1a64(np.frombuffer(b"a", np.uint8)) == 0xaf63dc4c8601ec8c


# ------------------------------------------------------------------------------- small fixtures
def test_histogram_small_golden(golden_small):
    g = golden_small
    assert (oracle.histogram256(g["hist_in"]) == g["hist256_out"]).all()
    assert (oracle.histogram64(g["hist_in"]) == g["hist64_out"]).all()
    # the identity the fused multi-GPU path relies on: hist64[v] = sum of hist256[4v..4v+3]
    assert (g["hist256_out"].reshape(64, 4).sum(1, dtype=np.uint32) == g["hist64_out"]).all()


@pytest.mark.parametrize("L", [4, 32, 1024, 2048, 8192])
def test_scan_small_golden(golden_small, L):
    src = golden_small["scan_in"]
    assert (oracle.scan_exclusive(src, src.size // L, L) == golden_small["scan_out_L%d" % L]).all()


@pytest.mark.parametrize("name", ["d1", "d2", "d3"])
@pytest.mark.parametrize("direction", [1, 0])
def test_merge_sort_host_small_golden(golden_small, name, direction):
    g = golden_small
    keys = g["sort_%s_keys" % name]
    vals = np.arange(keys.size, dtype=np.uint32)
    mk, mv = oracle.merge_sort_host(keys, vals, direction)
    assert (mk == g["sort_%s_dir%d_msh_keys" % (name, direction)]).all()
    assert (mv == g["sort_%s_dir%d_msh_vals" % (name, direction)]).all()
    sk, sv = oracle.stable_sort_kv(keys, vals, direction)
    assert (sk == mk).all()
    if name in ("d1", "d2"):
        # keys unique inside every 1024-aligned tile => mergeSortHost coincides with the stable sort
        assert (sv == mv).all()
    else:
        # duplicates inside tiles: the reference's CPU emulation is NOT stable
        assert (sv != mv).any()
    ok, stable = oracle.validate_sorted_values(sk, sv, keys)
    assert ok == 1 and stable == 1
    if name == "d3":  # validateSortedKeys needs keys < numValues (mergeSort_validate.cpp:50-70)
        assert oracle.validate_sorted_keys(sk, keys, 64, direction) == 1


def test_stable_sort_matches_numpy():
    rng = np.random.default_rng(5)
    for n in (0, 1, 2, 3, 1000, 4097):
        k = rng.integers(0, 17, n, dtype=np.uint32)
        v = np.arange(n, dtype=np.uint32)
        sk, sv = oracle.stable_sort_kv(k, v, 1)
        order = np.argsort(k, kind="stable")
        assert (sk == k[order]).all() and (sv == v[order]).all()
        sk, sv = oracle.stable_sort_kv(k, v, 0)
        order = np.argsort(~k, kind="stable")  # descending, ties in source order
        assert (sk == k[order]).all() and (sv == v[order]).all()


# ------------------------------------------------------------------- the samples' default inputs
def test_kat_histogram(kat):
    d = oracle.sample_input("hist_bytes", kat["histogram256"]["bytes"], kat["histogram256"]["seed"])
    h = oracle.histogram256(d)
    assert (int(h[0]), int(h[255]), hx(h)) == (kat["histogram256"]["h0"], kat["histogram256"]["h255"],
                                               kat["histogram256"]["fnv"])
    # SURVEY.md 8c spot values
    assert (int(h[0]), int(h[255])) == (262441, 262682)
    h = oracle.histogram64(d)
    assert (int(h[0]), int(h[63]), hx(h)) == (kat["histogram64"]["h0"], kat["histogram64"]["h63"],
                                              kat["histogram64"]["fnv"])
    assert (int(h[0]), int(h[63])) == (1048397, 1049943)


def test_kat_scan(kat):
    n = kat["scan"]["n"]
    a = oracle.sample_input("scan_u32", n, kat["scan"]["seed"])
    survey_last = {4: 1983869552, 1024: 3998548728, 2048: 2063854727, 262144: 972374877}
    for case in kat["scan"]["cases"]:
        L = case["L"]
        s = oracle.scan_exclusive(a, n // L, L)
        assert int(s[-1]) == case["last"] and hx(s) == case["fnv"]
        if L in survey_last:
            assert int(s[-1]) == survey_last[L]


def test_kat_reduce(kat):
    for case in kat["reduce"]["cases"]:
        n = case["n"]
        ri = oracle.sample_input("reduce_i32", n, 1)
        assert oracle.reduce(ri) == case["i32"] == int(ri.astype(np.int64).sum())
        rf = oracle.sample_input("reduce_f32", n, 1)
        got = oracle.reduce(rf)
        assert float(np.float32(got)).hex() == case["f32_hex"]
        # Kahan fp32 is within 1 ulp-ish of the exact (fp64) sum
        assert abs(got - float(rf.astype(np.float64).sum())) <= 2e-7 * abs(got)
        rd = oracle.sample_input("reduce_f64", n, 1)
        assert oracle.reduce(rd).hex() == case["f64_hex"]


@pytest.mark.skipif(oracle.ref_reduce is None, reason="oracle/_ref/libref_reduce.so not built")
def test_reduceCPU_restatement_equals_the_reference_template(kat):
    """reduceCPU<T> compiled from the sample's own driver TU (oracle/ref_reduce_wrap.cpp) == the restatement,
    bit for bit, on the sample's rand() inputs (the committed KAT values) and on random data incl. int wrap."""
    for case in kat["reduce"]["cases"]:
        n = case["n"]
        ri = oracle.sample_input("reduce_i32", n, 1)
        assert oracle.ref_reduce.reduce(ri) == oracle.reduce(ri) == case["i32"]
        rf = oracle.sample_input("reduce_f32", n, 1)
        assert float(np.float32(oracle.ref_reduce.reduce(rf))).hex() == case["f32_hex"]
        rd = oracle.sample_input("reduce_f64", n, 1)
        assert oracle.ref_reduce.reduce(rd).hex() == case["f64_hex"]
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 1000, 65537, 1 << 20):
        xi = rng.integers(-2**31, 2**31, n, dtype=np.int64).astype(np.int32)
        assert oracle.ref_reduce.reduce(xi) == oracle.reduce(xi)
        xf = (rng.random(n, dtype=np.float32) - np.float32(0.3)) * np.float32(1e3)
        assert np.float32(oracle.ref_reduce.reduce(xf)) == np.float32(oracle.reduce(xf))
        xd = rng.standard_normal(n)
        assert oracle.ref_reduce.reduce(xd) == oracle.reduce(xd)


def test_reduce_int_wraps():
    # reduceCPU<int> is a wrapping 32-bit sum (SURVEY.md 8c KAT-RED)
    x = np.full(1 << 16, 0x7FFF, np.int32)
    want = ((0x7FFF << 16) + 2**31) % 2**32 - 2**31
    assert oracle.reduce(x) == want


def test_kat_merge_sort(kat):
    m = kat["mergeSort"]
    k = oracle.sample_input("sort_keys", m["n"], m["seed"])
    v = oracle.fill_values(m["n"])
    sk, sv = oracle.stable_sort_kv(k, v, m["dir"])
    assert hx(sk) == m["stable"]["fnv_keys"] and hx(sv) == m["stable"]["fnv_vals"]
    mk, mv = oracle.merge_sort_host(k, v, m["dir"])
    assert hx(mk) == m["mergeSortHost"]["fnv_keys"] and hx(mv) == m["mergeSortHost"]["fnv_vals"]
    assert int((mv != sv).sum()) == m["mergeSortHost_vals_differing_from_stable"] == 21848
    assert oracle.validate_sorted_keys(sk, k, 65536, 1) == 1
    assert oracle.validate_sorted_values(sk, sv, k) == (1, 1)
    assert oracle.validate_sorted_values(mk, mv, k) == (1, 0)  # the reference's own emulation: NOT stable


# ---------------------------------------------------------- against the compiled reference gold
@needs_ref
def test_restatement_equals_reference_gold_random():
    rng = np.random.default_rng(11)
    for nbytes in (0, 4, 64, 4096, 1 << 20):
        d = rng.integers(0, 256, nbytes, dtype=np.uint8)
        assert (oracle.histogram256(d) == oracle.ref.histogram256(d)).all()
        assert (oracle.histogram64(d) == oracle.ref.histogram64(d)).all()
    for batch, L in ((1, 1), (3, 5), (7, 1024), (1, 1 << 20), (64, 4096)):
        s = rng.integers(0, 2**32, batch * L, dtype=np.uint32)
        assert (oracle.scan_exclusive(s, batch, L) == oracle.ref.scan_exclusive(s, batch, L)).all()
    for n in (1, 2, 1000, 1024, 1025, 3 * 1024, 5000, 65536 + 17):
        for direction in (1, 0):
            k = rng.integers(0, 300, n, dtype=np.uint32)
            v = np.arange(n, dtype=np.uint32)
            a = oracle.merge_sort_host(k, v, direction)
            b = oracle.ref.merge_sort_host(k, v, direction)
            assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
            assert oracle.validate_sorted_keys(a[0], k, 300, direction) == \
                oracle.ref.validate_sorted_keys(a[0], k, 300, direction) == 1
            assert oracle.validate_sorted_values(a[0], a[1], k)[0] == oracle.ref.validate_sorted_values(a[0], a[1], k)


def test_generator_is_counter_based():
    # any window of the stream can be regenerated on its own (what lets 8 GPUs fill shards in place)
    for kind, param in (("bytes", 0), ("u31", 0), ("i32_byte", 0), ("f32_byte", 0), ("key_unique", 0),
                        ("key_tile_unique", 2048), ("key_mod", 65536), ("f64_byte", 0)):
        whole = oracle.generate(kind, 0, 5000, 2009, param)
        part = oracle.generate(kind, 1237, 2000, 2009, param)
        assert (whole[1237:3237] == part).all()
    k = oracle.generate("key_unique", 0, 1 << 16, 3)
    assert np.unique(k).size == k.size
    k = oracle.generate("key_tile_unique", 0, 8192, 0, 2048)
    for t in range(8):
        assert np.unique(k[t * 1024:(t + 1) * 1024]).size == 1024


@needs_ref
def test_scalar_prod_restatement_equals_reference_gold():
    """scalarProd_cpu.cpp:32-46."""
    rng = np.random.default_rng(12)
    for vn, en in ((1, 1), (3, 7), (256, 4096), (5, 100003)):
        a = rng.random(vn * en, dtype=np.float32)
        b = (rng.random(vn * en, dtype=np.float32) - 0.5).astype(np.float32)
        assert (oracle.scalar_prod(a, b, vn, en) == oracle.ref.scalar_prod(a, b, vn, en)).all()


def test_scan_siblings_oracle():
    """Inclusive scan (shfl_scan CPUverify) and the ordered filter (warpAggregatedAtomicsCG host loop)."""
    import oracle
    rng = np.random.default_rng(1)
    src = rng.integers(0, 1 << 32, 10007, dtype=np.uint64).astype(np.uint32)
    assert (oracle.scan_inclusive(src) == np.cumsum(src.astype(np.uint64)).astype(np.uint32)).all()
    ex = oracle.scan_exclusive(src, 1, src.size)
    assert (oracle.scan_inclusive(src) == ex + src).all()
    x = rng.integers(-3, 4, 5000, dtype=np.int64).astype(np.int32)
    assert (oracle.filter_positive(x) == x[x > 0]).all()
    assert oracle.filter_positive(np.zeros(0, dtype=np.int32)).size == 0


def test_integral_image_oracle_matches_the_samples_checks():
    """shfl_scan's own acceptance checks on its own input (all-ones 1920x1080: row sums i+1, corner w*h,
    shfl_scan.cu:193-205, :372-381) and numpy's double cumsum on random images."""
    import oracle
    w, h = 1920, 1080
    out = oracle.integral_image(np.ones((h, w), np.uint8))
    assert (out[0] == np.arange(1, w + 1)).all() and int(out[-1, -1]) == w * h
    assert (out[:, 0] == np.arange(1, h + 1)).all()
    rng = np.random.default_rng(7)
    for hh, ww in ((1, 1), (3, 5), (37, 129), (200, 333)):
        img = rng.integers(0, 256, (hh, ww), dtype=np.uint8)
        want = np.cumsum(np.cumsum(img.astype(np.uint64), axis=1), axis=0).astype(np.uint32)
        assert (oracle.integral_image(img) == want).all()
    assert oracle.integral_image(np.zeros((0, 4), np.uint8)).size == 0


def test_full_size_mergeSortHost_fixture_is_what_its_script_writes():
    """tests/golden/kat_sort_2p28.json: hashes of the reference's mergeSortHost at 2^28 pairs (5 minutes of CPU, so it
    is not re-run here); its first and last pairs can be checked against the generator in a moment: with distinct
    keys the first output pair is the smallest key with its index."""
    import json
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat_sort_2p28.json")) as f:
        m = json.load(f)
    assert m["n"] == 1 << 28 and m["kind"] == "key_unique" and m["dir"] == 1
    assert len(m["fnv_keys"]) == 16 and len(m["fnv_vals"]) == 16
    for key, idx in (m["first"], m["last"]):
        assert int(oracle.generate("key_unique", idx, 1, m["seed"], 0)[0]) == key
