#!/usr/bin/env python3
"""Generate the committed golden vectors from the REFERENCE's own gold code.

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py
It compiles the reference gold sources unmodified (oracle/build.py -> oracle/_ref/libref_gold.so),
runs them on (a) the samples' own seeded default inputs and (b) small seeded cases, and writes

  tests/golden/kat.json      known answers on the samples' default inputs: spot values + FNV-1a-64
                             of the little-endian output bytes (hash function: oracle_fnv1a64)
  tests/golden/small.npz     small input/output pairs (a few KB each) for the CPU test-suite

Nothing here is read from /root/reference at test time; the fixtures travel with the repo.
The only function of the path with no linkable reference symbol is reduceCPU<T> (a template
inside the driver TU, reduction/reduction.cpp:167-180); its known answers come from the restated
Kahan loop in oracle/oracle.c and are marked "source": "restatement".
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def hx(a):
    return "%016x" % oracle.fnv1a64(a)


def main():
    ref = oracle.ref
    assert ref is not None, "reference gold not available: run inside the build container"
    kat = {"hash": "fnv1a64 over little-endian output bytes (oracle_fnv1a64)", "generated_by": "reference gold"}

    # ---- histogram: histogram/main.cpp:92-96 (srand(2009); rand()%256; 64 MiB) --------------
    d = oracle.sample_input("hist_bytes", 64 << 20, 2009)
    h256, h64 = ref.histogram256(d), ref.histogram64(d)
    kat["histogram256"] = {"bytes": 64 << 20, "seed": 2009, "h0": int(h256[0]), "h255": int(h256[255]),
                           "fnv": hx(h256)}
    kat["histogram64"] = {"bytes": 64 << 20, "seed": 2009, "h0": int(h64[0]), "h63": int(h64[63]), "fnv": hx(h64)}

    # ---- scan: scan/main.cpp:45-56 (N = 13*2^20/2, srand(2009); rand()) ----------------------
    n = 13 * 1048576 // 2
    a = oracle.sample_input("scan_u32", n, 2009)
    kat["scan"] = {"n": n, "seed": 2009, "cases": []}
    for L in (4, 64, 1024, 2048, 65536, 262144):
        s = ref.scan_exclusive(a, n // L, L)
        kat["scan"]["cases"].append({"L": L, "last": int(s[-1]), "fnv": hx(s)})

    # ---- reduce: reduction/reduction.cpp:483-491 (never seeded => glibc seed 1) --------------
    kat["reduce"] = {"source": "restatement", "cases": []}
    for logn in (20, 24):
        ri = oracle.sample_input("reduce_i32", 1 << logn, 1)
        rf = oracle.sample_input("reduce_f32", 1 << logn, 1)
        rd = oracle.sample_input("reduce_f64", 1 << logn, 1)
        assert oracle.reduce(ri) == int(ri.astype(np.int64).sum())  # no wrap at these sizes
        kat["reduce"]["cases"].append({"n": 1 << logn, "i32": oracle.reduce(ri),
                                       "f32_hex": float(np.float32(oracle.reduce(rf))).hex(),
                                       "f64_hex": oracle.reduce(rd).hex()})

    # ---- mergeSort: mergeSort/main.cpp:46-71 (N = 4 Mi, srand(2009); rand()%65536; val=i) ----
    N = 4 << 20
    k = oracle.sample_input("sort_keys", N, 2009)
    v = np.arange(N, dtype=np.uint32)
    mk, mv = ref.merge_sort_host(k, v, 1)
    order = np.argsort(k, kind="stable").astype(np.uint32)  # independent statement of "stable"
    kat["mergeSort"] = {"n": N, "seed": 2009, "dir": 1,
                        "mergeSortHost": {"fnv_keys": hx(mk), "fnv_vals": hx(mv)},
                        "stable": {"fnv_keys": hx(k[order]), "fnv_vals": hx(order)},
                        "mergeSortHost_vals_differing_from_stable": int((mv != order).sum()),
                        "validateSortedKeys": ref.validate_sorted_keys(mk, k, 65536, 1),
                        "validateSortedValues": ref.validate_sorted_values(mk, mv, k)}

    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1, sort_keys=True)

    # ---- small fixtures -------------------------------------------------------------------------
    rng = np.random.default_rng(20091)
    small = {}
    hb = rng.integers(0, 256, 4096, dtype=np.uint8)
    hb[:512] = 7  # a skewed run
    small["hist_in"], small["hist256_out"], small["hist64_out"] = hb, ref.histogram256(hb), ref.histogram64(hb)
    sc = rng.integers(0, 2**32, 8192, dtype=np.uint32)
    small["scan_in"] = sc
    for L in (4, 32, 1024, 2048, 8192):
        small["scan_out_L%d" % L] = ref.scan_exclusive(sc, sc.size // L, L)
    # sort: D1 (all unique), D2 (unique inside every 1024 tile, duplicated across), D3 (duplicates in tiles)
    n = 8192
    d1 = oracle.generate("key_unique", 0, n, 7)
    d2 = oracle.generate("key_tile_unique", 0, n, 0, 2048)
    d3 = rng.integers(0, 50, n, dtype=np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    for name, keys in (("d1", d1), ("d2", d2), ("d3", d3)):
        small["sort_%s_keys" % name] = keys
        for direction in (1, 0):
            rk, rv = ref.merge_sort_host(keys, vals, direction)
            small["sort_%s_dir%d_msh_keys" % (name, direction)] = rk
            small["sort_%s_dir%d_msh_vals" % (name, direction)] = rv
    np.savez_compressed(os.path.join(HERE, "small.npz"), **small)
    print("wrote kat.json and small.npz (%d arrays)" % len(small))


if __name__ == "__main__":
    main()
