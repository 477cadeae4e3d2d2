#!/usr/bin/env python3
"""The reference's own mergeSortHost (mergeSort/mergeSort_host.cpp:288-363, compiled unmodified into
oracle/_ref/libref_gold.so) run ONCE at the full 2^28 pairs of BASELINE.json configs[4], on the dataset where it
coincides with the stable order (all keys distinct: "key_unique", the counter-based generator shared by oracle.c and
csrc/generate.cu).  Takes a few minutes on one core and ~8 GiB of host memory, so the result is committed as hashes:

    python tests/golden/make_golden_sort_2p28.py   ->  tests/golden/kat_sort_2p28.json

tests/test_gpu_parity.py::test_sort_full_size_matches_mergeSortHost_hashes regenerates the same input on the GPU and
compares the hashes of the GPU sort's output (FNV-1a-64 over the little-endian output bytes, oracle_fnv1a64)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

N, SEED = 1 << 28, 7


def main():
    ref = oracle.ref
    assert ref is not None, "reference gold not available: run inside the build container"
    keys = oracle.generate("key_unique", 0, N, SEED, 0)
    vals = np.arange(N, dtype=np.uint32)
    t0 = time.time()
    mk, mv = ref.merge_sort_host(keys, vals, 1)
    dt = time.time() - t0
    assert bool((mk[1:] > mk[:-1]).all()), "keys are distinct: strictly ascending"
    assert bool((keys[mv] == mk).all()), "every value still names its key"
    out = {"n": N, "seed": SEED, "kind": "key_unique", "dir": 1, "generated_by": "reference mergeSortHost",
           "seconds_single_thread": round(dt, 1),
           "fnv_keys": "%016x" % oracle.fnv1a64(mk), "fnv_vals": "%016x" % oracle.fnv1a64(mv),
           "first": [int(mk[0]), int(mv[0])], "last": [int(mk[-1]), int(mv[-1])]}
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "kat_sort_2p28.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(out)


if __name__ == "__main__":
    main()
