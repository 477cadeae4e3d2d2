#!/usr/bin/env python3
"""bench.py -- throughput of the primitive hot path on B200, one JSON line on stdout (rank 0).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--skip-table]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (BASELINE.json configs[1]): one step = reduce<int> + reduce<float> SUM over 2^28
elements each, resident in HBM (2 GiB of algorithmic bytes per step, inputs 8x larger than L2).
At N GPUs every rank reduces its own 2^28-element shards and the shard sums are combined with NCCL
(weak scaling).  `value` = bytes all ranks reduced / max-over-ranks device time.
`e2e` = the same step through the host-buffer C ABI (b200_reduce_*_host): pinned host inputs, H2D inside
the timed region, result read back.  `primitives` carries the same measurement for every other primitive
of the path (histogram64/256, scan, key-value sort) with its own roofline fraction.
`--impl reference` times the samples' CPU gold path (reduceCPU, restated in oracle/) on the host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "HBM GB/s (% of B200 peak) per primitive; sort Gkeys/s at 1/2/4/8 GPUs"
N_ELEMS = 1 << 28
HIST_BYTES_PER_RANK = 4 << 30  # configs[3]: 32 GiB over 8 GPUs; timed AND checked against the gold at this size
WORKLOAD = "reduce<int> + reduce<float> SUM over 2^28 elements each per GPU (BASELINE configs[1])"
OVERLAP = os.environ.get("B200_BENCH_OVERLAP", "0") == "1"  # async all_gather behind the next step's kernels
HBM_NOMINAL = 8000.0
HBM_FALLBACK = 6650.0  # /opt/skills/guides/B200_PROFILING.md, only when MEASURED_PEAKS.json is absent


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return HBM_FALLBACK, "fallback (B200_PROFILING.md)"


def base_config(n_gpus):
    return {"workload": WORKLOAD, "elements_per_gpu": N_ELEMS, "bytes_per_step_per_gpu": 2 * 4 * N_ELEMS,
            "l2_policy": "inputs (1 GiB each) larger than L2 (126 MB); no flush needed",
            "parallelism": "shard%d" % n_gpus, "seed": 1, "generator": "splitmix64 counter-based (oracle_generate twin)"}


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi poller (the recipe's clocks line) running while the GPU is under the bench load."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                   f[5:9]):
                    if v == "Active":
                        reasons.add(name)
            os.remove(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------ timing
def time_steps(torch, fn, steps, warmup, barrier=None, finish=None):
    """W untimed steps, then exactly K steps between CUDA events on the current stream; -> seconds.
    `finish` (optional) completes whatever the last step left in flight and runs INSIDE the timed region."""
    for _ in range(warmup):
        fn()
    if finish:
        finish()
    torch.cuda.synchronize()
    if barrier:
        barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    if finish:
        finish()
    e1.record()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    return e0.elapsed_time(e1) * 1e-3


def max_over_ranks(torch, dist, seconds, world):
    if world == 1:
        return seconds
    t = torch.tensor([seconds], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# -------------------------------------------------------------------------------- reference arm (CPU)
def reference_gold():
    """-> (callable reduce(np array), kind): the reference's OWN reduceCPU<T> (reduction.cpp:167-180, compiled from the
    sample's driver TU where it lies, oracle/_ref/libref_reduce.so) when that file exists, else the restatement."""
    import oracle
    if oracle.ref_reduce is not None:
        return oracle.ref_reduce.reduce, "reference"
    return oracle.reduce, "port"


def cpu_reduce_parallel(reduce_fn, np, pool, xi, xf, threads):
    """Chunk-parallel harness around the gold reduceCPU: one contiguous chunk per thread, partials added up."""
    bounds = np.linspace(0, xi.size, threads + 1).astype(np.int64)
    jobs = [pool.submit(reduce_fn, xi[bounds[i]:bounds[i + 1]]) for i in range(threads)]
    jobs += [pool.submit(reduce_fn, xf[bounds[i]:bounds[i + 1]]) for i in range(threads)]
    res = [j.result() for j in jobs]
    si = sum(res[:threads]) & 0xFFFFFFFF
    sf = float(np.float64(sum(res[threads:])))
    return si, sf


def run_reference(args):
    """The samples' CPU gold path for the headline workload -- reduceCPU<int> + reduceCPU<float> over the DECLARED
    2^28 elements each per step -- timed on the host cores with every thread it can use (the gold code itself is
    single-threaded: each thread runs it on one contiguous chunk and the partials are added).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import concurrent.futures

    import numpy as np

    import oracle
    reduce_fn, kind = reference_gold()
    threads = len(os.sched_getaffinity(0))
    xi = oracle.generate("i32_byte", 0, N_ELEMS, 1)
    xf = oracle.generate("f32_byte", 0, N_ELEMS, 1)
    # the gold exactly as shipped: one thread, one call per array, whole input
    t0 = time.perf_counter()
    one_i, one_f = reduce_fn(xi), reduce_fn(xf)
    single_s = time.perf_counter() - t0
    pool = concurrent.futures.ThreadPoolExecutor(max_workers=threads)
    for _ in range(args.warmup):
        cpu_reduce_parallel(reduce_fn, np, pool, xi, xf, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        par_i, par_f = cpu_reduce_parallel(reduce_fn, np, pool, xi, xf, threads)
    dt = time.perf_counter() - t0
    gbs = args.steps * 2 * 4 * N_ELEMS / dt / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i32+f32",
            "data": "synthetic", "config": base_config(args.gpus),
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": threads, "kind": kind,
                             "sample": "every step = reduceCPU<int> + reduceCPU<float> (Kahan, reduction.cpp:167-180) "
                                       "over the full 2^28 elements each (2 GiB), one contiguous chunk per thread, "
                                       "partials added; nothing sampled down",
                             "single_thread_as_shipped_GBs": 2 * 4 * N_ELEMS / single_s / 1e9,
                             "single_thread_as_shipped_seconds": single_s,
                             "chunked_equals_single_thread": {"int": (par_i & 0xFFFFFFFF) == (one_i & 0xFFFFFFFF),
                                                              "float_rel_diff": abs(par_f - one_f) / abs(one_f)}},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def cpu_gold_table(oracle, torch, prim):
    """The samples' CPU gold functions, single thread as shipped, at the sizes SURVEY.md 8(d) asks for: the whole
    64 MiB histogram input and one 2 GiB chunk of the 32 GiB config (x16 stated), reduceCPU<double> at 2^27,
    scanExclusiveHost at the full 2^28, mergeSortHost at 2^24 with the extrapolation to 2^28 stated.
    Inputs come from the device generator (bit-identical to oracle_generate, tests/test_gpu_parity.py) and are copied
    to the host: generating GiBs on one host core would take longer than the gold functions themselves.
    `kind` = "reference" when the function ran from oracle/_ref (the reference's own sources compiled unmodified),
    "port" for the restatement in oracle/oracle.c."""
    import numpy as np
    gold = oracle.ref if oracle.ref is not None else oracle
    kind = "reference" if oracle.ref is not None else "port"
    out = {}

    def timed(fn, reps=3):
        best = None
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        return best

    def device_generated(kind_, n, seed, dtype, param=0):
        d = torch.empty(n, dtype=dtype, device="cuda")
        prim.generate(d, kind_, 0, n, seed, param)
        h = d.cpu().numpy()
        del d
        return h

    data = device_generated("bytes", 2 << 30, 2009, torch.uint8)
    for name, fn in (("histogram256CPU", gold.histogram256), ("histogram64CPU", gold.histogram64)):
        dt = timed(lambda: fn(data[:64 << 20]))
        out[name] = {"GB/s": round((64 << 20) / dt / 1e9, 3), "seconds": round(dt, 4),
                     "sample": "64 MiB (the whole configs[0] input)", "cores": 1, "kind": kind}
        dt = timed(lambda: fn(data), 1)
        out[name + "_2GiB_chunk"] = {"GB/s": round(data.size / dt / 1e9, 3), "seconds": round(dt, 3), "cores": 1,
                                     "kind": kind, "sample": "one 2 GiB chunk of the 32 GiB configs[3] input; the whole "
                                     "config is 16 such chunks: x16 = %.1f s on one core" % (16 * dt)}
    del data
    rg, rkind = reference_gold()
    xd = device_generated("f64_byte", 1 << 27, 1, torch.float64)
    dt = timed(lambda: rg(xd), 2)
    out["reduceCPU<double>"] = {"GB/s": round(xd.nbytes / dt / 1e9, 3), "seconds": round(dt, 4), "cores": 1,
                                "kind": rkind, "sample": "2^27 doubles (1 GiB), Kahan"}
    del xd
    n = N_ELEMS
    src = device_generated("u31", n, 2009, torch.int32).view(np.uint32)
    dt = timed(lambda: gold.scan_exclusive(src, 1, n), 2)
    out["scanExclusiveHost"] = {"GB/s": round(8 * n / dt / 1e9, 3), "seconds": round(dt, 4), "cores": 1, "kind": kind,
                                "sample": "one array of the full 2^28 u32 (configs[2]), 8 B per element"}
    del src
    n = 1 << 24
    keys = device_generated("key_unique", n, 2009, torch.int32).view(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    dt = timed(lambda: gold.merge_sort_host(keys, vals, 1), 1)
    out["mergeSortHost"] = {"Gkeys/s": round(n / dt / 1e9, 5), "seconds": round(dt, 3), "cores": 1, "kind": kind,
                            "sample": "2^24 pairs, D1 unique keys (1/16 of configs[4])",
                            "extrapolated_2^28_seconds": round(16 * dt * 18 / 14, 1),
                            "extrapolation": "x16 pairs, and the merge levels (log2(N/1024): 14 at 2^24, 18 at 2^28) "
                                             "dominate the run time: x 18/14"}
    img = oracle.generate("bytes", 0, 4096 * 4096, 11).reshape(4096, 4096)
    dt = timed(lambda: oracle.integral_image(img))
    out["integral_image"] = {"GB/s": round(5 * img.size / dt / 1e9, 3), "seconds": round(dt, 4), "cores": 1,
                             "kind": "port", "sample": "4096 x 4096 u8 image, 5 B per pixel"}
    return out


# ---------------------------------------------------------------------------------- primitives table
def primitives_table(torch, dist, prim, mg, rank, world, peak, barrier, p2p=False):
    """GB/s (Gkeys/s for sort) of every primitive of the path, device-resident inputs, CUDA events."""
    table = {}

    def record(name, seconds_per_call, alg_bytes, extra=None, gpus=None):
        gpus = gpus or world  # rows measured on every rank independently (no sharding) pass gpus=1
        t = max_over_ranks(torch, dist, seconds_per_call, world)
        gbs = alg_bytes / t / 1e9
        row = {"GB/s": round(gbs, 1), "us_per_call": round(t * 1e6, 2), "algorithmic_bytes": alg_bytes,
               "frac_of_measured_peak": round(gbs / (peak * gpus), 4),
               "frac_of_nominal_8TBs": round(gbs / (HBM_NOMINAL * gpus), 4), "n_gpus": gpus}
        if extra:
            row.update(extra)
        table[name] = row

    # --- histogram, 64 MiB (BASELINE configs[0]); 4 rotating inputs = 256 MiB > L2 so every call reads HBM
    n64 = 64 << 20
    bufs = [torch.empty(n64, dtype=torch.uint8, device="cuda") for _ in range(4)]
    for i, b in enumerate(bufs):
        prim.generate(b, "bytes", i * n64, n64, 2009)
    h256 = torch.empty(256, dtype=torch.int32, device="cuda")
    h64 = torch.empty(64, dtype=torch.int32, device="cuda")
    it = [0]

    def step256():
        prim.histogram256(h256, bufs[it[0] & 3], n64)
        it[0] += 1

    def step64():
        prim.histogram64(h64, bufs[it[0] & 3], n64)
        it[0] += 1
    record("histogram256_64MiB", time_steps(torch, step256, 40, 8) / 40, n64, {"l2": "4 rotating inputs (256 MiB)"}, 1)
    record("histogram64_64MiB", time_steps(torch, step64, 40, 8) / 40, n64, {"l2": "4 rotating inputs (256 MiB)"}, 1)
    # the same 40 calls replayed from a CUDA graph (a 10 us kernel is comparable to the launch latency
    # of a Python/ctypes loop, ~11 us per call on this host)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(40):
                step256()
        record("histogram256_64MiB_cuda_graph", time_steps(torch, g.replay, 5, 2) / (5 * 40), n64,
               {"l2": "4 rotating inputs (256 MiB)", "note": "40 calls per graph replay"}, 1)
        del g
    torch.cuda.current_stream().wait_stream(side)
    del bufs
    # --- histogram, 4 GiB per GPU, both histograms in one pass (BASELINE configs[3] shape: 32 GiB over 8)
    big = HIST_BYTES_PER_RANK
    d = torch.empty(big, dtype=torch.uint8, device="cuda")
    prim.generate(d, "bytes", rank * big, big, 2009)

    def step_fused():
        mg.histogram_sharded(d, big)
    record("histogram64+256_fused_4GiB_per_gpu", time_steps(torch, step_fused, 5, 2, barrier) / 5, big * world,
           {"scaling": "weak", "collective": (("in-kernel NVLink peer-window all-reduce (256 x u32)" if p2p else
                                              "all_reduce(320 x u32)") if world > 1 else None),
            "note": "1 B per input byte for BOTH histograms (one pass)"})
    h256b = torch.empty(256, dtype=torch.int32, device="cuda")

    def step_256_big():
        prim.histogram256(h256b, d, big)
    record("histogram256_4GiB_single_gpu", time_steps(torch, step_256_big, 5, 2) / 5, big, None, 1)
    del d
    # --- scan, 2^28 u32 in total (BASELINE configs[2]: strong scaling, carry via all_gather of shard totals)
    n = N_ELEMS // world
    src = torch.empty(n, dtype=torch.int32, device="cuda")
    dst = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", rank * n, n, 2009)

    def step_scan():
        mg.scan_sharded(dst, src, n)
    second_pass = None
    if world > 1 and p2p:
        # Trial of the tile-prefix path under the timing loop's own conditions (calls back to back, no host sync between
        # them) with a SHORT peer timeout: a rank that ran ahead of its peers' reads would show up here as lost words
        # within a second instead of stalling the timed loop for the default 30 s per call.  On any loss every rank
        # switches to the chained second pass (the path of the earlier records) on fresh windows.
        from cuda_samples_b200 import api
        second_pass = "independent tiles seeded with tile prefixes (first pass: tile sums + exchange of the shard totals)"
        step_scan()  # first call: scratch is sized here, under the default timeout
        torch.cuda.synchronize()
        barrier()
        api.p2p_set_timeout_ms(250)
        for _ in range(24):  # as many calls as the timed region below issues
            step_scan()
        torch.cuda.synchronize()
        lost = torch.tensor([int(prim.p2p_errors())], dtype=torch.int64, device="cuda")
        dist.all_reduce(lost)
        if int(lost.item()):
            api.p2p_set_timeout_ms(30000)
            prim.lib.b200_tune_scan_shard_tiled(0)
            mg.disable_p2p()  # (the loss count of a window set is sticky: new windows)
            mg.enable_p2p()
            second_pass = "chained scan seeded with the carry (the tile-prefix path lost %d words in its trial)" % int(lost.item())
    t_scan = time_steps(torch, step_scan, 20, 4, barrier) / 20
    scan_extra = {"scaling": "strong", "collective": (("in-kernel NVLink peer-window exchange of shard totals" if p2p else
                                                       "all_gather(G x u32)") if world > 1 else None),
                  "second_pass": second_pass}
    if world > 1 and p2p:
        torch.cuda.synchronize()
        lost = torch.tensor([int(prim.p2p_errors())], dtype=torch.int64, device="cuda")
        dist.all_reduce(lost)
        api.p2p_set_timeout_ms(30000)
        scan_extra["lost_words_in_timed_loop"] = int(lost.item())  # 0, or the row below is void
        if int(lost.item()):  # (all-reduced: every rank takes this branch together) fresh windows for the sort rows
            prim.lib.b200_tune_scan_shard_tiled(0)
            mg.disable_p2p()
            mg.enable_p2p()
    record("scanExclusiveLarge_2^28", t_scan, 8 * N_ELEMS, scan_extra)
    if world == 1:
        def step_scan_short():
            prim.scanExclusiveShort(dst, src, n // 1024, 1024)
        record("scanExclusiveShort_L1024_2^28", time_steps(torch, step_scan_short, 20, 4) / 20, 8 * N_ELEMS)
        # scan siblings: inclusive scan (shfl_scan) and ordered stream compaction (filter_arr)

        def step_scan_incl():
            prim.scanInclusive(dst, src, 1, n)
        record("scanInclusive_2^28", time_steps(torch, step_scan_incl, 20, 4) / 20, 8 * N_ELEMS)
        prim.generate(src, "key_mod", 0, n, 7, 5)  # the sample's rand() % 5: 80 % of the elements are kept
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")

        def step_compact():
            prim.filter_arr(dst, cnt, src, n)
        t_c = time_steps(torch, step_compact, 20, 4) / 20
        kept = int(cnt.item())
        record("filter_arr_compaction_2^28", t_c, 4 * N_ELEMS + 4 * kept,
               {"kept": kept, "note": "4 B read per element + 4 B written per kept element, output in input order"})
        # integral image (shfl_scan's second test) at an HBM-sized image and at the sample's 1920 x 1080
        for ih, iw, reps in ((16384, 16384, 10), (1080, 1920, 50)):
            img = torch.empty(ih * iw, dtype=torch.uint8, device="cuda")
            prim.generate(img, "bytes", 0, ih * iw, 11)
            ii_out = dst[:ih * iw]

            def step_ii():
                prim.integral_image(ii_out, img, iw, ih)
            record("integral_image_%dx%d" % (iw, ih), time_steps(torch, step_ii, reps, 4) / reps, 5 * ih * iw,
                   {"note": "1 B read + 4 B written per pixel; the band sums re-read the image (6 B/pixel moved)"})
            del img
    del src, dst
    # --- reduce<double> (instantiated by the sample, reduction_kernel.cu:985-989): per-thread Kahan in fp64 is the
    #     one reduce variant with real ALU work per byte
    if world == 1:
        nd = 1 << 27
        xd = torch.empty(nd, dtype=torch.float64, device="cuda")
        prim.generate(xd, "f64_byte", 0, nd, 1)
        od = torch.empty(64, dtype=torch.float64, device="cuda")

        def step_f64():
            prim.reduce(nd, 256, 64, 7, xd, od)
        record("reduce<double>_2^27", time_steps(torch, step_f64, 20, 4) / 20, 8 * nd)
        del xd
    # --- reduction sibling: batched dot products, 4096 vectors x 65536 floats (2 GiB read)
    if world == 1:
        vn, en = 4096, 65536
        da = torch.empty(vn * en, dtype=torch.float32, device="cuda")
        db = torch.empty(vn * en, dtype=torch.float32, device="cuda")
        prim.generate(da, "f32_byte", 0, vn * en, 3)
        prim.generate(db, "f32_byte", 1 << 40, vn * en, 3)
        dc = torch.empty(vn, dtype=torch.float32, device="cuda")

        def step_dot():
            prim.scalarProd(dc, da, db, vn, en)
        record("scalarProd_4096x65536", time_steps(torch, step_dot, 10, 3) / 10, 8 * vn * en,
               {"note": "8 B per element pair"})
        del da, db
    # --- key-value sort, 2^28 pairs in total (BASELINE configs[4]), dataset D1 (all keys distinct)
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    sv = torch.arange(rank * n, rank * n + n, dtype=torch.int64, device="cuda").to(torch.int32)
    prim.generate(sk, "key_unique", rank * n, n, 2009)
    if world == 1:
        dk, dv, bk, bv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4))

        def step_sort():
            prim.mergeSort(dk, dv, bk, bv, sk, sv, n, 1)
    else:
        def step_sort():
            mg.sort_sharded(sk, sv, n, 1)
    ts = time_steps(torch, step_sort, 5, 2, barrier) / 5
    t = max_over_ranks(torch, dist, ts, world)
    table["mergeSort_kv_2^28"] = {
        "Gkeys/s": round(N_ELEMS / t / 1e9, 3), "ms_per_call": round(t * 1e3, 3), "n_gpus": world,
        "traffic_model_bytes_per_pair": 68, "GB/s_at_68B_per_pair": round(68 * N_ELEMS / t / 1e9, 1),
        "frac_of_measured_peak": round(68 * N_ELEMS / t / 1e9 / (peak * world), 4),
        "frac_of_nominal_8TBs": round(68 * N_ELEMS / t / 1e9 / (HBM_NOMINAL * world), 4),
        "scaling": "strong", "dataset": "D1 unique keys, val=i, ascending",
        "collective": (("sample sort, no NCCL call: samples exchanged through the NVLink peer windows by the splitter "
                        "kernel, partition pass storing into the peers' receive buffers over NVLink") if p2p else
                       "sample sort: all_gather + all_to_all") if world > 1 else None}
    if world > 1 and p2p:
        # ---- exchange phase alone, against the NVLink roofline (north_star: "NVLink roofline for the exchange phases")
        def step_exchange():
            mg.sort_exchange_only(sk, sv, n, 1)
        tx = max_over_ranks(torch, dist, time_steps(torch, step_exchange, 5, 2, barrier) / 5, world)
        meta = mg.sort_exchange_only(sk, sv, n, 1).cpu().numpy()
        torch.cuda.synchronize()
        m = meta[:world * world].reshape(world, world)
        sent = int(m[rank].sum() - m[rank, rank])
        st = torch.tensor([sent], dtype=torch.int64, device="cuda")
        dist.all_reduce(st)
        mean_sent = int(st.item()) / world
        probe = mg.nvlink_probe(sk, 4 * n, n)
        pt = torch.tensor([probe[k] for k in sorted(probe)], dtype=torch.float64, device="cuda")
        dist.all_reduce(pt, op=dist.ReduceOp.MIN)  # the slowest rank's figure per mode
        probe = {k: round(float(v), 1) for k, v in zip(sorted(probe), pt.tolist())}
        gbs = 8 * mean_sent / tx / 1e9
        table["mergeSort_kv_2^28"]["exchange"] = {
            "phase": "splitters on the device + bucket counts and their exchange + partition pass storing the runs into "
                     "the peers' receive buffers + node-wide barrier (everything of the sort that crosses NVLink)",
            "bytes_out_per_gpu": int(8 * mean_sent), "seconds": tx, "GB/s_per_gpu": round(gbs, 1),
            "frac_of_900_nominal": round(gbs / 900.0, 4),
            "nvlink_probe_GBs_per_direction_per_gpu": probe,
            "frac_of_measured_peer_copy": round(gbs / probe["cudaMemcpyAsync"], 4) if probe.get("cudaMemcpyAsync") else None,
            "note": "the phase also reads the shard twice from HBM (count, partition) and ranks every pair: its time "
                    "is not pure link time; the small exchanges of the other primitives are latency-bound "
                    "(<= 2 KiB per rank, ~8 us per exchange, tests/multi_gpu_check.py)"}
    if world == 1:
        def step_sort_desc():
            prim.mergeSort(dk, dv, bk, bv, sk, sv, n, 0)
        td = time_steps(torch, step_sort_desc, 5, 2) / 5
        table["mergeSort_kv_2^28_descending"] = {
            "Gkeys/s": round(N_ELEMS / td / 1e9, 3), "ms_per_call": round(td * 1e3, 3), "n_gpus": 1,
            "frac_of_measured_peak": round(68 * N_ELEMS / td / 1e9 / peak, 4), "dataset": "D1, sortDir = 0"}
        # the sample's own distribution (mergeSort/main.cpp:46-71: keys = rand() % 65536): the two upper digit
        # passes see one digit value only
        prim.generate(sk, "key_mod", 0, n, 2009, 65536)
        t3 = time_steps(torch, step_sort, 5, 2) / 5
        table["mergeSort_kv_2^28_keys_mod_65536"] = {
            "Gkeys/s": round(N_ELEMS / t3 / 1e9, 3), "ms_per_call": round(t3 * 1e3, 3), "n_gpus": 1,
            "frac_of_measured_peak": round(36 * N_ELEMS / t3 / 1e9 / peak, 4),
            "bytes_per_pair": "4 (digit histograms) + 2 x 16 (the two passes whose digit varies; the device-side "
                              "pass plan skips the other two)",
            "dataset": "D3-like: key = counter-based random % 65536 (the sample's rand() % 65536 shape), ascending"}
        # the sample's own size (mergeSort/main.cpp:46: N = 4 Mi pairs) with the same keys
        m = 4 << 20
        t4 = time_steps(torch, lambda: prim.mergeSort(dk[:m], dv[:m], bk[:m], bv[:m], sk[:m], sv[:m], m, 1), 50, 5) / 50
        table["mergeSort_kv_4Mi_keys_mod_65536"] = {
            "Gkeys/s": round(m / t4 / 1e9, 3), "us_per_call": round(t4 * 1e6, 1), "n_gpus": 1,
            "dataset": "the sample's own shape: 4 Mi pairs, key = random % 65536, ascending (L2-resident: 32 MiB of "
                       "pairs per buffer)"}
    if world == 1:
        # --- batched short-array sorts (sortingNetworks API): arrays <= 4096 pairs are sorted
        #     inside one CTA's shared memory: 16 B of traffic per pair.  Keys like the sample's (rand() % 65536).
        del sk, sv
        for batch, L in ((16384, 64), (256, 4096), (1 << 20, 64), (16384, 4096)):
            m = batch * L
            bsk = torch.empty(m, dtype=torch.int32, device="cuda")
            prim.generate(bsk, "key_mod", 0, m, 5, 65536)
            bsv = torch.arange(m, dtype=torch.int32, device="cuda")
            bdk, bdv = torch.empty_like(bsk), torch.empty_like(bsv)

            def step_bitonic():
                prim.bitonicSort(bdk, bdv, bsk, bsv, batch, L, 1)
            tb = time_steps(torch, step_bitonic, 20, 4) / 20
            table["bitonicSort_%dx%d" % (batch, L)] = {
                "Gpairs/s": round(m / tb / 1e9, 3), "us_per_call": round(tb * 1e6, 2), "n_gpus": 1,
                "GB/s_at_16B_per_pair": round(16 * m / tb / 1e9, 1),
                "frac_of_measured_peak": round(16 * m / tb / 1e9 / peak, 4),
                "note": "SM-bound, not HBM-bound: two stable radix passes (16-bit keys) + one on the array index, "
                        "~70 instructions per pair per pass; the sample's 2^20-pair shapes are one wave of CTAs"}
            del bsk, bsv, bdk, bdv
        torch.cuda.empty_cache()
        isolated_and_host_rows(torch, prim, table, peak)
    return table


def isolated_and_host_rows(torch, prim, table, peak):
    """(1) The 64 MiB histogram as ONE call between synchronisations (the rows above time 40 calls back to back, where
    call k+1 overlaps the drain of call k through programmatic dependent launch: a caller who histograms once gets
    this row instead).  (2) histogram / scan / sort through the host-buffer ABI (b200_*_host: pinned host memory in,
    host memory out, copies inside the call, wall clock) -- the callers' view of histogram/main.cpp:136-170,
    scan/main.cpp:59-61,87 and mergeSort/main.cpp:80-97 with their cudaMemcpy's counted."""
    import numpy as np
    n64 = 64 << 20
    bufs = [torch.empty(n64, dtype=torch.uint8, device="cuda") for _ in range(4)]
    for i, b in enumerate(bufs):
        prim.generate(b, "bytes", i * n64, n64, 2009)
    h256 = torch.empty(256, dtype=torch.int32, device="cuda")
    times = []
    for i in range(24):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        prim.histogram256(h256, bufs[i & 3], n64)
        e1.record()
        torch.cuda.synchronize()
        if i >= 4:
            times.append(e0.elapsed_time(e1) * 1e-3)
    t = float(np.median(times))
    table["histogram256_64MiB"]["mode"] = "back to back: 40 calls on one stream, consecutive calls overlap (PDL)"
    table["histogram64_64MiB"]["mode"] = table["histogram256_64MiB"]["mode"]
    table["histogram256_64MiB_isolated"] = {
        "GB/s": round(n64 / t / 1e9, 1), "us_per_call": round(t * 1e6, 2), "algorithmic_bytes": n64,
        "frac_of_measured_peak": round(n64 / t / 1e9 / peak, 4), "n_gpus": 1,
        "mode": "isolated: one call between device synchronisations, CUDA events around it (launch latency "
                "included), median of 20"}
    del bufs

    def wall(fn, reps=3):
        fn()
        best = None
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            dt = time.perf_counter() - t0
            best = dt if best is None or dt < best else best
        return best

    def pinned(nbytes, dtype):
        hb = prim.HostBuffer(nbytes)
        return hb, hb.array(dtype)

    def fill(host_arr, dev_tensor):
        torch.from_numpy(host_arr.view(np.uint8)).copy_(dev_tensor.view(torch.uint8))

    def host_row(name, t, h2d, d2h, pcie, units, unit_name, api, ok, gold_row):
        row = {"seconds": round(t, 5), "h2d_bytes": h2d, "d2h_bytes": d2h, "api": api,
               unit_name: round(units / t / 1e9, 3), "GB/s_h2d": round(h2d / t / 1e9, 2),
               "pcie_gbs_h2d_probe": round(pcie, 2),
               "frac_of_pcie": round(max(h2d, d2h) / t / 1e9 / pcie, 4) if pcie else None,
               "frac_of_pcie_note": "the larger direction's bytes / time over the same-run pinned H2D probe (the two "
                                    "directions can overlap: PCIe is full duplex)",
               "result_equals_device_path": bool(ok), "cpu_gold_row": gold_row, "n_gpus": 1}
        table[name] = row

    # histogram: 64 MiB and 4 GiB of pinned bytes -> 256 bins
    for label, nb in (("64MiB", 64 << 20), ("4GiB", 4 << 30)):
        d = torch.empty(nb, dtype=torch.uint8, device="cuda")
        prim.generate(d, "bytes", 0, nb, 2009)
        hb, ha = pinned(nb, np.uint8)
        fill(ha, d)
        want = torch.empty(256, dtype=torch.int32, device="cuda")
        prim.histogram256(want, d, nb)
        del d
        pcie = prim.host_probe_h2d(ha, nb, 3)
        got = [None]

        def step_h():
            got[0] = prim.histogram256_host(ha)
        t = wall(step_h)
        host_row("e2e_histogram256_" + label, t, nb, 1024, pcie, nb, "GB/s", "b200_histogram256_host",
                 (got[0] == want.cpu().numpy().view(np.uint32)).all(), "histogram256")
        hb.free()
    # scan: 2^28 u32 in, 2^28 u32 out
    n = N_ELEMS
    src = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", 0, n, 2009)
    hb_s, hs = pinned(4 * n, np.uint32)
    hb_d, hd = pinned(4 * n, np.uint32)
    fill(hs, src)
    want = torch.empty_like(src)
    prim.scanExclusiveLarge(want, src, 1, n)
    pcie = prim.host_probe_h2d(hs, 4 * n, 3)
    t = wall(lambda: prim.scan_exclusive_host(hd, hs, 1, n))
    src.copy_(torch.from_numpy(hd.view(np.int32)))
    host_row("e2e_scanExclusiveLarge_2^28", t, 4 * n, 4 * n, pcie, 8 * n, "GB/s_in+out", "b200_scan_exclusive_host",
             torch.equal(src, want), "scanExclusiveHost")
    del src, want
    hb_s.free(), hb_d.free()
    # key-value sort: 2^28 pairs in, 2^28 pairs out
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(sk, "key_unique", 0, n, 2009)
    sv = torch.arange(n, dtype=torch.int64, device="cuda").to(torch.int32)
    hbs = [pinned(4 * n, np.uint32) for _ in range(4)]
    (hsk, hsv, hdk, hdv) = [a for _, a in hbs]
    fill(hsk, sk), fill(hsv, sv)
    dk, dv, bk, bv = (torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(4))
    prim.mergeSort(dk, dv, bk, bv, sk, sv, n, 1)
    del bk, bv, sk, sv
    pcie = prim.host_probe_h2d(hsk, 4 * n, 3)
    t = wall(lambda: prim.sort_kv_host(hdk, hdv, hsk, hsv, n, 1), reps=2)
    ok = bool((torch.from_numpy(hdk.view(np.int32)) == dk.cpu()).all()) and \
        bool((torch.from_numpy(hdv.view(np.int32)) == dv.cpu()).all())
    host_row("e2e_mergeSort_kv_2^28", t, 8 * n, 8 * n, pcie, n, "Gkeys/s", "b200_sort_kv_host", ok, "mergeSortHost")
    del dk, dv
    for hb, _ in hbs:
        hb.free()


# ------------------------------------------------------------------------------------------- parity
def parity_block(torch, dist, prim, mg, oracle, rank, world, xi, xf, gpu_sums, e2e_sums):
    """GPU result vs the samples' CPU gold INSIDE the bench run, at every N (the reference drivers' own pattern:
    GPU vs gold in the same process, histogram/main.cpp:136-147, reduction.cpp:551-590).  Every rank checks its own
    slice against the gold run on the bytes its kernels consumed (device input copied back to the host, like the
    samples' h_Data / d_Data pair); cross-rank quantities (global bins, sums, scan carries) are combined from the
    per-rank gold values over the process group.  -> dict for the JSON line (rank 0), `failures` all-reduced."""
    import numpy as np
    gold = oracle.ref if oracle.ref is not None else oracle
    reduce_gold, rkind = reference_gold()
    out = {"gold": "oracle/_ref (reference sources compiled unmodified)" if oracle.ref is not None else "oracle port",
           "reduceCPU": rkind}
    fails = []

    def allsum_i64(values):
        t = torch.tensor(values, dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(t)
        return t.cpu().numpy()

    def allgather_f64(value):
        t = torch.tensor([float(np.asarray(value).reshape(-1)[0])], dtype=torch.float64, device="cuda")
        g = torch.empty(world, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_gather_into_tensor(g, t)
        else:
            g.copy_(t)
        return g.cpu().numpy()

    # ---- headline reduce, FULL size (2^28 per rank per type): int bit-exact, float within 1e-5 of Kahan
    hi, hf = xi.cpu().numpy(), xf.cpu().numpy()
    t0 = time.perf_counter()
    gi, gf = reduce_gold(hi), reduce_gold(hf)
    out["gold_reduce_seconds"] = time.perf_counter() - t0  # bench.py's cpu_baseline at N=1: one thread, as shipped
    want_i = int(allsum_i64([gi & 0xFFFFFFFF])[0]) & 0xFFFFFFFF
    want_f = float(np.float64(allgather_f64(gf)).sum())  # rank order, fp64
    got_i, got_f = gpu_sums
    out["reduce_int_bit_exact"] = (got_i & 0xFFFFFFFF) == want_i
    out["reduce_float_rel_err"] = abs(got_f - want_f) / abs(want_f)
    out["float_tolerance"] = 1e-5
    out["reduce_e2e_int_bit_exact"] = (e2e_sums[0] & 0xFFFFFFFF) == (gi & 0xFFFFFFFF)  # host-buffer path: per rank
    out["reduce_e2e_float_rel_err"] = abs(e2e_sums[1] - gf) / abs(gf)
    out["reduce_sizes"] = "2^28 int + 2^28 float per rank (the timed inputs); float gold = reduceCPU<float> (Kahan) " \
                          "per shard, shard sums added in fp64 in rank order"
    for k in ("reduce_int_bit_exact", "reduce_e2e_int_bit_exact"):
        if not out[k]:
            fails.append(k)
    for k in ("reduce_float_rel_err", "reduce_e2e_float_rel_err"):
        if not out[k] <= 1e-5:
            fails.append(k)
    del hi, hf

    # ---- histogram64 + histogram256 sharded at the TIMED size, 4 GiB per rank (32 GiB over 8 GPUs = configs[3]):
    #      bins bit-exact vs histogram256CPU/64CPU, whose byteCount is a uint (histogram_gold.cpp:32,49), so the gold
    #      runs over 1 GiB chunks and the u32 bins are added up
    nb = HIST_BYTES_PER_RANK
    d = torch.empty(nb, dtype=torch.uint8, device="cuda")
    prim.generate(d, "bytes", rank * nb, nb, 2009)
    h256, h64 = mg.histogram_sharded(d, nb)
    g256, g64 = np.zeros(256, np.int64), np.zeros(64, np.int64)
    t0 = time.perf_counter()
    for c0 in range(0, nb, 1 << 30):
        hd = d[c0:c0 + (1 << 30)].cpu().numpy()
        g256 += gold.histogram256(hd).astype(np.int64)
        g64 += gold.histogram64(hd).astype(np.int64)
        del hd
    out["gold_histogram_seconds"] = time.perf_counter() - t0
    want256 = allsum_i64(g256) & 0xFFFFFFFF
    want64 = allsum_i64(g64) & 0xFFFFFFFF
    out["histogram256_bit_exact"] = bool((h256.cpu().numpy().view(np.uint32) == want256).all())
    out["histogram64_bit_exact"] = bool((h64.cpu().numpy().view(np.uint32) == want64).all())
    out["histogram_sizes"] = "%d GiB per rank, %d GiB in all: the size the histogram row times" % (nb >> 30,
                                                                                                  (nb >> 30) * world)
    for k in ("histogram256_bit_exact", "histogram64_bit_exact"):
        if not out[k]:
            fails.append(k)
    del d

    # ---- scanExclusiveLarge, the FULL 2^28-element array split into contiguous shards: element-wise bit-exact
    n = N_ELEMS // world
    src = torch.empty(n, dtype=torch.int32, device="cuda")
    dst = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(src, "u31", rank * n, n, 2009)
    mg.scan_sharded(dst, src, n)
    hs = src.cpu().numpy().view(np.uint32)
    totals = np.zeros(world, np.int64)
    totals[rank] = int(hs.sum(dtype=np.uint64)) & 0xFFFFFFFF
    totals = allsum_i64(totals)
    carry = int(totals[:rank].sum()) & 0xFFFFFFFF
    want = gold.scan_exclusive(hs, 1, n)  # scanExclusiveHost of this shard ...
    want = (want + np.uint32(carry)).astype(np.uint32)  # ... seeded with the lower shards' totals (u32 wrap)
    out["scan_bit_exact"] = bool((dst.cpu().numpy().view(np.uint32) == want).all())
    out["scan_sizes"] = "one array of 2^28 u32 over %d shard(s), every element compared" % world
    if not out["scan_bit_exact"]:
        fails.append("scan_bit_exact")
    del src, dst, hs, want

    # ---- key-value sort: exact vs the stable gold at world x 2^21 pairs (3 datasets x 2 directions), and the
    #      full-size run through size-independent properties
    ns = 1 << 21
    sort_ok = True
    for kind, param in (("key_unique", 0), ("key_tile_unique", 1 << 12), ("key_mod", 1000)):
        keys = oracle.generate(kind, 0, world * ns, 7, param)
        vals = np.arange(world * ns, dtype=np.uint32)
        for direction in (1, 0):
            sk = torch.from_numpy(keys[rank * ns:(rank + 1) * ns].view(np.int32).copy()).cuda()
            sv = torch.from_numpy(vals[rank * ns:(rank + 1) * ns].view(np.int32).copy()).cuda()
            dk, dv, n_out, m = mg.sort_sharded(sk, sv, ns, direction)
            wk, wv = oracle.stable_sort_kv(keys, vals, direction)
            off = 0 if m is None else int(m[:, :rank].sum())
            gk, gv = dk[:n_out].cpu().numpy().view(np.uint32), dv[:n_out].cpu().numpy().view(np.uint32)
            ok = (m is None or int(m.sum()) == world * ns) and bool((gk == wk[off:off + n_out]).all()) and \
                bool((gv == wv[off:off + n_out]).all())
            if kind == "key_unique" and direction == 1 and world * ns <= (1 << 22):
                # D1 is where mergeSortHost coincides with the stable order: check against the reference's own
                # emulation too (mergeSort_host.cpp:288-363)
                mk, mv = gold.merge_sort_host(keys, vals, 1)
                ok = ok and bool((mk == wk).all()) and bool((mv == wv).all())
            sort_ok = sort_ok and ok
    out["sort_bit_exact"] = sort_ok
    out["sort_sizes"] = "%d x 2^21 pairs, D1/D2/D3-like datasets, both directions, vs the stable gold" % world
    if not sort_ok:
        fails.append("sort_bit_exact")
    # full size: sortedness + stability inside every rank, order across rank boundaries, and a pair-wise checksum
    n = N_ELEMS // world
    sk = torch.empty(n, dtype=torch.int32, device="cuda")
    prim.generate(sk, "key_mod", rank * n, n, 2009, 1 << 20)  # ~256 copies of every key: stability is exercised
    sv = torch.arange(rank * n, rank * n + n, dtype=torch.int64, device="cuda").to(torch.int32)
    dk, dv, n_out, m = mg.sort_sharded(sk, sv, n, 1)
    k64, v64 = dk[:n_out].to(torch.int64) & 0xFFFFFFFF, dv[:n_out].to(torch.int64) & 0xFFFFFFFF
    inside = bool((k64[1:] >= k64[:-1]).all()) and bool(((k64[1:] > k64[:-1]) | (v64[1:] > v64[:-1])).all())
    edge = torch.tensor([int(k64[0]) if n_out else -1, int(k64[-1]) if n_out else -1,
                         int(v64[0]) if n_out else -1, int(v64[-1]) if n_out else -1, n_out],
                        dtype=torch.int64, device="cuda")
    edges = torch.empty(world * 5, dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_gather_into_tensor(edges, edge)
    else:
        edges.copy_(edge)
    e = edges.cpu().numpy().reshape(world, 5)
    across, last = True, None
    for r in range(world):
        if e[r, 4] == 0:
            continue
        if last is not None and not (e[r, 0] > last[0] or (e[r, 0] == last[0] and e[r, 2] > last[1])):
            across = False
        last = (e[r, 1], e[r, 3])

    def pair_hash(k, v):  # wrap-around sum of a per-pair mix: pairs must stay together and none may be lost
        x = (k * 0x9E3779B1 + v * 0x85EBCA77 + 0x165667B1) & 0xFFFFFFFFFFFF
        return int(((x ^ (x >> 15)) * 0x2545F491 & 0xFFFFFFFFFFFF).sum().item()) & 0xFFFFFFFFFFFF
    h_in = pair_hash(sk.to(torch.int64) & 0xFFFFFFFF, sv.to(torch.int64) & 0xFFFFFFFF)
    h_out = pair_hash(k64, v64)
    sums = allsum_i64([h_in, h_out, n_out])
    out["sort_full_size_properties"] = inside and across and \
        (int(sums[0]) & 0xFFFFFFFFFFFF) == (int(sums[1]) & 0xFFFFFFFFFFFF) and int(sums[2]) == N_ELEMS
    out["sort_full_sizes"] = "2^28 pairs in total, keys mod 2^20 (every key ~256 times), val = global index: sorted + " \
                             "ties in index order inside every rank and across rank boundaries, pair checksum and " \
                             "count preserved"
    if not out["sort_full_size_properties"]:
        fails.append("sort_full_size_properties")
    del sk, sv, dk, dv, k64, v64
    if mg.p2p_enabled():
        torch.cuda.synchronize()
        out["p2p_errors"] = prim.p2p_errors()
        if out["p2p_errors"]:
            fails.append("p2p_errors")
    out["failures_this_rank"] = fails
    out["failures"] = int(allsum_i64([len(fails)])[0])
    out["all_ok"] = out["failures"] == 0
    return out


# --------------------------------------------------------------------------------------------- ours
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import cuda_samples_b200 as prim
    from cuda_samples_b200 import multi_gpu as mg

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    barrier = (lambda: dist.barrier()) if world > 1 else None
    # N > 1: the small exchanges are done by the producing kernels themselves over NVLink peer windows
    # (csrc/p2p.cuh); B200_BENCH_P2P=0 falls back to one NCCL collective per step for comparison
    p2p = world > 1 and os.environ.get("B200_BENCH_P2P", "1") == "1" and mg.enable_p2p()
    peak, peak_src = measured_peak()

    # ---- inputs, resident in HBM (each rank generates its own slice of the global stream)
    xi = torch.empty(N_ELEMS, dtype=torch.int32, device="cuda")
    xf = torch.empty(N_ELEMS, dtype=torch.float32, device="cuda")
    prim.generate(xi, "i32_byte", rank * N_ELEMS, N_ELEMS, 1)
    prim.generate(xf, "f32_byte", rank * N_ELEMS, N_ELEMS, 1)
    res = {"pending": None, "last": None}

    def step():
        # both shard reductions, then ONE all_gather of the two partials; the collective of step i runs on NCCL's
        # stream while the kernels of step i+1 execute, and is consumed (summed in rank order) one step later
        p = mg.reduce_many_sharded([xi, xf], [N_ELEMS, N_ELEMS], overlap=OVERLAP)
        if res["pending"] is not None:
            res["pending"].result()
        res["last"] = res["pending"] or p
        res["pending"] = p

    def finish():
        if res["pending"] is not None:
            res["pending"].result()
            res["last"] = res["pending"]
            res["pending"] = None

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = prim.launch_count()
    seconds = time_steps(torch, step, args.steps, args.warmup, barrier, finish)
    launches = prim.launch_count() - launches0 - 2 * args.warmup
    seconds = max_over_ranks(torch, dist, seconds, world)
    # keep the same load running ~1.5 s so the 100 ms clock poller sees it (untimed); the iteration count is
    # derived from the max-over-ranks time so that every rank issues the same number of exchanges
    for _ in range(max(1, int(1.5 / (seconds / args.steps) / 50))):
        for _ in range(50):
            step()
        finish()
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    gpu_i, gpu_f = res["last"].typed()

    bytes_per_step = 2 * 4 * N_ELEMS * world
    value = bytes_per_step * args.steps / seconds / 1e9
    per_launch_s = seconds / (2 * args.steps)  # the timed region holds only reduce_kernel launches (+NCCL at N>1)
    achieved = 4 * N_ELEMS / per_launch_s / 1e9

    # ---- e2e: host buffers in, result out, copies inside the timed region (C ABI: b200_reduce_*_host).
    # The host inputs are pinned buffers from b200_host_alloc (on the GPU's NUMA node when the platform allows);
    # the PCIe roofline is measured in the same run with a plain pinned cudaMemcpyAsync of the same buffer.
    hbuf_i, hbuf_f = prim.HostBuffer(4 * N_ELEMS), prim.HostBuffer(4 * N_ELEMS)
    hi, hf = hbuf_i.array(np.int32), hbuf_f.array(np.float32)
    torch.from_numpy(hi).copy_(xi)
    torch.from_numpy(hf).copy_(xf)
    torch.cuda.synchronize()
    host_info = hbuf_i.info()
    e2e_steps = max(1, min(args.steps, 10))
    for _ in range(2):
        prim.reduce_host(hi), prim.reduce_host(hf)
    if barrier:
        barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e_i, e_f = prim.reduce_host(hi), prim.reduce_host(hf)
    e2e_s = time.perf_counter() - t0
    e2e_s = max_over_ranks(torch, dist, e2e_s, world)
    e2e_value = bytes_per_step * e2e_steps / e2e_s / 1e9
    if barrier:
        barrier()
    # PCIe roofline of the same run: every rank copies its 1 GiB pinned buffer 4 times back to back, all ranks between
    # the same two barriers; bytes of all ranks / the slowest rank's time = what this host feeds N GPUs AT ONCE
    # (per-rank bests summed would count moments when the other ranks had already finished)
    pcie_best = prim.host_probe_h2d(hi, 4 * N_ELEMS, 3)
    if barrier:
        barrier()
    pcie_s = max_over_ranks(torch, dist, prim.host_probe_h2d_sustained(xi, hi, 4 * N_ELEMS, 4), world)
    pcie_gbs = 4 * 4 * N_ELEMS * world / pcie_s / 1e9
    pcie_t = torch.tensor([pcie_best], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(pcie_t)
    pcie_sum_of_bests = float(pcie_t.item())

    # ---- parity of every primitive against the CPU gold, inside the run, at every N
    import oracle
    try:
        parity = parity_block(torch, dist, prim, mg, oracle, rank, world, xi, xf, (gpu_i, gpu_f), (e_i, e_f))
    except Exception as e:  # a crash in the checker must not hide the measurement -- but it is not a pass either
        parity = {"error": repr(e), "all_ok": False}
    hbuf_i.free()
    hbuf_f.free()
    del hi, hf

    table = None
    if not args.skip_table:
        del xi, xf
        torch.cuda.empty_cache()
        try:
            table = primitives_table(torch, dist, prim, mg, rank, world, peak, barrier, p2p)
        except Exception as e:  # the headline must still be printed
            table = {"error": repr(e)}

    if p2p:
        mg.disable_p2p()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        if parity.get("all_ok") is not True:  # `failures` is all-reduced: every rank knows, every rank says so
            sys.exit(3)
        return

    # ---- CPU baseline on rank 0 (N=1 only): the gold path, single thread as shipped, whole workload once
    cpu = None
    if world == 1 and "gold_reduce_seconds" in parity:
        rkind = parity["reduceCPU"]
        dt = parity["gold_reduce_seconds"]  # timed inside parity_block on the very inputs the kernels reduced
        cpu = {"value": 2 * 4 * N_ELEMS / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": rkind,
               "sample": "reduceCPU<int> + reduceCPU<float> (Kahan, reduction.cpp:167-180%s) once over the full "
                         "2^28-element inputs, single thread as shipped" %
                         (" compiled from the sample's own TU" if rkind == "reference" else " restated in oracle/oracle.c")}

    # ---- the other primitives' gold functions on the same host cores, 1 thread each
    cpu_gold = None
    if world == 1 and not args.skip_table:
        try:
            cpu_gold = cpu_gold_table(oracle, torch, prim)
        except Exception as e:
            cpu_gold = {"error": repr(e)}

    # dram bytes per launch of reduce_kernel from the committed `ncu --set full` capture -- only if that capture was of
    # the kernel that runs now (profiles/summarize.py stores a hash of the kernel's sources beside the figure)
    traffic, traffic_source = None, None
    try:
        sys.path.insert(0, os.path.join(ROOT, "profiles"))
        from summarize import reduce_source_hash
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        if tj.get("_reduce_source_sha256") == reduce_source_hash():
            traffic, traffic_source = tj.get("reduce_kernel"), tj.get("_source")
        else:
            traffic_source = "profiles/traffic.json was captured from other sources of reduce_kernel: figure withheld"
    except Exception as e:
        traffic_source = "unavailable: %r" % (e,)
    cfg = base_config(world)  # identical to the reference arm's config: the same workload, nothing else in here
    notes = {"clock_probe": "same step looped ~1.5 s after the timed region for the 100 ms nvidia-smi poller"}
    if world > 1:
        notes["exchange"] = ("fused into the last reduce kernel of the step: P2P stores into NVLink peer windows + "
                             "rank-order fp64 sum (csrc/p2p.cuh), no NCCL call in the timed region") if p2p else \
            "one NCCL all_gather of the two shard partials per step"
    line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": seconds / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "i32+f32", "data": "synthetic", "config": cfg,
            "roofline": {"bound": "hbm", "kernel": "reduce_kernel<int|float>", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                         "frac_of_nominal_8TBs": achieved / HBM_NOMINAL, "traffic": traffic, "traffic_source": traffic_source,
                         "algorithmic_bytes_per_launch": 4 * N_ELEMS},
            "cpu_baseline": cpu, "cpu_gold": cpu_gold, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": 2 * 4 * N_ELEMS * world,
                    "d2h_bytes_per_step": 8 * world, "steps": e2e_steps,
                    "api": "b200_reduce_i32_host + b200_reduce_f32_host per rank (pinned host input from "
                           "b200_host_alloc, H2D chunks overlapped with the kernels, result word read back)",
                    "pcie_gbs": pcie_gbs, "frac_of_pcie": e2e_value / pcie_gbs if pcie_gbs else None,
                    "pcie_probe": "plain pinned cudaMemcpyAsync H2D of the same 1 GiB buffer, 4 copies back to back on "
                                  "all %d rank(s) between two barriers: total bytes / slowest rank's time" % world,
                    "pcie_gbs_sum_of_per_rank_bests": pcie_sum_of_bests,
                    "host_buffer": host_info},
            "gpu_launches": int(launches), "parity": parity, "notes": notes, "primitives": table}
    # The verdict in a key the driver keeps, and in the exit code (the samples' own drivers return their gold compare
    # as the process status: histogram/main.cpp:221-227, scan/main.cpp:181, mergeSort/main.cpp:100-118).
    problems = list(parity.get("failures_this_rank", [])) if parity.get("all_ok") is not True else []
    if parity.get("all_ok") is not True and not problems:
        problems.append(parity.get("error", "parity failed on another rank"))
    for name, row in (table or {}).items() if isinstance(table, dict) else ():
        if isinstance(row, dict) and row.get("lost_words_in_timed_loop"):
            problems.append("%s: lost_words_in_timed_loop=%s" % (name, row["lost_words_in_timed_loop"]))
    if isinstance(table, dict) and "error" in table:
        problems.append("primitives table: " + str(table["error"]))
    line["roofline"]["parity_ok"] = not problems
    line["roofline"]["parity_failures"] = problems
    line["roofline"]["parity_checked"] = "every primitive vs the samples' CPU gold inside this run, on all %d rank(s)" \
                                         " (see `parity`)" % world
    weakest = weakest_primitive(table)
    if weakest:
        line["roofline"]["weakest_primitive"] = weakest
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    if problems:
        sys.exit(3)


NORTH_STAR_ROWS = ("histogram256_64MiB", "histogram64_64MiB", "histogram64+256_fused_4GiB_per_gpu",
                   "scanExclusiveLarge_2^28", "scanExclusiveShort_L1024_2^28", "mergeSort_kv_2^28")


def weakest_primitive(table):
    """The north-star primitive furthest below its HBM roofline in this run (the headline `roofline` describes the
    reduce, the easiest kernel of the path)."""
    worst = None
    for name in NORTH_STAR_ROWS:
        row = table.get(name) if isinstance(table, dict) else None
        frac = row.get("frac_of_measured_peak") if isinstance(row, dict) else None
        if isinstance(frac, (int, float)) and (worst is None or frac < worst["frac"]):
            worst = {"name": name, "frac": frac, "n_gpus": row.get("n_gpus")}
    return worst


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner, torchrun notices) write to fd 1; the contract is ONE JSON line on
    stdout, so everything else is sent to stderr and the line is written to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--skip-table", action="store_true", help="headline only (no per-primitive table)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
