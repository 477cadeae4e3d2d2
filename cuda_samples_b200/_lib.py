"""ctypes binding of libb200prim.so (the C ABI declared in include/b200prim.h and the drop-in headers).

There is no CPU fallback: if the shared library is missing the import fails loudly, and every
entry point raises on a non-zero status.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200PRIM_LIB") or os.path.join(_HERE, "libb200prim.so")  # override: debug builds only

vp = ctypes.c_void_p
sz = ctypes.c_size_t
u32 = ctypes.c_uint32
u64 = ctypes.c_uint64
ci = ctypes.c_int

# name -> (restype, argtypes); every int-returning b200_* function is status-checked
SIGNATURES = {
    "b200_version": (ci, []),
    "b200_error_string": (ctypes.c_char_p, [ci]),
    "b200_device_info": (ci, [ctypes.POINTER(ci)] * 4),
    "b200_launch_count": (ctypes.c_ulonglong, []),
    "b200_release_scratch": (ci, []),
    "b200_histogram256_async": (ci, [vp, vp, sz, vp]),
    "b200_histogram64_async": (ci, [vp, vp, sz, vp]),
    "b200_histogram_fused_async": (ci, [vp, vp, vp, sz, vp]),
    "b200_reduce_i32_async": (ci, [vp, sz, vp, sz, vp]),
    "b200_reduce_f32_async": (ci, [vp, sz, vp, sz, vp]),
    "b200_reduce_f64_async": (ci, [vp, sz, vp, sz, vp]),
    "b200_reduce_partial_async": (ci, [ci, vp, sz, vp, vp]),
    "b200_scalar_prod_async": (ci, [vp, vp, vp, ci, ci, vp]),
    "b200_scan_exclusive_async": (ci, [vp, vp, sz, sz, vp, vp]),
    "b200_scan_total_async": (ci, [vp, vp, sz, vp]),
    "b200_scan_exclusive_noalloc_async": (ci, [vp, vp, sz, sz, vp, vp]),
    "b200_scan_inclusive_async": (ci, [vp, vp, sz, sz, vp, vp]),
    "b200_compact_positive_async": (ci, [vp, vp, vp, sz, vp]),
    "b200_integral_image_u8_async": (ci, [vp, vp, sz, sz, vp]),
    "b200_sort_kv_async": (ci, [vp, vp, vp, vp, vp, vp, sz, u32, vp]),
    "b200_sort_kv_dyn_async": (ci, [vp, vp, vp, vp, vp, vp, sz, vp, u32, vp]),
    "b200_partition_kv_async": (ci, [vp, vp, vp, vp, sz, vp, u32, u32, vp, vp]),
    "b200_segmented_sort_kv_async": (ci, [vp, vp, vp, vp, sz, sz, u32, vp]),
    "b200_generate_async": (ci, [vp, sz, sz, u64, ci, u32, vp]),
    "b200_histogram256_host": (ci, [vp, vp, sz]),
    "b200_histogram64_host": (ci, [vp, vp, sz]),
    "b200_reduce_i32_host": (ci, [vp, sz, vp]),
    "b200_reduce_f32_host": (ci, [vp, sz, vp]),
    "b200_reduce_f64_host": (ci, [vp, sz, vp]),
    "b200_scan_exclusive_host": (ci, [vp, vp, sz, sz]),
    "b200_sort_kv_host": (ci, [vp, vp, vp, vp, sz, u32]),
    "b200_p2p_window_create": (ci, [ctypes.c_char_p]),
    "b200_p2p_connect": (ci, [ci, ci, ctypes.c_char_p]),
    "b200_p2p_disconnect": (ci, []),
    "b200_p2p_world": (ci, []),
    "b200_p2p_errors": (ci, [ctypes.POINTER(ctypes.c_ulonglong)]),
    "b200_p2p_barrier_async": (ci, [vp]),
    "b200_p2p_allgather_async": (ci, [vp, sz, vp, vp]),
    "b200_reduce_partial_exchange_async": (ci, [ci, vp, sz, vp, ctypes.c_uint, ctypes.c_uint, vp, vp, vp, vp]),
    "b200_scan_total_exchange_async": (ci, [vp, vp, sz, vp, vp]),
    "b200_scan_exclusive_shard_async": (ci, [vp, vp, sz, vp, vp, vp]),
    "b200_tune_scan_shard_tiled": (ci, [ci]),
    "b200_histogram_fused_allreduce_async": (ci, [vp, vp, vp, sz, vp]),
    "b200_p2p_buffer_create": (ci, [sz, ctypes.POINTER(vp), ctypes.c_char_p]),
    "b200_p2p_buffer_open": (ci, [ctypes.c_char_p, ctypes.POINTER(vp)]),
    "b200_p2p_buffer_close": (ci, [vp]),
    "b200_p2p_buffer_free": (ci, [vp]),
    "b200_partition_exchange_kv_async": (ci, [vp, vp, sz, vp, u32, ctypes.POINTER(vp), ctypes.POINTER(vp), sz, vp, vp,
                                              vp]),
    "b200_scan_reserve": (ci, [sz]),
    "b200_sort_reserve": (ci, [sz]),
    "b200_p2p_set_timeout_ms": (ci, [ctypes.c_uint]),
    "b200_p2p_probe_store": (ci, [vp, vp, sz, ci, ci, ctypes.POINTER(ctypes.c_double), vp]),
    "b200_sort_splitters_async": (ci, [vp, sz, u32, vp, vp]),
    "b200_ctx_create": (ci, [ci, ctypes.POINTER(ci), ctypes.POINTER(vp)]),
    "b200_ctx_destroy": (ci, [vp]),
    "b200_ctx_num_devices": (ci, [vp]),
    "b200_ctx_stream": (ci, [vp, ci, ctypes.POINTER(vp)]),
    "b200_ctx_synchronize": (ci, [vp]),
    "b200_histogram_sharded": (ci, [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(vp),
                                    ctypes.POINTER(sz)]),
    "b200_reduce_sharded_i32": (ci, [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp]),
    "b200_reduce_sharded_f32": (ci, [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp]),
    "b200_reduce_sharded_f64": (ci, [vp, ctypes.POINTER(vp), ctypes.POINTER(sz), vp]),
    "b200_scan_exclusive_sharded": (ci, [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), ctypes.POINTER(sz)]),
    "b200_sort_kv_sharded": (ci, [vp, ctypes.POINTER(vp), ctypes.POINTER(vp), sz, ctypes.POINTER(vp),
                                  ctypes.POINTER(vp), ctypes.POINTER(sz), u32, ctypes.POINTER(sz)]),
    "b200_host_alloc": (ci, [ctypes.POINTER(vp), sz]),
    "b200_host_free": (ci, [vp]),
    "b200_host_alloc_info": (ci, [vp] + [ctypes.POINTER(ci)] * 4),
    "b200_host_probe_h2d": (ci, [vp, sz, ci, ctypes.POINTER(ctypes.c_double)]),
    "b200_host_probe_h2d_sustained": (ci, [vp, vp, sz, ci, ctypes.POINTER(ctypes.c_double)]),
    "b200_tune_host_path": (ci, [ci, sz]),
    "b200_tune_hist_cluster": (ci, [ci]),
    "b200_tune_hist_launch": (ci, [ci, ci]),
    "b200_tune_scan_total_reverse": (ci, [ci]),
    "b200_tune_scan_pdl": (ci, [ci]),
    "b200_tune_scan_tma": (ci, [ci]),
    "b200_tune_scan_backoff": (ci, [ci]),
    "b200_tune_sort_lookback": (ci, [ci]),
    "b200_tune_sort_pairs": (ci, [ci]),
    "b200_tune_sort_plan": (ci, [ci]),
    "b200_tune_sort_peer_bulk": (ci, [ci]),
    "b200_tune_scan_reload": (ci, [ci]),
    # drop-in entry points (the samples' own signatures)
    "initHistogram64": (None, []), "closeHistogram64": (None, []), "histogram64": (None, [vp, vp, u32]),
    "initHistogram256": (None, []), "closeHistogram256": (None, []), "histogram256": (None, [vp, vp, u32]),
    "initScan": (None, []), "closeScan": (None, []),
    "scanExclusiveShort": (sz, [vp, vp, u32, u32]), "scanExclusiveLarge": (sz, [vp, vp, u32, u32]),
    "initMergeSort": (None, []), "closeMergeSort": (None, []),
    "mergeSort": (None, [vp, vp, vp, vp, vp, vp, u32, u32]),
    "reduce": (None, [ci, ci, ci, vp, vp]), "reduceSinglePass": (None, [ci, ci, ci, vp, vp]),
    "reduceSinglePassMultiBlockCG": (None, [ci, ci, ci, vp, vp]), "scalarProd": (None, [vp, vp, vp, ci, ci]),
    "bitonicSort": (u32, [vp, vp, vp, vp, u32, u32, u32]), "oddEvenMergeSort": (None, [vp, vp, vp, vp, u32, u32, u32]),
    "_Z6reduceIiEviiiiPT_S1_": (None, [ci, ci, ci, ci, vp, vp]),
    "_Z6reduceIfEviiiiPT_S1_": (None, [ci, ci, ci, ci, vp, vp]),
    "_Z6reduceIdEviiiiPT_S1_": (None, [ci, ci, ci, ci, vp, vp]),
}
DATA_SYMBOLS = ["MAX_BATCH_ELEMENTS", "MIN_SHORT_ARRAY_SIZE", "MAX_SHORT_ARRAY_SIZE", "MIN_LARGE_ARRAY_SIZE",
                "MAX_LARGE_ARRAY_SIZE"]


class B200Error(RuntimeError):
    pass


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "cuda_samples_b200: %s is missing -- build it with `python cuda_samples_b200/build.py` "
            "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


def check(status, what):
    if status != 0:
        msg = lib.b200_error_string(status)
        raise B200Error("%s failed: CUDA error %d (%s)" % (what, status, msg.decode() if msg else "?"))


def const_u32(name):
    return ctypes.c_uint32.in_dll(lib, name).value
