"""Host-side mirror of the samples' entry points (same names, same argument meaning) over the C ABI.

Device buffers are anything with a `data_ptr()` (torch CUDA tensors) or a raw integer device pointer;
work is enqueued on torch's current CUDA stream.  PyTorch is only plumbing here (device memory,
streams); the arithmetic happens in libb200prim.so.

Reference interfaces mirrored (paths relative to /root/reference/Samples):
  histogram64 / histogram256 / init* / close*   2_Concepts_and_Techniques/histogram/histogram_common.h:75-82
  reduce(size, threads, blocks, whichKernel, d_idata, d_odata)
                                                2_Concepts_and_Techniques/reduction/reduction.h:32
  scanExclusiveShort / scanExclusiveLarge / initScan / closeScan
                                                2_Concepts_and_Techniques/scan/scan_common.h:41-55
  mergeSort / initMergeSort / closeMergeSort    0_Introduction/mergeSort/mergeSort_common.h:49-54
"""
import ctypes

import numpy as np

from ._lib import B200Error, check, const_u32, lib

HISTOGRAM64_BIN_COUNT = 64
HISTOGRAM256_BIN_COUNT = 256
MAX_BATCH_ELEMENTS = const_u32("MAX_BATCH_ELEMENTS")
MIN_SHORT_ARRAY_SIZE = const_u32("MIN_SHORT_ARRAY_SIZE")
MAX_SHORT_ARRAY_SIZE = const_u32("MAX_SHORT_ARRAY_SIZE")
MIN_LARGE_ARRAY_SIZE = const_u32("MIN_LARGE_ARRAY_SIZE")
MAX_LARGE_ARRAY_SIZE = const_u32("MAX_LARGE_ARRAY_SIZE")

GEN_KINDS = {"bytes": 0, "u31": 1, "i32_byte": 2, "f32_byte": 3, "key_unique": 4, "key_tile_unique": 5,
             "key_mod": 6, "f64_byte": 7}


def _dptr(x):
    if x is None:
        return None
    if isinstance(x, int):
        return ctypes.c_void_p(x)
    if hasattr(x, "data_ptr"):
        if hasattr(x, "is_cuda") and not x.is_cuda:
            raise B200Error("expected a CUDA tensor, got a host tensor")
        if hasattr(x, "is_contiguous") and not x.is_contiguous():
            raise B200Error("expected a contiguous tensor")
        return ctypes.c_void_p(x.data_ptr())
    raise TypeError("device buffer must be a CUDA tensor or an integer device pointer, got %r" % type(x))


def _hptr(x):
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise B200Error("expected a C-contiguous array")
        return ctypes.c_void_p(x.ctypes.data)
    if hasattr(x, "data_ptr"):
        if getattr(x, "is_cuda", False):
            raise B200Error("expected a host tensor, got a CUDA tensor")
        return ctypes.c_void_p(x.data_ptr())
    if isinstance(x, int):
        return ctypes.c_void_p(x)
    raise TypeError(type(x))


def _stream(stream=None):
    if stream is None:
        import torch
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return ctypes.c_void_p(stream)
    return ctypes.c_void_p(stream.cuda_stream)


def launch_count():
    return int(lib.b200_launch_count())


def device_info():
    v = [ctypes.c_int(0) for _ in range(4)]
    check(lib.b200_device_info(*[ctypes.byref(x) for x in v]), "b200_device_info")
    return {"num_sms": v[0].value, "smem_optin": v[1].value, "cc": (v[2].value, v[3].value)}


# ------------------------------------------------------------------------------------ histogram
def initHistogram64():
    lib.initHistogram64()


def initHistogram256():
    lib.initHistogram256()


def closeHistogram64():
    lib.closeHistogram64()


def closeHistogram256():
    lib.closeHistogram256()


def histogram256(d_Histogram, d_Data, byteCount, stream=None):
    """d_Histogram[256] (u32) <- byte counts of d_Data[0:byteCount]; overwritten, wraps mod 2^32."""
    check(lib.b200_histogram256_async(_dptr(d_Histogram), _dptr(d_Data), byteCount, _stream(stream)), "histogram256")


def histogram64(d_Histogram, d_Data, byteCount, stream=None):
    """d_Histogram[64] (u32) <- counts of (byte >> 2)."""
    check(lib.b200_histogram64_async(_dptr(d_Histogram), _dptr(d_Data), byteCount, _stream(stream)), "histogram64")


def histogram_fused(d_Histogram256, d_Histogram64, d_Data, byteCount, stream=None):
    """Both histograms from one pass over the data."""
    check(lib.b200_histogram_fused_async(_dptr(d_Histogram256), _dptr(d_Histogram64), _dptr(d_Data), byteCount,
                                         _stream(stream)), "histogram_fused")


# ------------------------------------------------------------------------------------ reduction
_REDUCE = {"torch.int32": "i32", "torch.float32": "f32", "torch.float64": "f64",
           "int32": "i32", "float32": "f32", "float64": "f64"}


def reduce(size, threads, blocks, whichKernel, d_idata, d_odata, stream=None):
    """reduce<T>: afterwards sum(d_odata[0:blocks]) == sum(d_idata[0:size]); here d_odata[0] holds the
    total and d_odata[1:blocks] zeros.  `threads` / `whichKernel` are accepted and ignored."""
    kind = _REDUCE.get(str(d_idata.dtype))
    if kind is None:
        raise TypeError("reduce: unsupported dtype %s" % d_idata.dtype)
    fn = getattr(lib, "b200_reduce_%s_async" % kind)
    check(fn(_dptr(d_idata), size, _dptr(d_odata), max(int(blocks), 1), _stream(stream)), "reduce<%s>" % kind)


def reduce_partial(d_idata, size, d_slot_f64, stream=None):
    """Shard partial of a sharded reduction: sum(d_idata[0:size]) in its own type, stored as a double."""
    kind = {"i32": 0, "f32": 1, "f64": 2}.get(_REDUCE.get(str(d_idata.dtype)))
    if kind is None:
        raise TypeError("reduce_partial: unsupported dtype %s" % d_idata.dtype)
    check(lib.b200_reduce_partial_async(kind, _dptr(d_idata), size, _dptr(d_slot_f64), _stream(stream)),
          "reduce_partial")


def reduceSinglePass(size, threads, blocks, d_idata, d_odata, stream=None):
    """threadFenceReduction / reductionMultiBlockCG entry point (float): d_odata[0] = sum, rest zero."""
    check(lib.b200_reduce_f32_async(_dptr(d_idata), size, _dptr(d_odata), max(int(blocks), 1), _stream(stream)),
          "reduceSinglePass")


def scalarProd(d_C, d_A, d_B, vectorN, elementN, stream=None):
    """scalarProd sample: d_C[v] = dot(d_A[v], d_B[v]) for vectorN vectors of elementN floats."""
    check(lib.b200_scalar_prod_async(_dptr(d_C), _dptr(d_A), _dptr(d_B), vectorN, elementN, _stream(stream)),
          "scalarProd")


# ----------------------------------------------------------------------------------------- scan
def initScan():
    lib.initScan()


def closeScan():
    lib.closeScan()


def scanExclusiveShort(d_Dst, d_Src, batchSize, arrayLength, stream=None):
    check(lib.b200_scan_exclusive_async(_dptr(d_Dst), _dptr(d_Src), batchSize, arrayLength, None, _stream(stream)),
          "scanExclusiveShort")
    return 512


def scanExclusiveLarge(d_Dst, d_Src, batchSize, arrayLength, stream=None, d_carry=None):
    """d_carry (optional, device u32): seed of element 0 -- the shard carry of the multi-GPU scan."""
    check(lib.b200_scan_exclusive_async(_dptr(d_Dst), _dptr(d_Src), batchSize, arrayLength, _dptr(d_carry),
                                        _stream(stream)), "scanExclusiveLarge")
    return 512


def scanInclusive(d_Dst, d_Src, batchSize, arrayLength, stream=None, d_carry=None):
    """Inclusive +scan (the shfl_scan sample's result): d_Dst[j] = sum d_Src[0..j] per array, wrapping."""
    check(lib.b200_scan_inclusive_async(_dptr(d_Dst), _dptr(d_Src), batchSize, arrayLength, _dptr(d_carry),
                                        _stream(stream)), "scanInclusive")


def filter_arr(dst, nres, src, n, stream=None):
    """warpAggregatedAtomicsCG's filter_arr(dst, nres, src, n): keeps src[i] > 0 -- here in input order."""
    check(lib.b200_compact_positive_async(_dptr(dst), _dptr(nres), _dptr(src), n, _stream(stream)), "filter_arr")


def integral_image(d_out, d_img, width, height, stream=None):
    """Summed-area table of a u8 image (shfl_scan sample, shuffle_integral_image_test): d_out u32[height*width]."""
    check(lib.b200_integral_image_u8_async(_dptr(d_out), _dptr(d_img), width, height, _stream(stream)),
          "integral_image")


def scan_total(d_total, d_Src, n, stream=None):
    check(lib.b200_scan_total_async(_dptr(d_total), _dptr(d_Src), n, _stream(stream)), "scan_total")


# ------------------------------------------------ fused compute + exchange over NVLink peer windows
def p2p_window_create():
    """-> the 64-byte CUDA IPC handle of this rank's exchange window (created on first call)."""
    buf = ctypes.create_string_buffer(64)
    check(lib.b200_p2p_window_create(buf), "p2p_window_create")
    return buf.raw


def p2p_connect(rank, world, handles):
    """handles: list of `world` 64-byte IPC handles in rank order (every rank's p2p_window_create())."""
    blob = b"".join(handles)
    assert len(blob) == 64 * world
    check(lib.b200_p2p_connect(rank, world, blob), "p2p_connect")


def p2p_disconnect():
    check(lib.b200_p2p_disconnect(), "p2p_disconnect")


def p2p_world():
    return int(lib.b200_p2p_world())


def p2p_errors():
    n = ctypes.c_ulonglong(0)
    check(lib.b200_p2p_errors(ctypes.byref(n)), "p2p_errors")
    return int(n.value)


def p2p_allgather(d_payload, nbytes, d_gathered, stream=None):
    check(lib.b200_p2p_allgather_async(_dptr(d_payload), nbytes, _dptr(d_gathered), _stream(stream)), "p2p_allgather")


def p2p_buffer_create(nbytes):
    """-> (device pointer as int, 64-byte IPC handle) of a cudaMalloc'ed buffer other ranks can map."""
    ptr, buf = ctypes.c_void_p(0), ctypes.create_string_buffer(64)
    check(lib.b200_p2p_buffer_create(nbytes, ctypes.byref(ptr), buf), "p2p_buffer_create")
    return int(ptr.value), buf.raw


def p2p_buffer_open(handle):
    ptr = ctypes.c_void_p(0)
    check(lib.b200_p2p_buffer_open(handle, ctypes.byref(ptr)), "p2p_buffer_open")
    return int(ptr.value)


def p2p_buffer_close(ptr):
    check(lib.b200_p2p_buffer_close(ctypes.c_void_p(ptr)), "p2p_buffer_close")


def p2p_buffer_free(ptr):
    check(lib.b200_p2p_buffer_free(ctypes.c_void_p(ptr)), "p2p_buffer_free")


def partition_exchange_kv(srcKey, srcVal, N, d_splitters, sortDir, recv_keys, recv_vals, capacity, d_matrix, d_info,
                          stream=None):
    """Stable partition by splitters whose bucket runs are stored straight into the destination ranks' receive
    buffers (recv_keys / recv_vals: lists of `world` int device pointers).  See include/b200prim.h."""
    w = len(recv_keys)
    ka = (ctypes.c_void_p * w)(*recv_keys)
    va = (ctypes.c_void_p * w)(*recv_vals) if recv_vals is not None else None  # None: recv_keys are pair buffers
    check(lib.b200_partition_exchange_kv_async(_dptr(srcKey), _dptr(srcVal), N, _dptr(d_splitters), sortDir, ka, va,
                                               capacity, _dptr(d_matrix), _dptr(d_info), _stream(stream)),
          "partition_exchange_kv")


def reduce_partial_exchange(d_idata, size, d_slots, index, count, d_gathered, d_sums, d_carry=None, stream=None):
    """reduce -> exchange in one kernel: fills d_slots[index], then all ranks' slots land in d_gathered
    ([world][count] doubles) and their rank-order sums in d_sums ([count])."""
    kind = {"i32": 0, "f32": 1, "f64": 2}.get(_REDUCE.get(str(d_idata.dtype)))
    if kind is None:
        raise TypeError("reduce_partial_exchange: unsupported dtype %s" % d_idata.dtype)
    check(lib.b200_reduce_partial_exchange_async(kind, _dptr(d_idata), size, _dptr(d_slots), index, count,
                                                 _dptr(d_gathered), _dptr(d_sums), _dptr(d_carry), _stream(stream)),
          "reduce_partial_exchange")


def scan_total_exchange(d_carry, d_Src, n, d_work, stream=None):
    """Shard total -> peers -> d_carry[0] = wrapping sum of the lower ranks' totals.  d_work: float64[>= 2 + world]."""
    check(lib.b200_scan_total_exchange_async(_dptr(d_carry), _dptr(d_Src), n, _dptr(d_work), _stream(stream)),
          "scan_total_exchange")


def scan_exclusive_shard(d_Dst, d_Src, n, d_work, d_carry, stream=None):
    """One shard of a sharded exclusive scan, both steps (tile sums + exchange, then the scan of independent tiles)."""
    check(lib.b200_scan_exclusive_shard_async(_dptr(d_Dst), _dptr(d_Src), n, _dptr(d_work), _dptr(d_carry),
                                              _stream(stream)), "scan_exclusive_shard")


def histogram_fused_allreduce(d_Histogram256, d_Histogram64, d_Data, byteCount, stream=None):
    """Both histograms of ALL ranks' shards on every rank; the bins are exchanged by the kernel itself."""
    check(lib.b200_histogram_fused_allreduce_async(_dptr(d_Histogram256), _dptr(d_Histogram64), _dptr(d_Data),
                                                   byteCount, _stream(stream)), "histogram_fused_allreduce")


# ------------------------------------------------------------------------------------ mergeSort
def initMergeSort():
    lib.initMergeSort()


def closeMergeSort():
    lib.closeMergeSort()


def mergeSort(dstKey, dstVal, bufKey, bufVal, srcKey, srcVal, N, sortDir, stream=None):
    """Stable key-value sort; result in dst*, buf* scratch, src* untouched; sortDir 1 asc / 0 desc."""
    check(lib.b200_sort_kv_async(_dptr(dstKey), _dptr(dstVal), _dptr(bufKey), _dptr(bufVal), _dptr(srcKey),
                                 _dptr(srcVal), N, sortDir, _stream(stream)), "mergeSort")


def sort_kv_dyn(dstKey, dstVal, bufKey, bufVal, srcKey, srcVal, capacity, d_n, sortDir, stream=None):
    """mergeSort over the first min(d_n[0], capacity) pairs, d_n read on the device (no host round trip)."""
    check(lib.b200_sort_kv_dyn_async(_dptr(dstKey), _dptr(dstVal), _dptr(bufKey), _dptr(bufVal), _dptr(srcKey),
                                     _dptr(srcVal), capacity, _dptr(d_n), sortDir, _stream(stream)), "sort_kv_dyn")


def bitonicSort(d_DstKey, d_DstVal, d_SrcKey, d_SrcVal, batchSize, arrayLength, dir, stream=None):
    """sortingNetworks sample: batchSize independent key-value sorts of arrayLength pairs (stable, any length)."""
    check(lib.b200_segmented_sort_kv_async(_dptr(d_DstKey), _dptr(d_DstVal), _dptr(d_SrcKey), _dptr(d_SrcVal),
                                           batchSize, arrayLength, dir, _stream(stream)), "bitonicSort")
    return 256


oddEvenMergeSort = bitonicSort


def partition_kv(dstKey, dstVal, srcKey, srcVal, N, d_splitters, numBuckets, sortDir, d_bucketCounts, stream=None):
    check(lib.b200_partition_kv_async(_dptr(dstKey), _dptr(dstVal), _dptr(srcKey), _dptr(srcVal), N,
                                      _dptr(d_splitters), numBuckets, sortDir, _dptr(d_bucketCounts),
                                      _stream(stream)), "partition_kv")


# ------------------------------------------------------------------------------------- generator
def generate(d_out, kind, first, n, seed, param=0, stream=None):
    check(lib.b200_generate_async(_dptr(d_out), first, n, seed, GEN_KINDS[kind], param, _stream(stream)), "generate")


# ---------------------------------------------------------------- host-buffer (end-to-end) variants
def histogram256_host(h_data):
    h = np.empty(256, np.uint32)
    n = h_data.nbytes if isinstance(h_data, np.ndarray) else h_data.numel() * h_data.element_size()
    check(lib.b200_histogram256_host(_hptr(h), _hptr(h_data), n), "histogram256_host")
    return h


def histogram64_host(h_data):
    h = np.empty(64, np.uint32)
    n = h_data.nbytes if isinstance(h_data, np.ndarray) else h_data.numel() * h_data.element_size()
    check(lib.b200_histogram64_host(_hptr(h), _hptr(h_data), n), "histogram64_host")
    return h


def reduce_host(h_in):
    name = str(h_in.dtype)
    kind = _REDUCE.get(name)
    if kind is None:
        raise TypeError("reduce_host: unsupported dtype %s" % name)
    n = h_in.size if isinstance(h_in, np.ndarray) else h_in.numel()
    out = np.zeros(1, {"i32": np.int32, "f32": np.float32, "f64": np.float64}[kind])
    check(getattr(lib, "b200_reduce_%s_host" % kind)(_hptr(h_in), n, _hptr(out)), "reduce_host")
    return out[0].item()


def scan_exclusive_host(h_dst, h_src, batchSize, arrayLength):
    check(lib.b200_scan_exclusive_host(_hptr(h_dst), _hptr(h_src), batchSize, arrayLength), "scan_exclusive_host")
    return h_dst


def sort_kv_host(h_dstKey, h_dstVal, h_srcKey, h_srcVal, N, sortDir):
    check(lib.b200_sort_kv_host(_hptr(h_dstKey), _hptr(h_dstVal), _hptr(h_srcKey), _hptr(h_srcVal), N, sortDir),
          "sort_kv_host")
    return h_dstKey, h_dstVal


# ------------------------------------------------------------------ pinned, NUMA-local host buffers
class HostBuffer:
    """Pinned host memory from b200_host_alloc (registered + device-mapped, on the current GPU's NUMA node when the
    platform allows).  `.array(dtype)` is a numpy view; free() (or the destructor) releases it."""

    def __init__(self, nbytes):
        p = ctypes.c_void_p(0)
        check(lib.b200_host_alloc(ctypes.byref(p), nbytes), "b200_host_alloc")
        self.ptr, self.nbytes = int(p.value), int(nbytes)

    def array(self, dtype=np.uint8):
        dt = np.dtype(dtype)
        raw = (ctypes.c_char * self.nbytes).from_address(self.ptr)
        a = np.frombuffer(raw, dtype=dt, count=self.nbytes // dt.itemsize)
        a.flags.writeable = True
        return a

    def info(self):
        v = [ctypes.c_int(-1) for _ in range(4)]
        check(lib.b200_host_alloc_info(ctypes.c_void_p(self.ptr), *[ctypes.byref(x) for x in v]), "b200_host_alloc_info")
        return {"gpu_numa_node": v[0].value, "mem_numa_node_first_page": v[1].value,
                "mem_numa_node_last_page": v[2].value,
                "placement": {0: "none possible", 1: "mbind", 2: "first touch on the node's CPUs"}.get(v[3].value)}

    def free(self):
        if self.ptr:
            check(lib.b200_host_free(ctypes.c_void_p(self.ptr)), "b200_host_free")
            self.ptr = 0

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def host_probe_h2d(h_src, nbytes, reps=5):
    """Plain pinned cudaMemcpyAsync host->device of nbytes, best of reps -> GB/s (the PCIe roofline of *_host)."""
    g = ctypes.c_double(0.0)
    check(lib.b200_host_probe_h2d(_hptr(h_src), nbytes, reps, ctypes.byref(g)), "b200_host_probe_h2d")
    return float(g.value)


def host_probe_h2d_sustained(d_dst, h_src, nbytes, reps=4):
    """reps back-to-back pinned H2D copies of nbytes into the device tensor d_dst -> seconds (first issue to last done)."""
    t = ctypes.c_double(0.0)
    check(lib.b200_host_probe_h2d_sustained(d_dst.data_ptr(), _hptr(h_src), nbytes, reps, ctypes.byref(t)),
          "b200_host_probe_h2d_sustained")
    return float(t.value)


def tune_host_path(mode, chunk_bytes=0):
    """Experiments only: mode 0 = staged through device chunks (default), 1 = zero-copy reads of pinned input."""
    check(lib.b200_tune_host_path(mode, chunk_bytes), "b200_tune_host_path")


# -------------------------------------------- sharded primitives, ONE process driving all devices (C ABI)
class ShardedContext:
    """b200_ctx: peer access + exchange windows + one stream per device, created once; the sharded entry points take
    one buffer per device.  Thin ctypes mirror of include/b200prim.h (csrc/sharded.cu); buffers are CUDA tensors (or
    integer device pointers) living on the matching device."""

    def __init__(self, devices):
        self.devices = list(devices)
        arr = (ctypes.c_int * len(self.devices))(*self.devices)
        h = ctypes.c_void_p(0)
        check(lib.b200_ctx_create(len(self.devices), arr, ctypes.byref(h)), "b200_ctx_create")
        self.handle = h

    def _ptrs(self, bufs):
        assert len(bufs) == len(self.devices)
        return (ctypes.c_void_p * len(bufs))(*[_dptr(b).value or 0 if b is not None else 0 for b in bufs])

    def _sizes(self, ns):
        assert len(ns) == len(self.devices)
        return (ctypes.c_size_t * len(ns))(*[int(n) for n in ns])

    def synchronize(self):
        check(lib.b200_ctx_synchronize(self.handle), "b200_ctx_synchronize")

    def histogram(self, hist256, hist64, data, nbytes):
        check(lib.b200_histogram_sharded(self.handle, self._ptrs(hist256), self._ptrs(hist64), self._ptrs(data),
                                         self._sizes(nbytes)), "b200_histogram_sharded")

    def reduce(self, xs, ns):
        kind = _REDUCE.get(str(xs[0].dtype))
        out = np.zeros(1, {"i32": np.int32, "f32": np.float32, "f64": np.float64}[kind])
        check(getattr(lib, "b200_reduce_sharded_%s" % kind)(self.handle, self._ptrs(xs), self._sizes(ns), _hptr(out)),
              "b200_reduce_sharded_%s" % kind)
        return out[0].item()

    def scan_exclusive(self, dsts, srcs, ns):
        check(lib.b200_scan_exclusive_sharded(self.handle, self._ptrs(dsts), self._ptrs(srcs), self._sizes(ns)),
              "b200_scan_exclusive_sharded")

    def sort_kv(self, dst_keys, dst_vals, capacity, src_keys, src_vals, ns, direction=1):
        """-> list of pairs received per device; raises B200Error (status 2) when `capacity` is too small."""
        n_out = (ctypes.c_size_t * len(self.devices))()
        check(lib.b200_sort_kv_sharded(self.handle, self._ptrs(dst_keys), self._ptrs(dst_vals), capacity,
                                       self._ptrs(src_keys), self._ptrs(src_vals), self._sizes(ns), direction, n_out),
              "b200_sort_kv_sharded")
        return [int(x) for x in n_out]

    def destroy(self):
        if self.handle:
            check(lib.b200_ctx_destroy(self.handle), "b200_ctx_destroy")
            self.handle = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def sort_splitters(srcKey, N, sortDir, d_splitters, stream=None):
    """Splitters of the sharded sort chosen on the device (one peer-window exchange); every rank must call it."""
    check(lib.b200_sort_splitters_async(_dptr(srcKey), N, sortDir, _dptr(d_splitters), _stream(stream)),
          "sort_splitters")


def scan_reserve(n):
    check(lib.b200_scan_reserve(n), "b200_scan_reserve")


def sort_reserve(n):
    check(lib.b200_sort_reserve(n), "b200_sort_reserve")


def p2p_probe_store(d_dst, d_src, nbytes, mode, reps=5, stream=None):
    """GB/s of writing nbytes from d_src (local) into d_dst (peer-mapped or local): mode 0 4-byte lane stores,
    1 16-byte lane stores, 2 cp.async.bulk from shared memory, 3 cudaMemcpyAsync."""
    g = ctypes.c_double(0.0)
    check(lib.b200_p2p_probe_store(_dptr(d_dst), _dptr(d_src), nbytes, mode, reps, ctypes.byref(g), _stream(stream)),
          "b200_p2p_probe_store")
    return float(g.value)


def p2p_set_timeout_ms(ms):
    check(lib.b200_p2p_set_timeout_ms(ms), "b200_p2p_set_timeout_ms")
