"""CPU oracle for the primitive hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (cuda_samples_b200/) never does.

Two checkers live here:
  * `oracle.*`  -- ctypes view of oracle/liboracle.so, the C restatement in oracle/oracle.c
                   (each function cites the reference file:line it follows);
  * `oracle.ref`-- ctypes view of oracle/_ref/libref_gold.so, the reference's own gold sources
                   compiled unmodified (histogram64CPU, histogram256CPU, scanExclusiveHost,
                   mergeSortHost, validateSortedKeys, validateSortedValues, fillValues), or None
                   when that file was never built;
  * `oracle.ref_reduce` -- the reference's own reduceCPU<T> / getNumBlocksAndThreads (they live inside the sample's
                   driver TU reduction.cpp, compiled where it lies through oracle/ref_reduce_wrap.cpp into
                   oracle/_ref/libref_reduce.so), or None.
Parity status: pinned (tests/test_oracle.py: restatement == reference gold on the samples'
seeded inputs, and both == the known answers in tests/golden/kat.json).
"""
import contextlib
import ctypes
import os
import sys

import numpy as np

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))

c_u32p = ctypes.POINTER(ctypes.c_uint32)
c_void = ctypes.c_void_p
u64 = ctypes.c_uint64
u32 = ctypes.c_uint32


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def _load():
    path = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "oracle.c")):
        path = _build.build_oracle()
    lib = ctypes.CDLL(path)
    lib.oracle_histogram256.argtypes = [c_void, c_void, u64]
    lib.oracle_histogram64.argtypes = [c_void, c_void, u64]
    lib.oracle_scan_exclusive.argtypes = [c_void, c_void, u64, u64, u32]
    lib.oracle_reduce_i32.argtypes = [c_void, u64]
    lib.oracle_reduce_i32.restype = ctypes.c_int32
    lib.oracle_reduce_f32.argtypes = [c_void, u64]
    lib.oracle_reduce_f32.restype = ctypes.c_float
    lib.oracle_reduce_f64.argtypes = [c_void, u64]
    lib.oracle_reduce_f64.restype = ctypes.c_double
    lib.oracle_scalar_prod.argtypes = [c_void, c_void, c_void, ctypes.c_int, ctypes.c_int]
    lib.oracle_stable_sort_kv.argtypes = [c_void, c_void, c_void, c_void, u64, u32]
    lib.oracle_merge_sort_host.argtypes = [c_void] * 6 + [u64, u32]
    lib.oracle_validate_sorted_keys.argtypes = [c_void, c_void, u64, u64, u32, u32]
    lib.oracle_validate_sorted_keys.restype = u32
    lib.oracle_validate_sorted_values.argtypes = [c_void, c_void, c_void, u64, u64, ctypes.POINTER(ctypes.c_int)]
    lib.oracle_validate_sorted_values.restype = ctypes.c_int
    lib.oracle_fill_values.argtypes = [c_void, u64]
    lib.oracle_fill_sample_input.argtypes = [c_void, u64, u32, ctypes.c_int]
    lib.oracle_generate.argtypes = [c_void, u64, u64, u64, ctypes.c_int, u32]
    lib.oracle_scan_inclusive.argtypes = [c_void, c_void, u64]
    lib.oracle_filter_positive.argtypes = [c_void, c_void, u64]
    lib.oracle_filter_positive.restype = u64
    lib.oracle_integral_image_u8.argtypes = [c_void, c_void, u32, u32]
    lib.oracle_fnv1a64.argtypes = [c_void, u64]
    lib.oracle_fnv1a64.restype = u64
    return lib


_lib = _load()


# ---------------------------------------------------------------------------------- restatement
def histogram256(data):
    data = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    h = np.empty(256, np.uint32)
    _lib.oracle_histogram256(_ptr(h), _ptr(data), data.size)
    return h


def histogram64(data):
    data = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    h = np.empty(64, np.uint32)
    _lib.oracle_histogram64(_ptr(h), _ptr(data), data.size)
    return h


def scan_exclusive(src, batch, length, carry=0):
    src = np.ascontiguousarray(src, dtype=np.uint32).reshape(-1)
    assert src.size == batch * length
    dst = np.empty_like(src)
    _lib.oracle_scan_exclusive(_ptr(dst), _ptr(src), batch, length, carry)
    return dst


def scan_inclusive(src):
    """Running sum of a u32/i32 array (shfl_scan.cu:157-161 CPUverify), wrapping."""
    src = np.ascontiguousarray(src).view(np.uint32)
    dst = np.empty_like(src)
    _lib.oracle_scan_inclusive(_ptr(dst), _ptr(src), src.size)
    return dst


def filter_positive(src):
    """Elements > 0 in input order (warpAggregatedAtomicsCG.cu:287-292 host loop)."""
    src = np.ascontiguousarray(src, dtype=np.int32)
    dst = np.empty_like(src)
    n = _lib.oracle_filter_positive(_ptr(dst), _ptr(src), src.size)
    return dst[:n].copy()


def integral_image(img):
    """Summed-area table of a 2-D u8 image (shfl_scan.cu:308-381 restated), u32, wrapping."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape
    dst = np.empty((h, w), np.uint32)
    if h and w:
        _lib.oracle_integral_image_u8(_ptr(dst), _ptr(img), w, h)
    return dst


def reduce(data):
    data = np.ascontiguousarray(data).reshape(-1)
    assert data.size > 0
    if data.dtype == np.int32:
        return int(_lib.oracle_reduce_i32(_ptr(data), data.size))
    if data.dtype == np.float32:
        return float(_lib.oracle_reduce_f32(_ptr(data), data.size))
    if data.dtype == np.float64:
        return float(_lib.oracle_reduce_f64(_ptr(data), data.size))
    raise TypeError(data.dtype)


def scalar_prod(a, b, vector_n, element_n):
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
    assert a.size == b.size == vector_n * element_n
    c = np.empty(vector_n, np.float32)
    _lib.oracle_scalar_prod(_ptr(c), _ptr(a), _ptr(b), vector_n, element_n)
    return c


def stable_sort_kv(keys, vals, direction=1):
    keys = np.ascontiguousarray(keys, dtype=np.uint32)
    vals = np.ascontiguousarray(vals, dtype=np.uint32)
    dk, dv = np.empty_like(keys), np.empty_like(vals)
    _lib.oracle_stable_sort_kv(_ptr(dk), _ptr(dv), _ptr(keys), _ptr(vals), keys.size, direction)
    return dk, dv


def merge_sort_host(keys, vals, direction=1):
    keys = np.ascontiguousarray(keys, dtype=np.uint32)
    vals = np.ascontiguousarray(vals, dtype=np.uint32)
    dk, dv = np.empty_like(keys), np.empty_like(vals)
    bk, bv = np.empty_like(keys), np.empty_like(vals)
    _lib.oracle_merge_sort_host(_ptr(dk), _ptr(dv), _ptr(bk), _ptr(bv), _ptr(keys), _ptr(vals), keys.size, direction)
    return dk, dv


def validate_sorted_keys(res_key, src_key, num_values, direction=1):
    res_key = np.ascontiguousarray(res_key, dtype=np.uint32)
    src_key = np.ascontiguousarray(src_key, dtype=np.uint32)
    return int(_lib.oracle_validate_sorted_keys(_ptr(res_key), _ptr(src_key), 1, res_key.size, num_values, direction))


def validate_sorted_values(res_key, res_val, src_key):
    """-> (correct, stable)"""
    res_key = np.ascontiguousarray(res_key, dtype=np.uint32)
    res_val = np.ascontiguousarray(res_val, dtype=np.uint32)
    src_key = np.ascontiguousarray(src_key, dtype=np.uint32)
    st = ctypes.c_int(0)
    ok = _lib.oracle_validate_sorted_values(_ptr(res_key), _ptr(res_val), _ptr(src_key), 1, res_key.size,
                                            ctypes.byref(st))
    return int(ok), int(st.value)


def fill_values(n):
    v = np.empty(n, np.uint32)
    _lib.oracle_fill_values(_ptr(v), n)
    return v


_SAMPLE_MODES = {"hist_bytes": (0, np.uint8), "scan_u32": (1, np.uint32), "reduce_i32": (2, np.int32),
                 "reduce_f32": (3, np.float32), "sort_keys": (4, np.uint32), "reduce_f64": (5, np.float64)}


def sample_input(kind, n, seed):
    """The samples' own glibc-rand() inputs (see oracle.c:oracle_fill_sample_input)."""
    mode, dt = _SAMPLE_MODES[kind]
    out = np.empty(n, dt)
    _lib.oracle_fill_sample_input(_ptr(out), n, seed, mode)
    return out


GEN_KINDS = {"bytes": (0, np.uint8), "u31": (1, np.uint32), "i32_byte": (2, np.int32), "f32_byte": (3, np.float32),
             "key_unique": (4, np.uint32), "key_tile_unique": (5, np.uint32), "key_mod": (6, np.uint32),
             "f64_byte": (7, np.float64)}


def generate(kind, first, n, seed, param=0):
    """Counter-based generator (CPU twin of cuda_samples_b200.generate)."""
    k, dt = GEN_KINDS[kind]
    out = np.empty(n, dt)
    _lib.oracle_generate(_ptr(out), first, n, seed, k, param)
    return out


def fnv1a64(arr):
    arr = np.ascontiguousarray(arr)
    return int(_lib.oracle_fnv1a64(_ptr(arr), arr.nbytes))


# ------------------------------------------------------------------------ reference gold (as is)
@contextlib.contextmanager
def _quiet_stdout():
    """mergeSortHost / the validators printf progress lines (mergeSort_host.cpp:310,318)."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    try:
        os.dup2(devnull, 1)
        yield
    finally:
        ctypes.CDLL(None).fflush(None)
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)


class _RefGold:
    """The reference's own compiled gold code (uint lengths, host pointers)."""

    def __init__(self, path):
        lib = ctypes.CDLL(path)
        self.path = path
        self._lib = lib
        lib.histogram64CPU.argtypes = [c_void, c_void, u32]
        lib.histogram256CPU.argtypes = [c_void, c_void, u32]
        lib.scanExclusiveHost.argtypes = [c_void, c_void, u32, u32]
        lib.mergeSortHost.argtypes = [c_void] * 6 + [u32, u32]
        lib.validateSortedKeys.argtypes = [c_void, c_void, u32, u32, u32, u32]
        lib.validateSortedKeys.restype = u32
        lib.validateSortedValues.argtypes = [c_void, c_void, c_void, u32, u32]
        lib.validateSortedValues.restype = ctypes.c_int
        lib.fillValues.argtypes = [c_void, u32]
        if hasattr(lib, "scalarProdCPU"):
            lib.scalarProdCPU.argtypes = [c_void, c_void, c_void, ctypes.c_int, ctypes.c_int]

    def histogram256(self, data):
        data = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        assert data.size % 4 == 0 and data.size < 2**32
        h = np.empty(256, np.uint32)
        self._lib.histogram256CPU(_ptr(h), _ptr(data), data.size)
        return h

    def histogram64(self, data):
        data = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        assert data.size % 4 == 0 and data.size < 2**32
        h = np.empty(64, np.uint32)
        self._lib.histogram64CPU(_ptr(h), _ptr(data), data.size)
        return h

    def scan_exclusive(self, src, batch, length):
        src = np.ascontiguousarray(src, dtype=np.uint32).reshape(-1)
        assert src.size == batch * length and length >= 1
        dst = np.empty_like(src)
        self._lib.scanExclusiveHost(_ptr(dst), _ptr(src), batch, length)
        return dst

    def scalar_prod(self, a, b, vector_n, element_n):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
        c = np.empty(vector_n, np.float32)
        self._lib.scalarProdCPU(_ptr(c), _ptr(a), _ptr(b), vector_n, element_n)
        return c

    def merge_sort_host(self, keys, vals, direction=1):
        keys = np.ascontiguousarray(keys, dtype=np.uint32)
        vals = np.ascontiguousarray(vals, dtype=np.uint32)
        dk, dv = np.empty_like(keys), np.empty_like(vals)
        bk, bv = np.empty_like(keys), np.empty_like(vals)
        with _quiet_stdout():
            self._lib.mergeSortHost(_ptr(dk), _ptr(dv), _ptr(bk), _ptr(bv), _ptr(keys), _ptr(vals), keys.size,
                                    direction)
        return dk, dv

    def validate_sorted_keys(self, res_key, src_key, num_values, direction=1):
        res_key = np.ascontiguousarray(res_key, dtype=np.uint32)
        src_key = np.ascontiguousarray(src_key, dtype=np.uint32)
        with _quiet_stdout():
            return int(self._lib.validateSortedKeys(_ptr(res_key), _ptr(src_key), 1, res_key.size, num_values,
                                                    direction))

    def validate_sorted_values(self, res_key, res_val, src_key):
        res_key = np.ascontiguousarray(res_key, dtype=np.uint32)
        res_val = np.ascontiguousarray(res_val, dtype=np.uint32)
        src_key = np.ascontiguousarray(src_key, dtype=np.uint32)
        with _quiet_stdout():
            return int(self._lib.validateSortedValues(_ptr(res_key), _ptr(res_val), _ptr(src_key), 1, res_key.size))


def _load_ref():
    path = os.path.join(_HERE, "_ref", "libref_gold.so")
    if not os.path.exists(path):
        if not _build.have_reference():
            return None
        path = _build.build_ref_gold()
    try:
        return _RefGold(path)
    except OSError:
        return None


ref = _load_ref()


class _RefReduce:
    """The reference's own reduceCPU<T> (reduction.cpp:167-180), compiled from the sample's driver TU where it lies
    (oracle/ref_reduce_wrap.cpp).  `size` is an int like in the sample: n < 2^31."""

    def __init__(self, path):
        lib = ctypes.CDLL(path)
        self.path = path
        self._lib = lib
        lib.ref_reduceCPU_int.argtypes = [c_void, ctypes.c_int]
        lib.ref_reduceCPU_int.restype = ctypes.c_int32
        lib.ref_reduceCPU_float.argtypes = [c_void, ctypes.c_int]
        lib.ref_reduceCPU_float.restype = ctypes.c_float
        lib.ref_reduceCPU_double.argtypes = [c_void, ctypes.c_int]
        lib.ref_reduceCPU_double.restype = ctypes.c_double
        lib.ref_getNumBlocksAndThreads.argtypes = [ctypes.c_int] * 4 + [ctypes.POINTER(ctypes.c_int)] * 2

    def reduce(self, data):
        data = np.ascontiguousarray(data).reshape(-1)
        assert 0 < data.size < 2**31
        if data.dtype == np.int32:
            return int(self._lib.ref_reduceCPU_int(_ptr(data), data.size))
        if data.dtype == np.float32:
            return float(self._lib.ref_reduceCPU_float(_ptr(data), data.size))
        if data.dtype == np.float64:
            return float(self._lib.ref_reduceCPU_double(_ptr(data), data.size))
        raise TypeError(data.dtype)

    def num_blocks_and_threads(self, which_kernel, n, max_blocks, max_threads):
        """getNumBlocksAndThreads (reduction.cpp:204-240); reads the device properties: needs a GPU."""
        b, t = ctypes.c_int(0), ctypes.c_int(0)
        self._lib.ref_getNumBlocksAndThreads(which_kernel, n, max_blocks, max_threads, ctypes.byref(b),
                                             ctypes.byref(t))
        return b.value, t.value


def _load_ref_reduce():
    path = os.path.join(_HERE, "_ref", "libref_reduce.so")
    if not os.path.exists(path):
        try:
            path = _build.build_ref_reduce()
        except Exception:
            return None
        if not path or not os.path.exists(path):
            return None
    try:
        return _RefReduce(path)
    except OSError:
        return None


ref_reduce = _load_ref_reduce()
