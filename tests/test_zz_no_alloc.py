"""b200_scan_exclusive_noalloc_async: the scan for callers that are not allowed to allocate device memory.

Some programs that link libb200prim.so cannot call cudaMalloc once they are running: a plug-in whose host application
reserved the whole GPU for its own memory pool, a service with a frozen memory budget, a real-time loop that must not
hit the device-wide synchronisation inside cudaMalloc / cudaFree.  For them the library has one scan entry point whose
look-back descriptors sit in a fixed array inside the library image instead of cudaMalloc'ed scratch
(include/b200prim.h).  The reference has nothing like it, so the yardstick here is the library's own allocating entry
point, b200_scan_exclusive_async (itself bit-exact against scanExclusiveHost in test_gpu_parity.py): integer work, so
the two must agree bit for bit.

What is checked, through the C ABI (ctypes), with torch only lending device memory and the stream:
  * two calls in a row with different data and different shapes -- the second call finds the first call's descriptors
    and ticket state in the shared array and must not be confused by them;
  * the largest request the array can serve (8192 tiles = one array of 2^26 elements) and the refusal of one tile more;
  * that a call leaves the amount of free device memory alone.
This file is named so that pytest collects it last: under `pytest -x` a failure here hides nothing else.  It passes
on its own in a fresh process (its first call does the per-device set-up) and after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

MAX_TILES = 8192  # B200_SCAN_NOALLOC_MAX_TILES
TILE = 8192


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from cuda_samples_b200 import _lib
    return _lib.lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _scan(fn, src, batch, length, carry=None):
    """Run one scan entry point on a device tensor; the output starts as garbage so every element must be written."""
    dst = torch.full_like(src, 0x5A5A5A5A)
    st = fn(ctypes.c_void_p(dst.data_ptr()), ctypes.c_void_p(src.data_ptr()), batch, length,
            ctypes.c_void_p(carry.data_ptr() if carry is not None else None), _stream())
    assert st == 0, st
    return dst


def _random_u32(n, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    # int32 storage, full 32-bit patterns: the scan wraps mod 2^32 whatever the sign interpretation
    return torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int32, device="cuda", generator=g)


# (batchSize, arrayLength): chained arrays with a ragged last tile, several medium arrays, packed short arrays,
# one array shorter than a tile with a non-power-of-two length
SHAPES = [(1, 1_000_003), (3, 100_000), (64, 1024), (5, 4099), (1, 3 * TILE)]


def test_two_calls_in_a_row_match_the_allocating_entry_point(lib):
    for k, (batch, length) in enumerate(SHAPES):
        a = _random_u32(batch * length, seed=100 + k)
        b = _random_u32(batch * length, seed=200 + k)
        want_a = _scan(lib.b200_scan_exclusive_async, a, batch, length)
        want_b = _scan(lib.b200_scan_exclusive_async, b, batch, length)
        # back to back on one stream, no synchronisation in between: call 2 starts from whatever call 1 left behind
        got_a = _scan(lib.b200_scan_exclusive_noalloc_async, a, batch, length)
        got_b = _scan(lib.b200_scan_exclusive_noalloc_async, b, batch, length)
        torch.cuda.synchronize()
        assert torch.equal(got_a, want_a), (batch, length, "first call")
        assert torch.equal(got_b, want_b), (batch, length, "second call")
    # and across shapes: a long chained array right after short ones, then the same data again
    a = _random_u32(2_500_000, seed=7)
    want = _scan(lib.b200_scan_exclusive_async, a, 1, a.numel())
    first = _scan(lib.b200_scan_exclusive_noalloc_async, a, 1, a.numel())
    second = _scan(lib.b200_scan_exclusive_noalloc_async, a, 1, a.numel())
    torch.cuda.synchronize()
    assert torch.equal(first, want) and torch.equal(second, want)
    # independent check of one case on the host (numpy's uint32 cumsum wraps the same way)
    h = a.cpu().numpy().view(np.uint32)
    ref = np.concatenate(([0], np.cumsum(h[:-1], dtype=np.uint32))).astype(np.uint32)
    assert np.array_equal(second.cpu().numpy().view(np.uint32), ref)


def test_carry_and_in_place(lib):
    a = _random_u32(777_777, seed=11)
    carry = torch.tensor([0x9E3779B9 - 2**32], dtype=torch.int32, device="cuda")
    want = _scan(lib.b200_scan_exclusive_async, a, 1, a.numel(), carry)
    got = _scan(lib.b200_scan_exclusive_noalloc_async, a, 1, a.numel(), carry)
    buf = a.clone()
    st = lib.b200_scan_exclusive_noalloc_async(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(buf.data_ptr()), 1,
                                               buf.numel(), ctypes.c_void_p(carry.data_ptr()), _stream())
    assert st == 0
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    assert torch.equal(buf, want)


def test_largest_request_and_refusal_beyond_it(lib):
    n = MAX_TILES * TILE  # 2^26 elements: every descriptor of the built-in array is in use
    a = _random_u32(n, seed=3)
    want = _scan(lib.b200_scan_exclusive_async, a, 1, n)
    got1 = _scan(lib.b200_scan_exclusive_noalloc_async, a, 1, n)
    got2 = _scan(lib.b200_scan_exclusive_noalloc_async, a, 1, n)
    torch.cuda.synchronize()
    assert torch.equal(got1, want) and torch.equal(got2, want)
    del got1, got2, want
    # one element more needs tile 8193: refused with cudaErrorInvalidValue, nothing launched, output untouched
    big = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    out = torch.full((n + 1,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    launches = lib.b200_launch_count()
    for batch, length in ((1, n + 1), (MAX_TILES + 1, TILE), (2, MAX_TILES * TILE // 2 + 1), (2**33, 1 << 20)):
        st = lib.b200_scan_exclusive_noalloc_async(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(big.data_ptr()),
                                                   batch, length, None, _stream())
        assert st == 1, (batch, length, st)  # cudaErrorInvalidValue
    assert lib.b200_launch_count() == launches
    torch.cuda.synchronize()
    assert int(out[0]) == 0x5A5A5A5A and int(out[-1]) == 0x5A5A5A5A
    # empty requests are accepted and do nothing, like the allocating entry point
    assert lib.b200_scan_exclusive_noalloc_async(None, None, 0, 1024, None, _stream()) == 0
    assert lib.b200_scan_exclusive_noalloc_async(None, None, 4, 0, None, _stream()) == 0


def test_a_call_takes_no_device_memory(lib):
    a = _random_u32(4_000_000, seed=5)
    dst = torch.empty_like(a)
    args = (ctypes.c_void_p(dst.data_ptr()), ctypes.c_void_p(a.data_ptr()), 1, a.numel(), None, _stream())
    assert lib.b200_scan_exclusive_noalloc_async(*args) == 0  # set-up and code loading, if still to do, happen here
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.b200_scan_exclusive_noalloc_async(*args) == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, (free_before, free_after)
