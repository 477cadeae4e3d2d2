// host_path.cu -- host-buffer entry points (b200_*_host): what a CPU caller of histogram256CPU /
// reduceCPU / scanExclusiveHost / mergeSortHost (host pointers in, host pointers out) switches to.
// The host<->device copies live INSIDE the call: inputs are streamed through two device staging
// chunks so the H2D copy engine, the SMs and the D2H copy engine overlap; bench.py's `e2e` figure
// times exactly these functions.  They block until the result is in host memory.
#include "common.cuh"
#include "internal.h"

#include <type_traits>

namespace b200 {

int reduce_chunk(int kind, const void *d_in, size_t n, void *d_out, Scratch *scratch, const void *chain_in,
                 void *chain_out, cudaStream_t stream); // reduce.cu

static size_t g_chunk_bytes = 64u << 20; // per staging buffer
// 0: stage the input through device memory (H2D copy engine + kernel per chunk) -- the default: measured at
//    0.99 of a plain pinned cudaMemcpyAsync (55.1-55.4 of 55.6 GB/s, profiles/r02a_e2e_probe.log)
// 1: zero-copy -- ONE kernel reads the pinned host buffer over PCIe itself (no staging copy, no HBM round trip):
//    51.4 GB/s, the SMs' 128-byte read requests use the link less well than the copy engine's;
//    pageable (unregistered) input always takes the staged path
static int g_host_mode = 0;
constexpr size_t kChunkBytes = 32u << 20; // scan staging (in + out halves)

struct Pipe {
    DeviceCtx   *c       = nullptr;
    cudaStream_t compute = nullptr, h2d = nullptr, d2h = nullptr;
    cudaEvent_t  ready[2], freed[2], done[2];
};

// The host-buffer entry points own their streams, staging chunks, reduce scratch and result slot (all persistent
// in the device context), so a call allocates nothing and never synchronises the whole device.  Contract: they
// block until the result is in host memory, and they must not run concurrently with each other on one device.
static int open_pipe(Pipe &p, size_t stage_bytes)
{
    int st = 0;
    p.c    = ctx(&st);
    if (!p.c)
        return st;
    DeviceCtx *c = p.c;
    if (!c->copy_stream[0]) {
        B200_TRY(cudaStreamCreateWithFlags(&c->copy_stream[0], cudaStreamNonBlocking));
        B200_TRY(cudaStreamCreateWithFlags(&c->copy_stream[1], cudaStreamNonBlocking));
        B200_TRY(cudaStreamCreateWithFlags(&c->compute_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 6; i++)
            B200_TRY(cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming));
        B200_TRY(cudaHostAlloc(&c->host_slot, 4096, cudaHostAllocMapped | cudaHostAllocPortable));
    }
    p.h2d     = c->copy_stream[0];
    p.d2h     = c->copy_stream[1];
    p.compute = c->compute_stream;
    bool fresh = false;
    for (int i = 0; i < 2; i++) {
        p.ready[i] = c->ev[i];
        p.freed[i] = c->ev[2 + i];
        p.done[i]  = c->ev[4 + i];
        if (stage_bytes) {
            bool f = false;
            B200_TRY((cudaError_t)ensure(c->stage[i], stage_bytes, p.compute, &f));
            fresh |= f;
        }
    }
    if (fresh)
        B200_TRY(cudaStreamSynchronize(p.compute)); // scratch memsets done before the copy streams touch it
    return 0;
}

// device-visible address of a host buffer when it is pinned / registered (zero-copy candidate), else nullptr
static const void *mapped_ptr(const void *h)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, h) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

template <typename T> __global__ void sum_rows_kernel(const T *rows, size_t nrows, size_t width, T *out)
{
    // out[j] = sum_i rows[i][j]  (fixed order: deterministic)
    for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < width; j += (size_t)gridDim.x * blockDim.x) {
        T s = T(0);
        for (size_t i = 0; i < nrows; i++)
            s += rows[i * width + j];
        out[j] = s;
    }
}
__global__ void next_carry_kernel(unsigned *carry_next, const unsigned *carry, const unsigned *dst_last,
                                  const unsigned *src_last)
{
    // exclusive scan value of the last element + the element itself = inclusive total incl. carry
    (void)carry;
    *carry_next = *dst_last + *src_last;
}

// Streams `nbytes` of host input through the two staging chunks; per chunk calls
// launch(chunk_index, d_chunk, chunk_bytes) on the compute stream.
template <typename F> static int stream_input(Pipe &p, const void *h_in, size_t nbytes, size_t chunk, F launch)
{
    const char *src = reinterpret_cast<const char *>(h_in);
    size_t      k   = 0;
    for (size_t off = 0; off < nbytes; off += chunk, k++) {
        const int    b = (int)(k & 1);
        const size_t m = nbytes - off < chunk ? nbytes - off : chunk;
        if (k >= 2)
            B200_TRY(cudaStreamWaitEvent(p.h2d, p.freed[b], 0));
        B200_TRY(cudaMemcpyAsync(p.c->stage[b].ptr, src + off, m, cudaMemcpyHostToDevice, p.h2d));
        B200_TRY(cudaEventRecord(p.ready[b], p.h2d));
        B200_TRY(cudaStreamWaitEvent(p.compute, p.ready[b], 0));
        int st = launch(k, p.c->stage[b].ptr, m);
        if (st)
            return st;
        B200_TRY(cudaEventRecord(p.freed[b], p.compute));
    }
    return 0;
}

// reduce: the typed result is written by the (last) kernel straight into the mapped pinned slot -- the only thing
// that crosses the bus on the way back is that one word, and there is no D2H copy call.
template <typename T> static int reduce_host(int kind, const T *h_in, size_t n, T *h_result)
{
    Pipe        p;
    const void *zc = g_host_mode == 1 ? mapped_ptr(h_in) : nullptr;
    B200_TRY((cudaError_t)open_pipe(p, zc ? 0 : g_chunk_bytes));
    DeviceCtx *c = p.c;
    // host_work: [0, 64 KiB) reduce scratch (ticket + per-CTA partials) | two 8-byte chain slots
    B200_TRY((cudaError_t)ensure(c->host_work, (64u << 10) + 4096, p.compute));
    char *chain     = reinterpret_cast<char *>(c->host_work.ptr) + (64u << 10);
    T    *slot      = reinterpret_cast<T *>(c->host_slot);
    T    *slot_dev  = nullptr;
    B200_TRY(cudaHostGetDevicePointer((void **)&slot_dev, slot, 0));
    int st = 0;
    if (n == 0) {
        *h_result = T(0);
        return 0;
    }
    if (zc) {
        st = reduce_chunk(kind, zc, n, slot_dev, &c->host_work, nullptr, nullptr, p.compute);
    }
    else {
        const size_t chunk   = g_chunk_bytes;
        const size_t nchunks = (n * sizeof(T) + chunk - 1) / chunk;
        st = stream_input(p, h_in, n * sizeof(T), chunk, [&](size_t k, void *d, size_t bytes) {
            return reduce_chunk(kind, d, bytes / sizeof(T), k + 1 == nchunks ? slot_dev : nullptr, &c->host_work,
                                k ? chain + 8 * ((k - 1) & 1) : nullptr, chain + 8 * (k & 1), p.compute);
        });
    }
    if (st)
        return st;
    B200_TRY(cudaStreamSynchronize(p.compute));
    *h_result = *slot;
    return 0;
}

template <typename Launch> static int hist_host(uint32_t *h_hist, int bins, const void *h_data, size_t n, Launch hist_async)
{
    Pipe        p;
    const void *zc = g_host_mode == 1 ? mapped_ptr(h_data) : nullptr;
    B200_TRY((cudaError_t)open_pipe(p, zc ? 0 : g_chunk_bytes));
    DeviceCtx   *c       = p.c;
    const size_t chunk   = g_chunk_bytes;
    const size_t nchunks = zc ? 1 : (n + chunk - 1) / chunk;
    // rows live behind the reduce scratch of host_work
    const size_t rows_off = (64u << 10) + 4096;
    B200_TRY((cudaError_t)ensure(c->host_work, rows_off + (nchunks + 1) * bins * sizeof(unsigned), p.compute));
    unsigned *d_part = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(c->host_work.ptr) + rows_off);
    unsigned *d_total = d_part + nchunks * bins;
    int       st      = 0;
    if (zc || n == 0) {
        st = hist_async(d_total, zc ? zc : h_data, n, (b200_stream_t)p.compute);
    }
    else {
        st = stream_input(p, h_data, n, chunk, [&](size_t k, void *d, size_t bytes) {
            return hist_async(d_part + k * bins, d, bytes, (b200_stream_t)p.compute);
        });
        if (!st) {
            sum_rows_kernel<unsigned><<<1, 256, 0, p.compute>>>(d_part, nchunks, bins, d_total);
            count_launch();
            st = (int)cudaGetLastError();
        }
    }
    if (st)
        return st;
    B200_TRY(cudaMemcpyAsync(h_hist, d_total, bins * sizeof(unsigned), cudaMemcpyDeviceToHost, p.compute));
    B200_TRY(cudaStreamSynchronize(p.compute));
    return 0;
}

// One array (or a run of whole arrays) of `n` elements scanned chunk by chunk with a device-side carry.
// stage[b] = [ in chunk | out chunk ].
static int scan_host_run(Pipe &p, uint32_t *h_dst, const uint32_t *h_src, size_t n, size_t batch_in_chunk,
                         size_t L, unsigned *d_carry /* 2 u32, [0] zeroed */)
{
    const size_t half = kChunkBytes; // bytes of the in half; out half follows
    const size_t elems_per_chunk = batch_in_chunk ? batch_in_chunk * L : half / 4;
    size_t       k = 0;
    for (size_t off = 0; off < n; off += elems_per_chunk, k++) {
        const int    b = (int)(k & 1);
        const size_t m = n - off < elems_per_chunk ? n - off : elems_per_chunk;
        char        *d_in  = reinterpret_cast<char *>(p.c->stage[b].ptr);
        char        *d_out = d_in + half;
        if (k >= 2)
            B200_TRY(cudaStreamWaitEvent(p.h2d, p.freed[b], 0)); // scan of chunk k-2 has consumed the in half
        B200_TRY(cudaMemcpyAsync(d_in, h_src + off, m * 4, cudaMemcpyHostToDevice, p.h2d));
        B200_TRY(cudaEventRecord(p.ready[b], p.h2d));
        B200_TRY(cudaStreamWaitEvent(p.compute, p.ready[b], 0));
        if (k >= 2)
            B200_TRY(cudaStreamWaitEvent(p.compute, p.done[b], 0)); // D2H of chunk k-2 has drained the out half
        int st;
        if (batch_in_chunk) { // whole arrays: no carry between chunks
            st = b200_scan_exclusive_async((uint32_t *)d_out, (const uint32_t *)d_in, m / L, L, nullptr,
                                           (b200_stream_t)p.compute);
        }
        else { // slice of one long array: carry in d_carry[k&1], next carry to d_carry[(k+1)&1]
            st = b200_scan_exclusive_async((uint32_t *)d_out, (const uint32_t *)d_in, 1, m, d_carry + (k & 1),
                                           (b200_stream_t)p.compute);
            if (!st) {
                next_carry_kernel<<<1, 1, 0, p.compute>>>(d_carry + ((k + 1) & 1), d_carry + (k & 1),
                                                          (const unsigned *)d_out + (m - 1),
                                                          (const unsigned *)d_in + (m - 1));
                count_launch();
            }
        }
        if (st)
            return st;
        B200_TRY(cudaGetLastError());
        B200_TRY(cudaEventRecord(p.freed[b], p.compute));
        B200_TRY(cudaStreamWaitEvent(p.d2h, p.freed[b], 0));
        B200_TRY(cudaMemcpyAsync(h_dst + off, d_out, m * 4, cudaMemcpyDeviceToHost, p.d2h));
        B200_TRY(cudaEventRecord(p.done[b], p.d2h));
    }
    return 0;
}

} // namespace b200

using namespace b200;

extern "C" int b200_reduce_i32_host(const int32_t *h_in, size_t n, int32_t *h_result)
{
    return reduce_host<int32_t>(0, h_in, n, h_result);
}
extern "C" int b200_reduce_f32_host(const float *h_in, size_t n, float *h_result)
{
    return reduce_host<float>(1, h_in, n, h_result);
}
extern "C" int b200_reduce_f64_host(const double *h_in, size_t n, double *h_result)
{
    return reduce_host<double>(2, h_in, n, h_result);
}
// experiments (not part of the stable ABI): mode 0 staged / 1 zero-copy; chunk_bytes > 0 sets the staging chunk
extern "C" int b200_tune_host_path(int mode, size_t chunk_bytes)
{
    g_host_mode = mode;
    if (chunk_bytes)
        g_chunk_bytes = chunk_bytes;
    return 0;
}
extern "C" int b200_histogram256_host(uint32_t *h_hist, const void *h_data, size_t byteCount)
{
    return hist_host(h_hist, 256, h_data, byteCount, b200_histogram256_async);
}
extern "C" int b200_histogram64_host(uint32_t *h_hist, const void *h_data, size_t byteCount)
{
    return hist_host(h_hist, 64, h_data, byteCount, b200_histogram64_async);
}

extern "C" int b200_scan_exclusive_host(uint32_t *h_dst, const uint32_t *h_src, size_t batch, size_t L)
{
    if (batch == 0 || L == 0)
        return 0;
    Pipe p;
    B200_TRY((cudaError_t)open_pipe(p, 2 * kChunkBytes));
    B200_TRY((cudaError_t)ensure(p.c->host_work, (64u << 10) + 4096, p.compute));
    unsigned *d_carry = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(p.c->host_work.ptr) + (64u << 10) + 64);
    const size_t per_chunk = kChunkBytes / 4;
    int          st        = 0;
    if (L <= per_chunk) {
        st = scan_host_run(p, h_dst, h_src, batch * L, per_chunk / L, L, d_carry);
    }
    else {
        for (size_t a = 0; a < batch && !st; a++) {
            B200_TRY(cudaMemsetAsync(d_carry, 0, 2 * sizeof(unsigned), p.compute));
            st = scan_host_run(p, h_dst + a * L, h_src + a * L, L, 0, L, d_carry);
            // the event slots are reused by the next array: drain first
            if (!st && a + 1 < batch) {
                B200_TRY(cudaStreamSynchronize(p.d2h));
                B200_TRY(cudaStreamSynchronize(p.compute));
            }
        }
    }
    if (st)
        return st;
    B200_TRY(cudaStreamSynchronize(p.d2h));
    B200_TRY(cudaStreamSynchronize(p.compute));
    return 0;
}

extern "C" int b200_sort_kv_host(uint32_t *h_dstKey, uint32_t *h_dstVal, const uint32_t *h_srcKey,
                                 const uint32_t *h_srcVal, size_t N, uint32_t sortDir)
{
    if (N == 0)
        return 0;
    Pipe p;
    B200_TRY((cudaError_t)open_pipe(p, 4096));
    uint32_t *d = nullptr; // srcK srcV bufK bufV dstK dstV
    B200_TRY(cudaMalloc(&d, 6 * N * sizeof(uint32_t)));
    uint32_t *sK = d, *sV = d + N, *bK = d + 2 * N, *bV = d + 3 * N, *dK = d + 4 * N, *dV = d + 5 * N;
    // keys on one copy stream, values on the other; the sort waits for both
    B200_TRY(cudaMemcpyAsync(sK, h_srcKey, N * 4, cudaMemcpyHostToDevice, p.h2d));
    B200_TRY(cudaEventRecord(p.ready[0], p.h2d));
    B200_TRY(cudaMemcpyAsync(sV, h_srcVal, N * 4, cudaMemcpyHostToDevice, p.h2d));
    B200_TRY(cudaEventRecord(p.ready[1], p.h2d));
    B200_TRY(cudaStreamWaitEvent(p.compute, p.ready[0], 0));
    B200_TRY(cudaStreamWaitEvent(p.compute, p.ready[1], 0));
    int st = b200_sort_kv_async(dK, dV, bK, bV, sK, sV, N, sortDir, (b200_stream_t)p.compute);
    if (!st) {
        cudaError_t e = cudaMemcpyAsync(h_dstKey, dK, N * 4, cudaMemcpyDeviceToHost, p.compute);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(h_dstVal, dV, N * 4, cudaMemcpyDeviceToHost, p.compute);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(p.compute);
        st = (int)e;
    }
    cudaFree(d);
    return st;
}
