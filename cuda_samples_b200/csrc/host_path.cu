// host_path.cu -- host-buffer entry points (b200_*_host): what a CPU caller of histogram256CPU /
// reduceCPU / scanExclusiveHost / mergeSortHost (host pointers in, host pointers out) switches to.
// The host<->device copies live INSIDE the call: inputs are streamed through two device staging
// chunks so the H2D copy engine, the SMs and the D2H copy engine overlap; bench.py's `e2e` figure
// times exactly these functions.  They block until the result is in host memory.
#include "common.cuh"
#include "internal.h"

#include <type_traits>

namespace b200 {

constexpr size_t kChunkBytes = 32u << 20; // per staging buffer

struct Pipe {
    DeviceCtx   *c       = nullptr;
    cudaStream_t compute = nullptr, h2d = nullptr, d2h = nullptr;
    cudaEvent_t  ready[2], freed[2], done[2];
};

static int open_pipe(Pipe &p, size_t stage_bytes)
{
    int st = 0;
    p.c    = ctx(&st);
    if (!p.c)
        return st;
    DeviceCtx *c = p.c;
    // the library scratch is shared with whatever the caller enqueued through the *_async entry points
    B200_TRY(cudaDeviceSynchronize());
    if (!c->copy_stream[0]) {
        B200_TRY(cudaStreamCreateWithFlags(&c->copy_stream[0], cudaStreamNonBlocking));
        B200_TRY(cudaStreamCreateWithFlags(&c->copy_stream[1], cudaStreamNonBlocking));
        B200_TRY(cudaStreamCreateWithFlags(&c->compute_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 6; i++)
            B200_TRY(cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming));
    }
    p.h2d     = c->copy_stream[0];
    p.d2h     = c->copy_stream[1];
    p.compute = c->compute_stream;
    for (int i = 0; i < 2; i++) {
        p.ready[i] = c->ev[i];
        p.freed[i] = c->ev[2 + i];
        p.done[i]  = c->ev[4 + i];
        B200_TRY((cudaError_t)ensure(c->stage[i], stage_bytes, p.compute));
    }
    B200_TRY(cudaStreamSynchronize(p.compute)); // scratch memsets done before the copy streams touch it
    return 0;
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
__global__ void sum_rows_f32_kernel(const float *rows, size_t nrows, float *out)
{
    // chunk totals of a float reduction are added up in double
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0.0;
        for (size_t i = 0; i < nrows; i++)
            s += (double)rows[i];
        out[0] = (float)s;
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

template <typename T, typename Launch> static int reduce_host(const T *h_in, size_t n, T *h_result, Launch reduce_async)
{
    Pipe p;
    B200_TRY((cudaError_t)open_pipe(p, kChunkBytes));
    const size_t nchunks = (n * sizeof(T) + kChunkBytes - 1) / kChunkBytes;
    T           *d_part  = nullptr;
    B200_TRY(cudaMallocAsync(&d_part, (nchunks + 2) * sizeof(T), p.compute));
    B200_TRY(cudaMemsetAsync(d_part, 0, (nchunks + 2) * sizeof(T), p.compute));
    int st = stream_input(p, h_in, n * sizeof(T), kChunkBytes, [&](size_t k, void *d, size_t bytes) {
        return reduce_async(reinterpret_cast<const T *>(d), bytes / sizeof(T), d_part + k, 1, (b200_stream_t)p.compute);
    });
    if (st)
        return st;
    T *d_total = d_part + nchunks;
    if (std::is_same<T, float>::value)
        sum_rows_f32_kernel<<<1, 32, 0, p.compute>>>((const float *)d_part, nchunks, (float *)d_total);
    else
        sum_rows_kernel<T><<<1, 32, 0, p.compute>>>(d_part, nchunks, 1, d_total);
    count_launch();
    B200_TRY(cudaGetLastError());
    B200_TRY(cudaMemcpyAsync(h_result, d_total, sizeof(T), cudaMemcpyDeviceToHost, p.compute));
    B200_TRY(cudaFreeAsync(d_part, p.compute));
    B200_TRY(cudaStreamSynchronize(p.compute));
    return 0;
}

template <typename Launch> static int hist_host(uint32_t *h_hist, int bins, const void *h_data, size_t n, Launch hist_async)
{
    Pipe p;
    B200_TRY((cudaError_t)open_pipe(p, kChunkBytes));
    size_t    nchunks = (n + kChunkBytes - 1) / kChunkBytes;
    unsigned *d_part  = nullptr;
    B200_TRY(cudaMallocAsync(&d_part, (nchunks + 1) * bins * sizeof(unsigned), p.compute));
    B200_TRY(cudaMemsetAsync(d_part, 0, (nchunks + 1) * bins * sizeof(unsigned), p.compute));
    int st = stream_input(p, h_data, n, kChunkBytes, [&](size_t k, void *d, size_t bytes) {
        return hist_async(d_part + k * bins, d, bytes, (b200_stream_t)p.compute);
    });
    if (st)
        return st;
    unsigned *d_total = d_part + nchunks * bins;
    sum_rows_kernel<unsigned><<<1, 256, 0, p.compute>>>(d_part, nchunks, bins, d_total);
    count_launch();
    B200_TRY(cudaGetLastError());
    B200_TRY(cudaMemcpyAsync(h_hist, d_total, bins * sizeof(unsigned), cudaMemcpyDeviceToHost, p.compute));
    B200_TRY(cudaFreeAsync(d_part, p.compute));
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
    return reduce_host<int32_t>(h_in, n, h_result, b200_reduce_i32_async);
}
extern "C" int b200_reduce_f32_host(const float *h_in, size_t n, float *h_result)
{
    return reduce_host<float>(h_in, n, h_result, b200_reduce_f32_async);
}
extern "C" int b200_reduce_f64_host(const double *h_in, size_t n, double *h_result)
{
    return reduce_host<double>(h_in, n, h_result, b200_reduce_f64_async);
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
    unsigned *d_carry = nullptr;
    B200_TRY(cudaMallocAsync(&d_carry, 2 * sizeof(unsigned), p.compute));
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
    B200_TRY(cudaFreeAsync(d_carry, p.compute));
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
