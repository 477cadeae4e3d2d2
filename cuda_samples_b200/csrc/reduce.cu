// reduce.cu -- single-pass sum reduction for sm_100a.
//
// Replaces reduce<T>() and its ten teaching kernels (Samples/2_Concepts_and_Techniques/reduction/
// reduction_kernel.cu:104-617, launcher :633-982) plus the caller's multi-pass finish
// (reduction.cpp:292-318) with ONE launch:
//   * persistent grid (SMs x kCtasPerSm CTAs), grid-stride over 16-byte vectors, kUnroll
//     independent ld.global.nc.L1::no_allocate.v4 in flight per thread before any arithmetic;
//   * int: wrapping u32 adds, warp finish with redux.sync (__reduce_add_sync);
//     float: every batch of 4*kUnroll values is summed as a depth-log2 fp32 tree and then added to
//     a per-thread fp64 accumulator, so the rounding error does not grow with n;
//     double: per-thread Kahan (compensated) fp64 accumulators;
//   * per-CTA partial -> global scratch -> last CTA (atomic ticket + __threadfence) adds the
//     partials in a fixed order: deterministic run to run, no floating-point atomics;
//   * d_out[0] = total, d_out[1..out_len) = 0 so every caller path of the sample still adds up.
// Roofline: HBM-bound, sizeof(T) algorithmic bytes per element (read once, nothing written).
#include "common.cuh"
#include "internal.h"
#include "p2p.cuh"

#ifndef B200_REDUCE_KEEP_L2
#define B200_REDUCE_KEEP_L2 0
#endif
namespace b200 {

constexpr int kReduceThreads = 512;
constexpr int kReduceCtasPerSm = 3;
constexpr int kReduceUnroll = 4;

// ---- per-thread accumulators ---------------------------------------------------------------------
struct LaneI32 {
    typedef unsigned Part;
    typedef int      Elem;
    unsigned s = 0;
    __device__ __forceinline__ void add_vecs(const uint4 (&v)[kReduceUnroll])
    {
#pragma unroll
        for (int k = 0; k < kReduceUnroll; k++)
            s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
    }
    __device__ __forceinline__ void add_vec(uint4 v) { s += (v.x + v.y) + (v.z + v.w); }
    __device__ __forceinline__ void add_scalar(int x) { s += (unsigned)x; }
    __device__ __forceinline__ Part partial() const { return s; }
};

struct LaneF32 {
    typedef double Part;
    typedef float  Elem;
    double s = 0.0;
    static __device__ __forceinline__ float sum4(uint4 v)
    {
        return (__uint_as_float(v.x) + __uint_as_float(v.y)) + (__uint_as_float(v.z) + __uint_as_float(v.w));
    }
    __device__ __forceinline__ void add_vecs(const uint4 (&v)[kReduceUnroll])
    {
        float t[kReduceUnroll];
#pragma unroll
        for (int k = 0; k < kReduceUnroll; k++)
            t[k] = sum4(v[k]);
#pragma unroll
        for (int w = kReduceUnroll / 2; w >= 1; w >>= 1)
#pragma unroll
            for (int k = 0; k < w; k++)
                t[k] = t[k] + t[k + w];
        s += (double)t[0];
    }
    __device__ __forceinline__ void add_vec(uint4 v) { s += (double)sum4(v); }
    __device__ __forceinline__ void add_scalar(float x) { s += (double)x; }
    __device__ __forceinline__ Part partial() const { return s; }
};

struct LaneF64 {
    typedef double Part;
    typedef double Elem;
    // two Kahan accumulators (one per double of a 16-byte vector)
    double s0 = 0.0, c0 = 0.0, s1 = 0.0, c1 = 0.0;
    static __device__ __forceinline__ void kahan(double &s, double &c, double x)
    {
        double y = __dsub_rn(x, c);
        double t = __dadd_rn(s, y);
        c        = __dsub_rn(__dsub_rn(t, s), y);
        s        = t;
    }
    __device__ __forceinline__ void add_vec(uint4 v)
    {
        kahan(s0, c0, __hiloint2double((int)v.y, (int)v.x));
        kahan(s1, c1, __hiloint2double((int)v.w, (int)v.z));
    }
    __device__ __forceinline__ void add_vecs(const uint4 (&v)[kReduceUnroll])
    {
#pragma unroll
        for (int k = 0; k < kReduceUnroll; k++)
            add_vec(v[k]);
    }
    __device__ __forceinline__ void add_scalar(double x) { kahan(s0, c0, x); }
    __device__ __forceinline__ Part partial() const { return (s0 - c0) + (s1 - c1); }
};

// ---- CTA-wide sum, fixed order (deterministic) ---------------------------------------------------
__device__ __forceinline__ unsigned warp_sum(unsigned v) { return __reduce_add_sync(0xffffffffu, v); }
__device__ __forceinline__ double   warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <typename Part> __device__ __forceinline__ Part block_sum(Part v, Part *smem /* >= 32 */)
{
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0)
        smem[warp] = v;
    __syncthreads();
    Part r = Part(0);
    if (warp == 0) {
        r = (lane < (blockDim.x >> 5)) ? smem[lane] : Part(0);
        r = warp_sum(r);
    }
    __syncthreads();
    return r; // valid in warp 0
}

// What the last CTA does after the local sum when the kernel also carries the node-wide exchange (kP2P):
// d_slots[0..count) are this rank's shard partials (doubles; this launch fills [index], earlier launches on the
// stream the others); they are stored into every peer's window, every rank's slots are gathered, and
// sums[k] = sum over ranks (rank order, fp64) of slot k.  carry (optional) = sum of slot [index] over the LOWER
// ranks as wrapping u32: the carry-in of this rank's shard in a sharded scan.
struct ReduceExchange {
    P2PArgs   p2p;
    double   *slots;
    double   *gathered; // [world][count]
    double   *sums;     // [count]
    unsigned *carry;    // nullable
    unsigned  index, count;
};

// scratch layout: [0] ticket (u32), [8] result slot of an in-place call, [16 ..) one 8-byte partial per CTA
template <typename L, bool kP2P>
__global__ void __launch_bounds__(kReduceThreads, kReduceCtasPerSm)
    reduce_kernel(const typename L::Elem *__restrict__ in, size_t n, typename L::Elem *__restrict__ out,
                  size_t out_len, unsigned *__restrict__ ticket, typename L::Part *__restrict__ partials,
                  double *__restrict__ out_f64, const typename L::Part *__restrict__ chain_in,
                  typename L::Part *__restrict__ chain_out, ReduceExchange ex, bool reverse)
{
    typedef typename L::Elem Elem;
    typedef typename L::Part Part;
    constexpr int kPerVec = 16 / sizeof(Elem);
    __shared__ Part s_part[32];
    __shared__ bool s_last;
    // A kernel behind this one that opted in to programmatic dependent launch (the chained scan of a sharded scan,
    // which needs this kernel's result -- the carry -- only when its first tile publishes) may start filling its
    // pipeline while the last CTA here is still exchanging the shard totals with the peers.
    pdl_launch_dependents();

    L acc;
    // split [0,n) into scalar head (to 16-byte alignment), vector body, scalar tail
    const uintptr_t addr = (uintptr_t)in;
    size_t head = ((16 - (addr & 15)) & 15) / sizeof(Elem);
    if (head > n)
        head = n;
    const size_t nvec = (n - head) / kPerVec;
    const size_t tail = head + nvec * kPerVec;
    const uint4 *vin  = reinterpret_cast<const uint4 *>(in + head);

    const size_t stride = (size_t)gridDim.x * kReduceThreads;
    size_t       v      = (size_t)blockIdx.x * kReduceThreads + threadIdx.x;
    // reverse: walk the array from its end to its start, so that its HEAD is what the L2 holds when the kernel ends
    // (the shard total of a sharded scan: the scan that follows starts at the head and finds it cached)
    auto at = [&](size_t idx) { return vin + (reverse ? nvec - 1 - idx : idx); };
#if B200_REDUCE_KEEP_L2
    const uint64_t pol = reverse ? kL2EvictLast : 0x1000000000000000ull; // (evict_normal)
#define ld_stream_u4(p) ld_stream_hint_u4(p, pol)
#endif
    // main loop: kUnroll independent 16-byte loads in flight per thread
    for (; v + (kReduceUnroll - 1) * stride < nvec; v += kReduceUnroll * stride) {
        uint4 r[kReduceUnroll];
#pragma unroll
        for (int k = 0; k < kReduceUnroll; k++)
            r[k] = ld_stream_u4(at(v + k * stride));
        acc.add_vecs(r);
    }
    if (v < nvec) { // last partial round: independent guarded loads (no one-at-a-time latency chain)
        uint4 r[kReduceUnroll - 1];
#pragma unroll
        for (int k = 0; k < kReduceUnroll - 1; k++)
            r[k] = (v + k * stride < nvec) ? ld_stream_u4(at(v + k * stride)) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int k = 0; k < kReduceUnroll - 1; k++)
            acc.add_vec(r[k]); // zeros are exact no-ops for every T
    }
#if B200_REDUCE_KEEP_L2
#undef ld_stream_u4
#endif
    if (blockIdx.x == 0) {
        for (size_t i = threadIdx.x; i < head; i += kReduceThreads)
            acc.add_scalar(in[i]);
        for (size_t i = tail + threadIdx.x; i < n; i += kReduceThreads)
            acc.add_scalar(in[i]);
    }

    Part p = block_sum<Part>(acc.partial(), s_part);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = p;
        __threadfence();
        unsigned t = atomicAdd(ticket, 1u);
        s_last     = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last)
        return;

    // last CTA: fixed-order sum of the per-CTA partials
    __threadfence();
    Part q = Part(0);
    for (unsigned i = threadIdx.x; i < gridDim.x; i += kReduceThreads)
        q += __ldcg(partials + i);
    q = block_sum<Part>(q, s_part);
    if (threadIdx.x == 0) {
        // chunked input (host_path.cu): the running total of the earlier chunks comes in un-rounded and goes out
        // un-rounded, so a float array streamed in chunks is still rounded to fp32 exactly once
        if (chain_in)
            q += *chain_in;
        if (chain_out)
            *chain_out = q;
        if (out_len > 0)
            out[0] = (Elem)q;
        if (out_f64) // shard partial for the multi-GPU combine: the typed result, widened exactly
            out_f64[0] = (double)(Elem)q;
        *ticket = 0; // ready for the next call (stream-ordered)
    }
    for (size_t i = 1 + threadIdx.x; i < out_len; i += kReduceThreads)
        out[i] = Elem(0);
    if (kP2P) { // compute -> collective in the same kernel: P2P stores over NVLink, no NCCL call
        __shared__ unsigned long long s_seq;
        __syncthreads(); // slot [index] (written by thread 0 just above) is visible to the CTA
        const unsigned           words = 2 * ex.count;
        const unsigned seq =
            p2p_exchange(ex.p2p, reinterpret_cast<const unsigned *>(ex.slots), words, &s_seq);
        auto peer = [&](unsigned r, unsigned k) {
            return __hiloint2double((int)p2p_peer_word(ex.p2p, seq, r, 2 * k + 1),
                                    (int)p2p_peer_word(ex.p2p, seq, r, 2 * k));
        };
        for (unsigned i = threadIdx.x; i < ex.p2p.world * ex.count; i += kReduceThreads)
            ex.gathered[i] = peer(i / ex.count, i % ex.count);
        if (threadIdx.x < ex.count) {
            double t = 0.0;
            for (unsigned r = 0; r < ex.p2p.world; r++)
                t += peer(r, threadIdx.x);
            ex.sums[threadIdx.x] = t;
        }
        if (ex.carry && threadIdx.x == 0) {
            unsigned cin = 0;
            for (unsigned r = 0; r < ex.p2p.rank; r++)
                cin += (unsigned)(long long)peer(r, ex.index);
            *ex.carry = cin;
        }
    }
}

template <typename L>
static int launch_reduce(const typename L::Elem *d_in, size_t n, typename L::Elem *d_out, size_t out_len,
                         cudaStream_t stream, double *d_out_f64 = nullptr, const ReduceExchange *ex = nullptr,
                         Scratch *scratch = nullptr, const void *chain_in = nullptr, void *chain_out = nullptr,
                         bool reverse = false)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const int max_grid = c->num_sms * kReduceCtasPerSm;
    Scratch  &work     = scratch ? *scratch : c->reduce;
    B200_TRY((cudaError_t)ensure(work, 16 + sizeof(double) * (size_t)max_grid, stream));
    constexpr int kPerVec = 16 / sizeof(typename L::Elem);
    size_t        nvec    = n / kPerVec + 1;
    size_t        want    = (nvec + kReduceThreads - 1) / kReduceThreads;
    int           grid    = (int)(want < (size_t)max_grid ? want : (size_t)max_grid);
    if (grid < 1)
        grid = 1;
    typedef typename L::Part Part;
    typedef typename L::Elem Elem;
    // In place (threadFenceReduction.cu:247 calls reduce(s, threads, blocks, d_odata, d_odata)): the kernel reads its
    // input through the read-only path, so an output that overlaps the input goes through a scratch slot and is
    // copied into place by stream-ordered copies behind the kernel.
    Elem *alias_out = nullptr;
    if (d_out && out_len && (uintptr_t)d_out < (uintptr_t)(d_in + n) && (uintptr_t)(d_out + out_len) > (uintptr_t)d_in) {
        alias_out = d_out;
        d_out     = reinterpret_cast<Elem *>(reinterpret_cast<char *>(work.ptr) + 8);
    }
    const size_t kernel_out_len = alias_out ? 1 : out_len;
    unsigned *ticket   = reinterpret_cast<unsigned *>(work.ptr);
    auto     *partials = reinterpret_cast<Part *>(reinterpret_cast<char *>(work.ptr) + 16);
    if (ex)
        reduce_kernel<L, true><<<grid, kReduceThreads, 0, stream>>>(d_in, n, d_out, kernel_out_len, ticket, partials,
                                                                   d_out_f64, (const Part *)chain_in,
                                                                   (Part *)chain_out, *ex, reverse);
    else
        reduce_kernel<L, false><<<grid, kReduceThreads, 0, stream>>>(d_in, n, d_out, kernel_out_len, ticket, partials,
                                                                    d_out_f64, (const Part *)chain_in,
                                                                    (Part *)chain_out, ReduceExchange(), reverse);
    count_launch();
    B200_TRY(cudaGetLastError());
    // what this kernel writes while a dependent it has already released may be running (internal.h)
    pdl_note_output(c->device, d_out, kernel_out_len * sizeof(Elem));
    pdl_note_output(c->device, d_out_f64, sizeof(double));
    pdl_note_output(c->device, chain_out, sizeof(Part));
    if (ex) {
        pdl_note_output(c->device, ex->gathered, sizeof(double) * ex->count * ex->p2p.world);
        pdl_note_output(c->device, ex->sums, sizeof(double) * ex->count);
        pdl_note_output(c->device, ex->carry, sizeof(unsigned));
    }
    if (alias_out) {
        B200_TRY(cudaMemcpyAsync(alias_out, d_out, sizeof(Elem), cudaMemcpyDeviceToDevice, stream));
        if (out_len > 1)
            B200_TRY(cudaMemsetAsync(alias_out + 1, 0, (out_len - 1) * sizeof(Elem), stream));
    }
    return 0;
}

int reduce_preload()
{
    cudaFuncAttributes fa;
    B200_TRY(cudaFuncGetAttributes(&fa, reduce_kernel<LaneI32, true>));
    B200_TRY(cudaFuncGetAttributes(&fa, reduce_kernel<LaneF32, true>));
    B200_TRY(cudaFuncGetAttributes(&fa, reduce_kernel<LaneF64, true>));
    B200_TRY(cudaFuncGetAttributes(&fa, reduce_kernel<LaneI32, false>));
    B200_TRY(cudaFuncGetAttributes(&fa, reduce_kernel<LaneF32, false>));
    B200_TRY(cudaFuncGetAttributes(&fa, reduce_kernel<LaneF64, false>));
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    return ensure(c->reduce, 16 + sizeof(double) * (size_t)c->num_sms * kReduceCtasPerSm, 0);
}

// One chunk of a host array streamed through the device (host_path.cu): own scratch (so the caller's *_async work on
// other streams is not disturbed), running total chained through 8-byte slots, typed result written to d_out when
// given (the last chunk; d_out may be mapped pinned host memory).  kind: 0 i32, 1 f32, 2 f64.
int reduce_chunk(int kind, const void *d_in, size_t n, void *d_out, Scratch *scratch, const void *chain_in,
                 void *chain_out, cudaStream_t stream)
{
    const size_t out_len = d_out ? 1 : 0;
    switch (kind) {
    case 0:
        return launch_reduce<LaneI32>((const int *)d_in, n, (int *)d_out, out_len, stream, nullptr, nullptr, scratch,
                                      chain_in, chain_out);
    case 1:
        return launch_reduce<LaneF32>((const float *)d_in, n, (float *)d_out, out_len, stream, nullptr, nullptr,
                                      scratch, chain_in, chain_out);
    case 2:
        return launch_reduce<LaneF64>((const double *)d_in, n, (double *)d_out, out_len, stream, nullptr, nullptr,
                                      scratch, chain_in, chain_out);
    }
    return (int)cudaErrorInvalidValue;
}

} // namespace b200

using namespace b200;

extern "C" int b200_reduce_i32_async(const int32_t *d_in, size_t n, int32_t *d_out, size_t out_len,
                                     b200_stream_t stream)
{
    return launch_reduce<LaneI32>(d_in, n, d_out, out_len, (cudaStream_t)stream);
}
extern "C" int b200_reduce_f32_async(const float *d_in, size_t n, float *d_out, size_t out_len, b200_stream_t stream)
{
    return launch_reduce<LaneF32>(d_in, n, d_out, out_len, (cudaStream_t)stream);
}
extern "C" int b200_reduce_f64_async(const double *d_in, size_t n, double *d_out, size_t out_len,
                                     b200_stream_t stream)
{
    return launch_reduce<LaneF64>(d_in, n, d_out, out_len, (cudaStream_t)stream);
}

// Shard partial of a multi-GPU reduction: the sum of d_in[0..n) in its own type (wrapping for int), written as a
// double to *d_slot so that partials of different types share one all_gather buffer.  kind: 0 i32, 1 f32, 2 f64.
extern "C" int b200_reduce_partial_async(int kind, const void *d_in, size_t n, double *d_slot, b200_stream_t stream)
{
    switch (kind) {
    case 0: return launch_reduce<LaneI32>((const int *)d_in, n, nullptr, 0, (cudaStream_t)stream, d_slot);
    case 1: return launch_reduce<LaneF32>((const float *)d_in, n, nullptr, 0, (cudaStream_t)stream, d_slot);
    case 2: return launch_reduce<LaneF64>((const double *)d_in, n, nullptr, 0, (cudaStream_t)stream, d_slot);
    }
    return (int)cudaErrorInvalidValue;
}

// The same, fused with the node-wide exchange over the NVLink peer windows (p2p.cuh; b200_p2p_connect first).
// d_slots[0..count): this rank's shard partials, slot [index] is produced by this launch.  On return (stream
// order) d_gathered[r * count + k] = rank r's slot k, d_sums[k] = their sum in rank order (fp64), and, when
// d_carry is not null, *d_carry = wrapping u32 sum of slot [index] over the ranks below this one.
extern "C" int b200_reduce_partial_exchange_async(int kind, const void *d_in, size_t n, double *d_slots,
                                                  unsigned index, unsigned count, double *d_gathered, double *d_sums,
                                                  uint32_t *d_carry, b200_stream_t stream)
{
    const P2PArgs *a = p2p_args();
    if (!a || a->world == 0 || count == 0 || index >= count || 8u * count > kP2PSlotBytes ||
        count > (unsigned)kReduceThreads)
        return (int)cudaErrorInvalidValue;
    ReduceExchange ex;
    ex.p2p      = *a;
    ex.slots    = d_slots;
    ex.gathered = d_gathered;
    ex.sums     = d_sums;
    ex.carry    = d_carry;
    ex.index    = index;
    ex.count    = count;
    double *slot = d_slots + index;
    switch (kind) {
    case 0: return launch_reduce<LaneI32>((const int *)d_in, n, nullptr, 0, (cudaStream_t)stream, slot, &ex);
    case 1: return launch_reduce<LaneF32>((const float *)d_in, n, nullptr, 0, (cudaStream_t)stream, slot, &ex);
    case 2: return launch_reduce<LaneF64>((const double *)d_in, n, nullptr, 0, (cudaStream_t)stream, slot, &ex);
    }
    return (int)cudaErrorInvalidValue;
}

// Sharded long scan, step 1 fused with its exchange: shard total -> every peer -> *d_carry = wrapping sum of the
// lower ranks' totals (the d_carry argument of b200_scan_exclusive_async).  d_work: >= 2 + world doubles.
static bool g_scan_total_reverse = true;
extern "C" int b200_tune_scan_total_reverse(int on)
{
    g_scan_total_reverse = on != 0;
    return 0;
}
// ---------------------------------------------------------------------------------------------------------------
// Sharded long scan as reduce-then-scan over 8192-element tiles (16-byte aligned shards of n % 4 == 0 elements).
// A rank's carry needs the lower ranks' totals, and a total needs the whole shard: the shard is read twice whatever
// the scan is.  The chained scan spends its second pass waiting on look-backs (5.4 TB/s); here the FIRST pass leaves
// one sum per tile behind, its last CTA exchanges the shard total with the peers (NVLink windows, no collective call)
// and turns the tile sums into exclusive tile prefixes -- carry included -- in place, and the second pass
// (scan.cu: scan_tiled) is a scan with no dependency between tiles at all: the kernel of the sample's "short"
// arrays, 6.9 TB/s.  The first pass walks the shard back to front, so the second one finds the shard's head in L2.
// ---------------------------------------------------------------------------------------------------------------
constexpr int      kTileSumThreads = 256;
constexpr unsigned kTileSumTile    = 8192; // == the tile of scan_kernel (scan.cu: kScanTile)

__global__ void __launch_bounds__(kTileSumThreads)
    scan_tile_sums_kernel(const unsigned *__restrict__ src, size_t n, unsigned *__restrict__ tsum, unsigned ntiles,
                          unsigned *__restrict__ ticket, ReduceExchange ex, bool exchange)
{
    __shared__ unsigned           s_part[32];
    __shared__ unsigned           s_tot[kP2PMaxRanks]; // every rank's shard total of this exchange
    __shared__ unsigned           s_carry;
    __shared__ bool               s_last;
    __shared__ unsigned long long s_seq;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_launch_dependents(); // the scan behind this kernel may load and scan its first tiles under our tail
    {
        const unsigned tile  = ntiles - 1 - blockIdx.x;
        const size_t   base  = (size_t)tile * kTileSumTile;
        const size_t   rem   = n - base;
        const unsigned count = rem < kTileSumTile ? (unsigned)rem : kTileSumTile; // a multiple of 4
        const uint4   *v     = reinterpret_cast<const uint4 *>(src + base);
        uint4          r[kTileSumTile / 4 / kTileSumThreads];
#pragma unroll
        for (int k = 0; k < (int)(kTileSumTile / 4 / kTileSumThreads); k++) {
            const unsigned idx = k * kTileSumThreads + tid;
            r[k]               = (4 * idx < count) ? ld_stream_u4(v + idx) : make_uint4(0, 0, 0, 0);
        }
        unsigned s = 0;
#pragma unroll
        for (int k = 0; k < (int)(kTileSumTile / 4 / kTileSumThreads); k++)
            s += (r[k].x + r[k].y) + (r[k].z + r[k].w);
        s = __reduce_add_sync(0xffffffffu, s);
        if (lane == 0)
            s_part[warp] = s;
        __syncthreads();
        if (tid == 0) {
            unsigned t = 0;
#pragma unroll
            for (int w = 0; w < kTileSumThreads / 32; w++)
                t += s_part[w];
            tsum[tile] = t;
            __threadfence();
            s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        }
        __syncthreads();
        if (!s_last)
            return;
    }
    // ---- last CTA: shard total -> peers -> carry; tile sums -> exclusive tile prefixes (in place) ----------------
    __threadfence();
    const unsigned per = (ntiles + kTileSumThreads - 1) / kTileSumThreads; // a contiguous run of tiles per thread
    const unsigned b0 = tid * per < ntiles ? tid * per : ntiles, b1 = b0 + per < ntiles ? b0 + per : ntiles;
    unsigned       local = 0;
    for (unsigned b = b0; b < b1; b++)
        local += __ldcg(tsum + b);
    unsigned inc = local; // inclusive scan of the threads' run totals: lanes, then warps
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d)
            inc += t;
    }
    if (lane == 31)
        s_part[warp] = inc;
    __syncthreads();
    unsigned before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kTileSumThreads / 32; w++) {
        before += (w < (int)warp) ? s_part[w] : 0u;
        total += s_part[w];
    }
    unsigned run = before + inc - local;
    unsigned carry = 0;
    if (exchange) {
        if (tid == 0)
            ex.slots[ex.index] = (double)total;
        __syncthreads();
        const unsigned seq = p2p_exchange(ex.p2p, reinterpret_cast<const unsigned *>(ex.slots), 2 * ex.count, &s_seq);
        // EVERY rank waits for EVERY rank's total, although only the lower ones enter its carry: that wait is what
        // keeps the ranks in step.  The windows' slot-reuse rule (p2p.cuh) holds because no rank can start exchange
        // s + 2 before all ranks have published s + 1; a rank that read nobody's words (rank 0 needs none) would run
        // ahead through back-to-back calls and overwrite words its peers have not read yet.
        if (tid < ex.p2p.world)
            s_tot[tid] = (unsigned)(long long)__hiloint2double((int)p2p_peer_word(ex.p2p, seq, tid, 2 * ex.index + 1),
                                                               (int)p2p_peer_word(ex.p2p, seq, tid, 2 * ex.index));
        __syncthreads();
        if (tid == 0) {
            unsigned cin = 0;
            for (unsigned r = 0; r < ex.p2p.rank; r++)
                cin += s_tot[r];
            s_carry = cin;
            if (ex.carry)
                *ex.carry = cin;
        }
        __syncthreads();
        carry = s_carry;
    }
    run += carry;
    for (unsigned b = b0; b < b1; b++) {
        const unsigned x = __ldcg(tsum + b);
        tsum[b]          = run;
        run += x;
    }
    if (tid == 0)
        *ticket = 0; // ready for the next call (stream-ordered)
}

namespace b200 {
int scan_shard_preload()
{
    cudaFuncAttributes fa;
    B200_TRY(cudaFuncGetAttributes(&fa, scan_tile_sums_kernel));
    return 0;
}
} // namespace b200

static bool g_scan_shard_tiled = true;
extern "C" int b200_tune_scan_shard_tiled(int on)
{
    g_scan_shard_tiled = on != 0;
    return 0;
}

// One shard of a sharded exclusive scan, both steps: d_dst = exclusive scan of d_src seeded with the wrapping sum of
// the LOWER ranks' shards.  Connected peer windows (b200_p2p_connect / b200_ctx) carry the exchange; with a world of
// one it is a plain scan.  d_work: >= 2 + world doubles of device memory.  Shards that are not 16-byte aligned or
// whose length is not a multiple of 4 take the chained scan behind b200_scan_total_exchange_async.
extern "C" int b200_scan_exclusive_shard_async(uint32_t *d_dst, const uint32_t *d_src, size_t n, double *d_work,
                                               uint32_t *d_carry, b200_stream_t stream_)
{
    cudaStream_t   stream = (cudaStream_t)stream_;
    const P2PArgs *a      = p2p_args();
    if (!a || a->world == 0 || !d_work || !d_carry)
        return (int)cudaErrorInvalidValue;
    const bool tiled = g_scan_shard_tiled && n >= 4 * (size_t)kTileSumTile && (n & 3) == 0 &&
                       ((((uintptr_t)d_dst) | ((uintptr_t)d_src)) & 15) == 0 && n / kTileSumTile < 0x7fffffffull;
    if (!tiled) {
        B200_TRY((cudaError_t)b200_scan_total_exchange_async(d_carry, d_src, n, d_work, stream_));
        return n ? b200_scan_exclusive_async(d_dst, d_src, 1, n, d_carry, stream_) : 0;
    }
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const unsigned ntiles = (unsigned)((n + kTileSumTile - 1) / kTileSumTile);
    B200_TRY((cudaError_t)ensure(c->scan_tiles, 16 + 4 * (size_t)ntiles, stream));
    unsigned *ticket = reinterpret_cast<unsigned *>(c->scan_tiles.ptr);
    unsigned *tsum   = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(c->scan_tiles.ptr) + 16);
    ReduceExchange ex;
    ex.p2p      = *a;
    ex.slots    = d_work;
    ex.gathered = d_work + 2;
    ex.sums     = d_work + 1;
    ex.carry    = d_carry;
    ex.index    = 0;
    ex.count    = 1;
    scan_tile_sums_kernel<<<ntiles, kTileSumThreads, 0, stream>>>(d_src, n, tsum, ntiles, ticket, ex, true);
    B200_TRY(cudaGetLastError());
    count_launch();
    pdl_note_output(c->device, d_work, sizeof(double) * (2 + a->world));
    pdl_note_output(c->device, d_carry, sizeof(uint32_t));
    return scan_tiled(d_dst, d_src, n, tsum, stream);
}

extern "C" int b200_scan_total_exchange_async(uint32_t *d_carry, const uint32_t *d_src, size_t n, double *d_work,
                                              b200_stream_t stream)
{
    const P2PArgs *a = p2p_args();
    if (!a || a->world == 0)
        return (int)cudaErrorInvalidValue;
    ReduceExchange ex;
    ex.p2p      = *a;
    ex.slots    = d_work;
    ex.gathered = d_work + 2;
    ex.sums     = d_work + 1;
    ex.carry    = d_carry;
    ex.index    = 0;
    ex.count    = 1;
    // the shard is read back to front: its head is still in the L2 when the chained scan starts there
    return launch_reduce<LaneI32>((const int *)d_src, n, nullptr, 0, (cudaStream_t)stream, d_work, &ex, nullptr, nullptr,
                                  nullptr, g_scan_total_reverse);
}
