// histogram.cu -- byte histograms (256 bins, 64 bins, or both in one pass) for sm_100a.
//
// Replaces histogram256Kernel + mergeHistogram256Kernel (Samples/2_Concepts_and_Techniques/
// histogram/histogram256.cu:54-126) and histogram64Kernel + mergeHistogram64Kernel
// (histogram64.cu:67-158) with ONE launch and no partial-histogram buffer:
//   * one persistent CTA of 1024 threads per SM, grid-stride 16-byte streaming loads, kUnroll in
//     flight per thread;
//   * counting: a shared-memory table laid out [bin][lane] (u32).  Lane l of every warp only ever
//     touches column l, so bank == lane and every shared atomic of a warp is bank-conflict-free
//     whatever the data is (uniform random, all-equal, ...) -- the reference's warp-private
//     256-word tables serialise on same-bank bins (histogram256.cu:44-52);
//   * merge: each warp folds 8 rows with redux.sync; non-leader CTAs of a thread-block cluster add
//     their bins into the leader's shared memory through DSMEM (red.shared::cluster), the leader
//     issues one red.global per bin into a zeroed accumulator; the last leader (atomic ticket)
//     copies the accumulator to d_Histogram (overwrite semantics of the sample) and re-zeroes it.
//   * histogram64 bin = byte >> 2 (the TOP six bits, histogram_gold.cpp:41-44); the fused mode
//     derives it exactly from the 256-bin totals (hist64[v] = sum hist256[4v..4v+3]).
// Roofline: HBM-bound by bytes (1 B per input byte) but paced by shared-atomic throughput
// (one ATOMS lane-op per input byte); see DESIGN.md.
#include <cooperative_groups.h>

#include "common.cuh"
#include "internal.h"
#include "p2p.cuh"

namespace cg = cooperative_groups;

namespace b200 {

#ifndef B200_HIST_THREADS
#define B200_HIST_THREADS 1024
#endif
constexpr int kHistThreadsBig = B200_HIST_THREADS;       // 1 CTA of 1024 threads per SM
constexpr int kHistThreadsSmall = 512;                   // short inputs: two calls' CTAs fit on an SM together
#ifndef B200_HIST_UNROLL
#define B200_HIST_UNROLL 8
#endif
#ifndef B200_HIST_EARLY
#define B200_HIST_EARLY 0
#endif
constexpr int kHistUnroll  = B200_HIST_UNROLL;

// kHistFusedP2P: the fused mode whose last CTA also all-reduces the 256 bins over the NVLink peer windows (p2p.cuh)
enum HistMode { kHist256 = 0, kHist64 = 1, kHistFused = 2, kHistFusedP2P = 3 };

// scratch layout (u32 words): [0] ticket, [4..260) acc256 (the peer-exchange mode's u32 accumulator), [260..324)
// unused, from word 328: 256 + 64 u64 accumulators {contributions so far << 40 | sum} of the single-GPU modes
constexpr int kHistAcc256 = 4;
constexpr int kHistAcc64  = 260;
constexpr int kHistAccU64 = 328;
constexpr int kHistWords  = kHistAccU64 + 2 * (256 + 64);
constexpr unsigned long long kHistOne = 1ull << 40; // a cluster's bin count is < 2^32 and there are < 2^8 clusters

template <int BINS, int SH> __device__ __forceinline__ void count_word(unsigned *table, unsigned lane4, unsigned w)
{
    // byte k of w -> bin = (byte >> SH); table byte offset = bin * 128 + lane * 4
    constexpr unsigned kMask = (unsigned)(BINS - 1) << 7;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int s   = k * 8 + SH - 7;
        unsigned  off = (s >= 0) ? (w >> (s >= 0 ? s : 0)) : (w << (s < 0 ? -s : 0));
        off           = (off & kMask) | lane4;
        atomicAdd(reinterpret_cast<unsigned *>(reinterpret_cast<char *>(table) + off), 1u);
    }
}

template <int BINS, int SH> __device__ __forceinline__ void count_vec(unsigned *table, unsigned lane4, uint4 v)
{
    count_word<BINS, SH>(table, lane4, v.x);
    count_word<BINS, SH>(table, lane4, v.y);
    count_word<BINS, SH>(table, lane4, v.z);
    count_word<BINS, SH>(table, lane4, v.w);
}

// Programmatic dependent launch (the sample's driver calls histogram256() 17 times back to back,
// histogram/main.cpp:161-170): every launch carries cudaLaunchAttributeProgrammaticStreamSerialization, the kernel
// releases its dependents at its very top (griddepcontrol.launch_dependents) and waits for its predecessor
// (griddepcontrol.wait) only before it touches what consecutive calls share -- the {count,sum} accumulators and the
// output bins.  Call k+1 therefore zeroes its table and streams its input while call k folds, merges and drains; with
// the 512-thread geometry used for short inputs the two calls' CTAs are co-resident on every SM.

template <int MODE, int kHistThreads>
__global__ void __launch_bounds__(kHistThreads, 1024 / kHistThreads)
    histogram_kernel(const unsigned char *__restrict__ data, size_t n, unsigned *__restrict__ out256,
                     unsigned *__restrict__ out64, unsigned *__restrict__ scratch, P2PArgs p2p)
{
    pdl_launch_dependents();
    constexpr int BINS = (MODE == kHist64) ? 64 : 256;
    constexpr int SH   = (MODE == kHist64) ? 2 : 0;
    __shared__ unsigned table[BINS * 32];
    __shared__ unsigned merged[BINS]; // the cluster leader's copy collects the whole cluster
    __shared__ bool     s_last;

    cg::cluster_group cluster = cg::this_cluster();
    const unsigned    crank   = cluster.block_rank();
    const unsigned    csize   = cluster.num_blocks();
    const unsigned    lane    = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned    lane4   = lane * 4;

    const uintptr_t addr = (uintptr_t)data;
    size_t          head = (16 - (addr & 15)) & 15;
    if (head > n)
        head = n;
    const size_t nvec   = (n - head) / 16;
    const size_t tail   = head + nvec * 16;
    const uint4 *vin    = reinterpret_cast<const uint4 *>(data + head);
    const size_t stride = (size_t)gridDim.x * kHistThreads;
    size_t       v      = (size_t)blockIdx.x * kHistThreads + threadIdx.x;
#if B200_HIST_EARLY
    // the first round of loads is in flight while the table is zeroed (short inputs: the fill latency counts)
    uint4      r0[kHistUnroll];
    const bool early = v + (kHistUnroll - 1) * stride < nvec;
    if (early) {
#pragma unroll
        for (int k = 0; k < kHistUnroll; k++)
            r0[k] = ld_stream_u4(vin + v + k * stride);
    }
#endif
    for (int i = threadIdx.x; i < BINS * 32; i += kHistThreads)
        table[i] = 0;
    if (threadIdx.x < BINS)
        merged[threadIdx.x] = 0;
    __syncthreads();

    // ---- count ----------------------------------------------------------------------------------
#if B200_HIST_EARLY
    if (early) {
#pragma unroll
        for (int k = 0; k < kHistUnroll; k++)
            count_vec<BINS, SH>(table, lane4, r0[k]);
        v += kHistUnroll * stride;
    }
#endif
    for (; v + (kHistUnroll - 1) * stride < nvec; v += kHistUnroll * stride) {
        uint4 r[kHistUnroll];
#pragma unroll
        for (int k = 0; k < kHistUnroll; k++)
            r[k] = ld_stream_u4(vin + v + k * stride);
#pragma unroll
        for (int k = 0; k < kHistUnroll; k++)
            count_vec<BINS, SH>(table, lane4, r[k]);
    }
    if (v < nvec) { // last partial round: keep the loads independent (no one-at-a-time latency chain)
        uint4 r[kHistUnroll - 1];
#pragma unroll
        for (int k = 0; k < kHistUnroll - 1; k++)
            if (v + k * stride < nvec)
                r[k] = ld_stream_u4(vin + v + k * stride);
#pragma unroll
        for (int k = 0; k < kHistUnroll - 1; k++)
            if (v + k * stride < nvec)
                count_vec<BINS, SH>(table, lane4, r[k]);
    }
    if (blockIdx.x == 0) { // ragged head / tail bytes
        for (size_t i = threadIdx.x; i < head; i += kHistThreads)
            atomicAdd(&table[(unsigned)(data[i] >> SH) * 32 + lane], 1u);
        for (size_t i = tail + threadIdx.x; i < n; i += kHistThreads)
            atomicAdd(&table[(unsigned)(data[i] >> SH) * 32 + lane], 1u);
    }
    __syncthreads();

    // ---- fold the 32 lane columns: warp w owns rows [w*BINS/32, (w+1)*BINS/32) -------------------
    constexpr int kRowsPerWarp = BINS / (kHistThreads / 32);
    unsigned      mine         = 0; // lane r (< kRowsPerWarp) keeps the total of row warp*kRowsPerWarp + r
#pragma unroll
    for (int r = 0; r < kRowsPerWarp; r++) {
        unsigned t = __reduce_add_sync(0xffffffffu, table[(warp * kRowsPerWarp + r) * 32 + lane]);
        if (lane == r)
            mine = t;
    }
    // ---- cluster merge through DSMEM ---------------------------------------------------------------
    if (csize > 1)
        cluster.sync(); // every CTA's `merged` is zeroed before anyone adds remotely
    unsigned *leader = (csize > 1) ? cluster.map_shared_rank(merged, 0) : merged;
    if (lane < kRowsPerWarp)
        atomicAdd(&leader[warp * kRowsPerWarp + lane], mine);
    if (csize > 1)
        cluster.sync();
    else
        __syncthreads();
    if (crank != 0)
        return;
    pdl_wait(); // the previous call on the stream has completed: accumulators re-armed, its output written

    // ---- leader, single-GPU modes: ONE 64-bit atomic per bin carries the count and the number of clusters that
    // have contributed; the thread that completes a bin writes it out and re-arms the accumulator.  No fence, no
    // ticket, no second pass over the accumulator: the finish is a single L2 round trip.
    if (MODE != kHistFusedP2P) {
        unsigned long long *acc64 = reinterpret_cast<unsigned long long *>(scratch + kHistAccU64);
        const unsigned      nclus = gridDim.x / csize;
        if (threadIdx.x < BINS) {
            const unsigned      mine = merged[threadIdx.x];
            unsigned long long *a    = acc64 + ((MODE == kHist64) ? 256 : 0) + threadIdx.x;
            const unsigned long long old = atomicAdd(a, kHistOne | mine);
            if ((unsigned)(old >> 40) == nclus - 1) {
                *a = 0;
                (MODE == kHist64 ? out64 : out256)[threadIdx.x] = (unsigned)old + mine;
            }
            if (MODE == kHistFused) { // 4 adjacent 256-bins make one 64-bin (exact: histogram_gold.cpp:41-44)
                unsigned s4 = mine + __shfl_xor_sync(0xffffffffu, mine, 1);
                s4 += __shfl_xor_sync(0xffffffffu, s4, 2);
                if ((threadIdx.x & 3) == 0) {
                    unsigned long long      *b    = acc64 + 256 + (threadIdx.x >> 2);
                    const unsigned long long old4 = atomicAdd(b, kHistOne | s4);
                    if ((unsigned)(old4 >> 40) == nclus - 1) {
                        *b                       = 0;
                        out64[threadIdx.x >> 2] = (unsigned)old4 + s4;
                    }
                }
            }
        }
        return;
    }

    // ---- leader, peer-exchange mode: one global reduction per bin, the last leader (ticket) runs the exchange ------
    unsigned *ticket = scratch;
    unsigned *acc    = scratch + ((MODE == kHist64) ? kHistAcc64 : kHistAcc256);
    if (threadIdx.x < BINS)
        atomicAdd(&acc[threadIdx.x], merged[threadIdx.x]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        s_last     = (t == gridDim.x / csize - 1);
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    if (MODE == kHistFusedP2P) { // histogram -> all-reduce in the same kernel: bins to every peer, sum of all ranks back
        __shared__ unsigned long long s_seq;
        const unsigned               seq = p2p_exchange(p2p, acc, BINS, &s_seq);
        __syncthreads(); // every thread has read its share of acc before it is re-zeroed below
        if (threadIdx.x < BINS) {
            unsigned total = 0;
            for (unsigned r = 0; r < p2p.world; r++)
                total += p2p_peer_word(p2p, seq, r, threadIdx.x); // u32 wrap-add: order-independent, bit-exact
            acc[threadIdx.x]    = 0;
            out256[threadIdx.x] = total;
            unsigned s4         = total + __shfl_xor_sync(0xffffffffu, total, 1);
            s4 += __shfl_xor_sync(0xffffffffu, s4, 2);
            if ((threadIdx.x & 3) == 0)
                out64[threadIdx.x >> 2] = s4;
        }
    }
    else if (threadIdx.x < BINS) {
        unsigned total   = __ldcg(&acc[threadIdx.x]);
        acc[threadIdx.x] = 0;
        if (MODE == kHist64) {
            out64[threadIdx.x] = total;
        }
        else {
            out256[threadIdx.x] = total;
            if (MODE == kHistFused) {
                // 4 adjacent 256-bins make one 64-bin
                unsigned s4 = total + __shfl_xor_sync(0xffffffffu, total, 1);
                s4 += __shfl_xor_sync(0xffffffffu, s4, 2);
                if ((threadIdx.x & 3) == 0)
                    out64[threadIdx.x >> 2] = s4;
            }
        }
    }
    if (threadIdx.x == 0)
        *ticket = 0;
}

static int g_hist_cluster = 0; // CTAs per cluster (1, 2, 4 or 8; 2 packs 148 SMs exactly); 0 = by input size
static int g_hist_pdl     = 1; // programmatic dependent launch between consecutive calls on a stream
static int g_hist_small   = 1; // inputs under 128 MiB: 0 = 1024-thread CTAs, 1 / 2 = that many 512-thread CTAs per SM
                               // (with 1, the CTAs of two consecutive calls are co-resident on every SM)

template <int MODE, int THREADS>
static int launch_hist_t(DeviceCtx *c, unsigned *out256, unsigned *out64, const void *d_data, size_t n,
                         cudaStream_t stream, const P2PArgs *p2p, int cluster)
{
    // persistent: one CTA per SM, but never more CTAs than there are 16 KiB slabs of input
    size_t       want  = (n / 16 + THREADS - 1) / THREADS;
    const size_t slots = (size_t)c->num_sms * (THREADS == kHistThreadsSmall ? g_hist_small : 1);
    int          grid  = (int)(want < slots ? want : slots);
    grid               = (grid / cluster) * cluster;
    if (grid < cluster)
        grid = cluster;

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(grid);
    cfg.blockDim           = dim3(THREADS);
    cfg.dynamicSmemBytes   = 0;
    cfg.stream             = stream;
    cudaLaunchAttribute at[2];
    at[0].id               = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    at[1].id               = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs              = at;
    // the kernel streams `data` before its griddepcontrol.wait: only with the attribute when no early-releasing kernel
    // in front can still be writing that range (internal.h: programmatic-launch guard)
    cfg.numAttrs           = (g_hist_pdl && pdl_early_read_ok(c->device, d_data, n)) ? 2 : 1;
    B200_TRY(cudaLaunchKernelEx(&cfg, histogram_kernel<MODE, THREADS>, (const unsigned char *)d_data, n, out256, out64,
                                (unsigned *)c->hist.ptr, p2p ? *p2p : P2PArgs()));
    count_launch();
    pdl_note_output(c->device, out256, 256 * sizeof(unsigned));
    pdl_note_output(c->device, out64, 64 * sizeof(unsigned));
    return 0;
}

template <int MODE>
static int launch_hist(unsigned *out256, unsigned *out64, const void *d_data, size_t n, cudaStream_t stream,
                       const P2PArgs *p2p = nullptr)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY((cudaError_t)ensure(c->hist, kHistWords * sizeof(unsigned), stream));
    // short inputs skip the two cluster barriers of the DSMEM merge (64 MiB: 16.9 vs 18.2 us); from 128 MiB on the
    // merge is free and halves the global atomics
    const int cluster = g_hist_cluster ? g_hist_cluster : (n < (128u << 20) ? 1 : 2);
    if (g_hist_small && cluster == 1 && n < (128u << 20))
        return launch_hist_t<MODE, kHistThreadsSmall>(c, out256, out64, d_data, n, stream, p2p, cluster);
    return launch_hist_t<MODE, kHistThreadsBig>(c, out256, out64, d_data, n, stream, p2p, cluster);
}

int hist_preload()
{
    cudaFuncAttributes fa;
    B200_TRY(cudaFuncGetAttributes(&fa, histogram_kernel<kHistFusedP2P, kHistThreadsBig>));
    B200_TRY(cudaFuncGetAttributes(&fa, histogram_kernel<kHistFusedP2P, kHistThreadsSmall>));
    B200_TRY(cudaFuncGetAttributes(&fa, histogram_kernel<kHistFused, kHistThreadsBig>));
    B200_TRY(cudaFuncGetAttributes(&fa, histogram_kernel<kHistFused, kHistThreadsSmall>));
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    return ensure(c->hist, kHistWords * sizeof(unsigned), 0);
}

int set_hist_cluster(int cluster)
{
    if (cluster != 0 && cluster != 1 && cluster != 2 && cluster != 4 && cluster != 8)
        return (int)cudaErrorInvalidValue;
    g_hist_cluster = cluster;
    return 0;
}

} // namespace b200

using namespace b200;

extern "C" int b200_histogram256_async(uint32_t *d_hist256, const void *d_data, size_t byteCount,
                                       b200_stream_t stream)
{
    return launch_hist<kHist256>(d_hist256, nullptr, d_data, byteCount, (cudaStream_t)stream);
}
extern "C" int b200_histogram64_async(uint32_t *d_hist64, const void *d_data, size_t byteCount, b200_stream_t stream)
{
    return launch_hist<kHist64>(nullptr, d_hist64, d_data, byteCount, (cudaStream_t)stream);
}
extern "C" int b200_histogram_fused_async(uint32_t *d_hist256, uint32_t *d_hist64, const void *d_data,
                                          size_t byteCount, b200_stream_t stream)
{
    return launch_hist<kHistFused>(d_hist256, d_hist64, d_data, byteCount, (cudaStream_t)stream);
}
// tuning knob used by bench.py / tests (not part of the stable ABI)
extern "C" int b200_tune_hist_cluster(int cluster) { return set_hist_cluster(cluster); }
// pdl: 1 = consecutive calls on a stream overlap through programmatic dependent launch (default), 0 = plain launches;
// small_ctas: inputs under 128 MiB run with 0 = 1024-thread CTAs, 1 (default) / 2 = one / two 512-thread CTAs per SM
extern "C" int b200_tune_hist_launch(int pdl, int small_ctas)
{
    g_hist_pdl   = pdl != 0;
    g_hist_small = small_ctas < 0 ? 0 : (small_ctas > 2 ? 2 : small_ctas);
    return 0;
}
// Both histograms of the concatenation of ALL ranks' shards, on every rank, in one kernel per rank: the last CTA
// stores its 256 bins into every peer's window over NVLink and sums the G payloads (b200_p2p_connect first).
extern "C" int b200_histogram_fused_allreduce_async(uint32_t *d_hist256, uint32_t *d_hist64, const void *d_data,
                                                    size_t byteCount, b200_stream_t stream)
{
    const P2PArgs *a = p2p_args();
    if (!a || a->world == 0)
        return (int)cudaErrorInvalidValue;
    return launch_hist<kHistFusedP2P>(d_hist256, d_hist64, d_data, byteCount, (cudaStream_t)stream, a);
}
