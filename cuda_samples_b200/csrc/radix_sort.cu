// radix_sort.cu -- stable key-value u32 sort (onesweep LSD radix, 8-bit digits) for sm_100a.
//
// Replaces mergeSort() and its four kernels (Samples/0_Introduction/mergeSort/mergeSort.cu:
// mergeSortSharedKernel :108, generateSampleRanksKernel :182, mergeRanksAndIndicesKernel :237,
// mergeElementaryIntervalsKernel :326, host loop :471-537: 1 + 4*log2(N/1024) launches and
// ~16 B/pair of traffic per merge level, N <= 2^22) with
//   1 memset + 1 digit-histogram kernel (reads the keys once for all four digits and leaves the
//   exclusive digit offsets behind) + 4 scatter passes, each a single chained-scan kernel:
//     * tiles of 8192 pairs are taken in ticket order; keys are loaded warp-striped;
//     * ranking is stable by construction: inside a warp, items are ranked in index order with a
//       ballot-built match mask per 8-bit digit (rank = running warp-private digit counter + number
//       of lower lanes holding the same digit), warps are ordered by a per-digit scan over warps;
//     * the tile's 256 digit counts go to a look-back descriptor row (2 status bits + 30-bit
//       value per digit); thread d walks back over earlier tiles for digit d;
//     * keys, then values, are staged through shared memory in digit order so that each digit's
//       run leaves the SM as one contiguous store.
//   Passes ping-pong src -> buf -> dst -> buf -> dst: src is never written, the result lands in dst.
//   sortDir == 0 (descending) sorts on ~key; ties keep source order in both directions.
// The same pass kernel with a splitter-based "digit" is the local partition step of the multi-GPU
// sample sort (b200_partition_kv_async).
// Traffic model: (4 + 16*4) = 68 B per pair; HBM-bound in bytes, issue-bound in practice.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int      kSortThreads = 512;
constexpr int      kSortWarps   = kSortThreads / 32;
constexpr int      kSortItems   = 16;
constexpr unsigned kSortTile    = kSortThreads * kSortItems; // 8192 pairs
constexpr int      kRadix       = 256;
constexpr unsigned kDescA = 1u << 30, kDescP = 2u << 30, kDescValue = (1u << 30) - 1;

// shared memory of the pass kernel (dynamic, > 48 KiB)
struct SortSmem {
    unsigned buf[kSortTile];              // keys, then values, in digit order
    unsigned whist[kSortWarps][kRadix];   // warp-private digit counters -> per-warp scatter bases
    unsigned gofs[kRadix];                // global index of a digit's first slot minus its tile-local start
    unsigned wsum[kRadix / 32];
    unsigned splitters[kRadix];
    unsigned tile;
};

// "digit" of a key: radix byte of (key ^ flip), or the bucket index among ascending splitters
template <bool kBucket> struct DigitOp {
    unsigned        shift, flip, nspl;
    const unsigned *spl; // shared memory
    __device__ __forceinline__ unsigned operator()(unsigned key) const
    {
        const unsigned o = key ^ flip;
        if (!kBucket)
            return (o >> shift) & 255u;
        // number of splitters <= o  (upper bound), nspl <= 255
        unsigned lo = 0, hi = nspl;
        while (lo < hi) {
            unsigned mid = (lo + hi) >> 1;
            if (spl[mid] <= o)
                lo = mid + 1;
            else
                hi = mid;
        }
        return lo;
    }
};

// lanes of the warp holding the same 8-bit digit (ballot per bit)
__device__ __forceinline__ unsigned match_digit(unsigned d)
{
    unsigned mask = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool     bit  = (d >> b) & 1u;
        const unsigned vote = __ballot_sync(0xffffffffu, bit);
        mask &= bit ? vote : ~vote;
    }
    return mask;
}

// ---------------------------------------------------------------------------------------------------
// Digit histograms of all passes in one read of the keys; the last CTA turns each 256-bin histogram
// into exclusive offsets in place.  ghist: [passes][256], zero on entry.
// ---------------------------------------------------------------------------------------------------
template <bool kBucket>
__global__ void __launch_bounds__(kSortThreads)
    sort_hist_kernel(const unsigned *__restrict__ keys, size_t n, unsigned flip, const unsigned *__restrict__ splitters,
                     unsigned nspl, unsigned *__restrict__ ghist, unsigned *__restrict__ ticket,
                     unsigned *__restrict__ counts_out)
{
    constexpr int       P = kBucket ? 1 : 4;
    __shared__ unsigned h[P][kRadix];
    __shared__ unsigned s_spl[kRadix];
    __shared__ bool     s_last;
    for (int i = threadIdx.x; i < P * kRadix; i += kSortThreads)
        (&h[0][0])[i] = 0;
    if (kBucket && threadIdx.x < nspl)
        s_spl[threadIdx.x] = splitters[threadIdx.x];
    __syncthreads();
    DigitOp<kBucket> op{0, flip, nspl, s_spl};

    auto count = [&](unsigned key) {
        if (kBucket) {
            atomicAdd(&h[0][op(key)], 1u);
        }
        else {
            const unsigned o = key ^ flip;
            atomicAdd(&h[0][o & 255u], 1u);
            atomicAdd(&h[1 % P][(o >> 8) & 255u], 1u);
            atomicAdd(&h[2 % P][(o >> 16) & 255u], 1u);
            atomicAdd(&h[3 % P][o >> 24], 1u);
        }
    };
    const uintptr_t addr = (uintptr_t)keys;
    size_t          head = ((16 - (addr & 15)) & 15) / 4;
    if (head > n)
        head = n;
    const size_t nvec   = (n - head) / 4;
    const size_t tail   = head + nvec * 4;
    const uint4 *vin    = reinterpret_cast<const uint4 *>(keys + head);
    const size_t stride = (size_t)gridDim.x * kSortThreads;
    for (size_t v = (size_t)blockIdx.x * kSortThreads + threadIdx.x; v < nvec; v += stride) {
        uint4 r = ld_stream_u4(vin + v);
        count(r.x);
        count(r.y);
        count(r.z);
        count(r.w);
    }
    if (blockIdx.x == 0) {
        for (size_t i = threadIdx.x; i < head; i += kSortThreads)
            count(keys[i]);
        for (size_t i = tail + threadIdx.x; i < n; i += kSortThreads)
            count(keys[i]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < P * kRadix; i += kSortThreads) {
        unsigned c = (&h[0][0])[i];
        if (c)
            atomicAdd(&ghist[i], c);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        s_last     = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    // warp p scans histogram p: lane l owns entries [8l, 8l+8)
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < P) {
        unsigned *g = ghist + warp * kRadix + lane * 8;
        unsigned  c[8], sum = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            c[i] = __ldcg(g + i);
            sum += c[i];
        }
        if (kBucket && counts_out)
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (lane * 8 + i <= nspl)
                    counts_out[lane * 8 + i] = c[i];
        unsigned inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d)
                inc += t;
        }
        unsigned run = inc - sum;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            g[i] = run;
            run += c[i];
        }
    }
    if (threadIdx.x == 0)
        *ticket = 0;
}

// ---------------------------------------------------------------------------------------------------
// One scatter pass (chained scan over tiles).
// ---------------------------------------------------------------------------------------------------
template <bool kBucket>
__global__ void __launch_bounds__(kSortThreads, 2)
    sort_pass_kernel(const unsigned *__restrict__ kin, const unsigned *__restrict__ vin, unsigned *__restrict__ kout,
                     unsigned *__restrict__ vout, unsigned n, unsigned shift, unsigned flip,
                     const unsigned *__restrict__ splitters, unsigned nspl, const unsigned *__restrict__ gbase,
                     unsigned *__restrict__ ticket, unsigned *__restrict__ desc)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem      &S    = *reinterpret_cast<SortSmem *>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0)
        S.tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < kSortWarps * kRadix; i += kSortThreads)
        (&S.whist[0][0])[i] = 0;
    if (kBucket && tid < nspl)
        S.splitters[tid] = splitters[tid];
    __syncthreads();
    const unsigned   tile  = S.tile;
    const unsigned   base  = tile * kSortTile;
    const unsigned   valid = (n - base < kSortTile) ? (n - base) : kSortTile;
    DigitOp<kBucket> op{shift, flip, nspl, S.splitters};

    // ---- load keys (warp-striped: item i of lane l is element warp*512 + i*32 + l of the tile) --------
    unsigned       key[kSortItems];
    const unsigned wbase = warp * (kSortItems * 32) + lane;
    const unsigned pad   = ~flip; // (pad ^ flip) = 0xffffffff: last digit in every pass, ranks after all valid keys
    if (valid == kSortTile) {
#pragma unroll
        for (int i = 0; i < kSortItems; i++)
            key[i] = __ldg(kin + base + wbase + i * 32);
    }
    else {
#pragma unroll
        for (int i = 0; i < kSortItems; i++)
            key[i] = (wbase + i * 32 < valid) ? __ldg(kin + base + wbase + i * 32) : pad;
    }

    // ---- rank inside the warp, in index order -----------------------------------------------------------
    unsigned rank[kSortItems];
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const unsigned d      = op(key[i]);
        const unsigned mask   = match_digit(d);
        const unsigned leader = __ffs(mask) - 1;
        unsigned       pre    = 0;
        if (lane == leader) {
            pre              = S.whist[warp][d];
            S.whist[warp][d] = pre + __popc(mask);
        }
        pre     = __shfl_sync(0xffffffffu, pre, leader);
        rank[i] = pre + __popc(mask & lanemask_lt());
        __syncwarp();
    }

    // values: issue the loads now, they are consumed after the look-back
    unsigned val[kSortItems];
    if (valid == kSortTile) {
#pragma unroll
        for (int i = 0; i < kSortItems; i++)
            val[i] = __ldg(vin + base + wbase + i * 32);
    }
    else {
#pragma unroll
        for (int i = 0; i < kSortItems; i++)
            val[i] = (wbase + i * 32 < valid) ? __ldg(vin + base + wbase + i * 32) : 0u;
    }
    __syncthreads();

    // ---- per digit: order the warps, publish the tile count, look back --------------------------------
    unsigned count = 0, tile_start = 0;
    if (tid < kRadix) {
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            unsigned t       = S.whist[w][tid];
            S.whist[w][tid]  = count;
            count += t;
        }
        st_relaxed_u32(desc + (size_t)tile * kRadix + tid, (tile == 0 ? kDescP : kDescA) | count);
        // exclusive scan of the 256 counts -> where each digit starts inside the tile
        unsigned inc = count;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d)
                inc += t;
        }
        if (lane == 31)
            S.wsum[warp] = inc;
        tile_start = inc - count;
    }
    __syncthreads();
    if (tid < kRadix) {
        unsigned woff = 0;
#pragma unroll
        for (int w = 0; w < kRadix / 32; w++)
            woff += (w < (int)warp) ? S.wsum[w] : 0u;
        tile_start += woff;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++)
            S.whist[w][tid] += tile_start;
        unsigned excl = 0;
        if (tile > 0) {
            const unsigned *dp = desc + tid;
            for (long long t = (long long)tile - 1;; t--) {
                unsigned d;
                do {
                    d = ld_relaxed_u32(dp + (size_t)t * kRadix);
                } while ((d >> 30) == 0);
                excl += d & kDescValue;
                if (d & kDescP)
                    break;
            }
            st_relaxed_u32(desc + (size_t)tile * kRadix + tid, kDescP | (excl + count));
        }
        S.gofs[tid] = __ldg(gbase + tid) + excl - tile_start;
    }
    __syncthreads();

    // ---- keys: to shared memory in digit order, then out -----------------------------------------------
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const unsigned d = op(key[i]);
        rank[i] += S.whist[warp][d]; // tile-local slot
        S.buf[rank[i]] = key[i];
    }
    __syncthreads();
    unsigned gidx[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const unsigned j = tid + k * kSortThreads;
        const unsigned x = S.buf[j];
        const unsigned d = op(x);
        gidx[k]          = S.gofs[d] + j;
        if (j < valid)
            kout[gidx[k]] = x;
    }
    __syncthreads();
    // ---- values: same slots ------------------------------------------------------------------------------
#pragma unroll
    for (int i = 0; i < kSortItems; i++)
        S.buf[rank[i]] = val[i];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSortItems; k++) {
        const unsigned j = tid + k * kSortThreads;
        if (j < valid)
            vout[gidx[k]] = S.buf[j];
    }
}

// scratch layout (bytes): [0,16) hist ticket | [16, 16+4 KiB) digit offsets [4][256]
//                         | [4112, 4128) pass tickets | descriptors [passes][tiles][256] from 4128
constexpr size_t kSortOffHist = 16, kSortOffTickets = 16 + 4 * kRadix * 4, kSortOffDesc = kSortOffTickets + 16;

template <bool kBucket> static int prepare_pass_kernel()
{
    static bool done = false;
    if (!done) {
        B200_TRY(cudaFuncSetAttribute(sort_pass_kernel<kBucket>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)sizeof(SortSmem)));
        done = true;
    }
    return 0;
}

} // namespace b200

using namespace b200;

extern "C" int b200_sort_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, uint32_t *d_bufKey, uint32_t *d_bufVal,
                                  const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t N, uint32_t sortDir,
                                  b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N == 0)
        return 0;
    if (N >= (1ull << 30))
        return (int)cudaErrorInvalidValue; // 30-bit look-back values; shard larger inputs
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY((cudaError_t)prepare_pass_kernel<false>());
    const unsigned tiles      = (unsigned)((N + kSortTile - 1) / kSortTile);
    const size_t   desc_bytes = (size_t)tiles * kRadix * 4;
    const size_t   need       = kSortOffDesc + 4 * desc_bytes;
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    char *base = reinterpret_cast<char *>(c->sort.ptr);
    B200_TRY(cudaMemsetAsync(base, 0, need, stream));
    unsigned *hticket = reinterpret_cast<unsigned *>(base);
    unsigned *ghist   = reinterpret_cast<unsigned *>(base + kSortOffHist);
    unsigned *tickets = reinterpret_cast<unsigned *>(base + kSortOffTickets);
    const unsigned flip = sortDir ? 0u : 0xffffffffu;

    size_t want = (N / 4 + kSortThreads - 1) / kSortThreads;
    size_t cap  = (size_t)c->num_sms * 4;
    int    grid = (int)(want < cap ? want : cap);
    if (grid < 1)
        grid = 1;
    sort_hist_kernel<false><<<grid, kSortThreads, 0, stream>>>(d_srcKey, N, flip, nullptr, 0, ghist, hticket, nullptr);
    B200_TRY(cudaGetLastError());

    const uint32_t *kin = d_srcKey, *vin = d_srcVal;
    for (int p = 0; p < 4; p++) {
        uint32_t *kout = (p & 1) ? d_dstKey : d_bufKey;
        uint32_t *vout = (p & 1) ? d_dstVal : d_bufVal;
        unsigned *desc = reinterpret_cast<unsigned *>(base + kSortOffDesc + p * desc_bytes);
        sort_pass_kernel<false><<<tiles, kSortThreads, sizeof(SortSmem), stream>>>(
            kin, vin, kout, vout, (unsigned)N, 8u * p, flip, nullptr, 0, ghist + p * kRadix, tickets + p, desc);
        B200_TRY(cudaGetLastError());
        kin = kout;
        vin = vout;
    }
    count_launch(5);
    return 0;
}

extern "C" int b200_partition_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, const uint32_t *d_srcKey,
                                       const uint32_t *d_srcVal, size_t N, const uint32_t *d_splitters,
                                       uint32_t numBuckets, uint32_t sortDir, uint32_t *d_bucketCounts,
                                       b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (numBuckets < 1 || numBuckets > kRadix || N >= (1ull << 30))
        return (int)cudaErrorInvalidValue;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    if (N == 0) {
        B200_TRY(cudaMemsetAsync(d_bucketCounts, 0, numBuckets * 4, stream));
        return 0;
    }
    B200_TRY((cudaError_t)prepare_pass_kernel<true>());
    const unsigned tiles      = (unsigned)((N + kSortTile - 1) / kSortTile);
    const size_t   desc_bytes = (size_t)tiles * kRadix * 4;
    const size_t   need       = kSortOffDesc + desc_bytes;
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    char *base = reinterpret_cast<char *>(c->sort.ptr);
    B200_TRY(cudaMemsetAsync(base, 0, need, stream));
    unsigned *hticket = reinterpret_cast<unsigned *>(base);
    unsigned *ghist   = reinterpret_cast<unsigned *>(base + kSortOffHist);
    unsigned *tickets = reinterpret_cast<unsigned *>(base + kSortOffTickets);
    unsigned *desc    = reinterpret_cast<unsigned *>(base + kSortOffDesc);
    const unsigned flip = sortDir ? 0u : 0xffffffffu;
    const unsigned nspl = numBuckets - 1;

    size_t want = (N / 4 + kSortThreads - 1) / kSortThreads;
    size_t cap  = (size_t)c->num_sms * 4;
    int    grid = (int)(want < cap ? want : cap);
    if (grid < 1)
        grid = 1;
    sort_hist_kernel<true><<<grid, kSortThreads, 0, stream>>>(d_srcKey, N, flip, d_splitters, nspl, ghist, hticket,
                                                            d_bucketCounts);
    B200_TRY(cudaGetLastError());
    sort_pass_kernel<true><<<tiles, kSortThreads, sizeof(SortSmem), stream>>>(
        d_srcKey, d_srcVal, d_dstKey, d_dstVal, (unsigned)N, 0u, flip, d_splitters, nspl, ghist, tickets, desc);
    B200_TRY(cudaGetLastError());
    count_launch(2);
    return 0;
}
