// sorting_networks.cu -- batched short-array key-value sorts on the radix-sort kernels.
//
// Replaces bitonicSort() / oddEvenMergeSort() of Samples/2_Concepts_and_Techniques/sortingNetworks
// (sortingNetworks_common.h:45-55; bitonicSort.cu, oddEvenMergeSort.cu): batchSize independent arrays of
// arrayLength (key, value) u32 pairs, dir != 0 ascending / == 0 descending.  The samples' networks need
// power-of-two lengths and are not stable; this is a stable segmented sort for any length:
//   1. stable sort of (key, index) over the whole batch            -> perm  (order of all pairs by key)
//   2. stable sort of (perm / arrayLength, perm) by the segment id  -> perm2 (LSD: key first, then segment)
//      -- one b200_partition_kv_async pass when batchSize <= 256, a full sort otherwise
//   3. gather keys and values through perm2.
// Same kernels as mergeSort (sort_hist4_kernel, sort_pass_kernel); three tiny kernels around them.
//
// Arrays of up to 4096 pairs -- the lengths the sorting networks are made for (bitonicSort.cu:42-96 sorts an array of
// up to SHARED_SIZE_LIMIT = 1024 in one CTA's shared memory) -- never leave the SM: seg_sort_kernel loads a tile of
// whole arrays (up to 4096 pairs) into shared memory, sorts it there with stable 8-bit radix passes and writes it
// once: 16 B of traffic per pair instead of the >= 150 B of the composition above.
//   * passes: one per key byte that actually VARIES inside the tile (OR/AND of the tile's keys; the sample's
//     keys are < 65536: two passes), then -- when the tile holds several arrays -- one pass on the array index,
//     which puts every array back in its place (LSD order: key bytes first, array index last);
//   * a pass is the tile-local part of sort_pass_kernel: per-warp digit counts (one shared RED per pair), the
//     digit owners scan them into scatter positions, eight-ballot match ranking, the pair goes to its slot in
//     place (every thread holds its pairs in registers across the barrier);
//   * stable, any array length, both directions (descending = ascending on ~key).
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int      kSegThreads = 512;
constexpr int      kSegWarps   = kSegThreads / 32;
constexpr int      kSegItems   = 8;
constexpr unsigned kSegTile    = kSegThreads * kSegItems; // 4096 pairs
constexpr unsigned kSegMaxArrays = 255;                    // per tile: the array index is an 8-bit digit, 255 = padding

struct SegSmem {
    uint2         pair[kSegTile];
    unsigned char seg[kSegTile];
    unsigned      wcnt[kSegWarps][256];
    unsigned      wsum[8];
    unsigned      bits[2 * kSegWarps]; // OR / AND of the keys per warp
};

__device__ __forceinline__ unsigned seg_match_digit(unsigned d)
{
    unsigned mask = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool     p = (d >> b) & 1u;
        const unsigned v = __ballot_sync(0xffffffffu, p);
        mask &= p ? v : ~v;
    }
    return mask;
}

__global__ void __launch_bounds__(kSegThreads, 2)
    seg_sort_kernel(unsigned *__restrict__ dstK, unsigned *__restrict__ dstV, const unsigned *__restrict__ srcK,
                    const unsigned *__restrict__ srcV, size_t batch, unsigned L, unsigned apc /* arrays per CTA */,
                    unsigned magic /* ceil(2^32 / L) */, unsigned flip)
{
    extern __shared__ __align__(16) unsigned char seg_raw[];
    SegSmem       &S    = *reinterpret_cast<SegSmem *>(seg_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = lanemask_lt();
    const size_t   a0 = (size_t)blockIdx.x * apc;
    const unsigned na = (unsigned)(batch - a0 < apc ? batch - a0 : apc);
    const unsigned nv = na * L; // valid pairs of this tile
    const size_t   g0 = a0 * L;

    // ---- load (coalesced), find the key bits that vary inside the tile ------------------------------------------
    unsigned key[kSegItems], val[kSegItems], sid[kSegItems];
    unsigned all_or = 0, all_and = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < kSegItems; i++) {
        const unsigned e = warp * (kSegItems * 32) + i * 32 + lane;
        if (e < nv) {
            key[i] = __ldg(srcK + g0 + e);
            val[i] = __ldg(srcV + g0 + e);
            sid[i] = __umulhi(e, magic); // e / L, exact for e, L <= 4096
            all_or |= key[i] ^ flip;
            all_and &= key[i] ^ flip;
        }
        else {
            key[i] = ~flip; // ordered value 0xffffffff: behind every real key of its digit in every pass
            val[i] = 0;
            sid[i] = 255;
        }
    }
    all_or  = __reduce_or_sync(0xffffffffu, all_or);
    all_and = __reduce_and_sync(0xffffffffu, all_and);
    if (lane == 0) {
        S.bits[warp]             = all_or;
        S.bits[kSegWarps + warp] = all_and;
    }
    __syncthreads();
    all_or = 0, all_and = 0xffffffffu;
#pragma unroll
    for (int w = 0; w < kSegWarps; w++) {
        all_or |= S.bits[w];
        all_and &= S.bits[kSegWarps + w];
    }
    const unsigned varying = nv ? (all_or ^ all_and) : 0u;
    unsigned      *mycnt   = S.wcnt[warp];

    // ---- passes: key bytes that vary (least significant first), then the array index -----------------------------
    for (int pass = 0; pass < 5; pass++) {
        if (pass < 4 ? ((varying >> (8 * pass)) & 255u) == 0 : apc == 1)
            continue; // (tile-uniform)
        if (pass > 0 || true) {
#pragma unroll
            for (int k = 0; k < 8; k++)
                mycnt[k * 32 + lane] = 0;
            __syncwarp();
        }
        auto digit = [&](int i) { return pass < 4 ? (((key[i] ^ flip) >> (8 * pass)) & 255u) : sid[i]; };
#pragma unroll
        for (int i = 0; i < kSegItems; i++)
            atomicAdd(mycnt + digit(i), 1u);
        __syncthreads();
        unsigned count = 0, start = 0;
        if (tid < 256) {
#pragma unroll
            for (int w = 0; w < kSegWarps; w++)
                count += S.wcnt[w][tid];
            unsigned inc = count;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d)
                    inc += t;
            }
            if (lane == 31)
                S.wsum[warp] = inc;
            start = inc - count;
        }
        __syncthreads();
        if (tid < 256) {
#pragma unroll
            for (int w = 0; w < 8; w++)
                start += (w < (int)warp) ? S.wsum[w] : 0u;
#pragma unroll
            for (int w = 0; w < kSegWarps; w++) {
                const unsigned cw = S.wcnt[w][tid];
                S.wcnt[w][tid]    = start;
                start += cw;
            }
        }
        __syncthreads();
        // rank in index order (warp, item, lane) and scatter in place: every pair of the tile sits in registers
#pragma unroll
        for (int i = 0; i < kSegItems; i++) {
            const unsigned d     = digit(i);
            const unsigned mask  = seg_match_digit(d);
            const unsigned below = mask & lt;
            unsigned       pre   = 0;
            if (below == 0)
                pre = atomicAdd(mycnt + d, (unsigned)__popc(mask));
            pre                 = __shfl_sync(0xffffffffu, pre, __ffs(mask) - 1);
            const unsigned slot = pre + __popc(below);
            S.pair[slot]        = make_uint2(key[i], val[i]);
            S.seg[slot]         = (unsigned char)sid[i];
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kSegItems; i++) {
            const unsigned e  = warp * (kSegItems * 32) + i * 32 + lane;
            const uint2    pr = S.pair[e];
            key[i]            = pr.x;
            val[i]            = pr.y;
            sid[i]            = S.seg[e];
        }
        __syncthreads(); // (the counter rows are re-zeroed and the slots rewritten by the next pass)
    }
    // ---- store: position e of the sorted tile is pair e of the tile's arrays -------------------------------------
#pragma unroll
    for (int i = 0; i < kSegItems; i++) {
        const unsigned e = warp * (kSegItems * 32) + i * 32 + lane;
        if (e < nv) {
            dstK[g0 + e] = key[i];
            dstV[g0 + e] = val[i];
        }
    }
}

__global__ void iota_kernel(unsigned *out, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = (unsigned)i;
}
__global__ void segment_of_kernel(unsigned *seg, const unsigned *perm, size_t n, unsigned L)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        seg[i] = perm[i] / L;
}
__global__ void iota_splitters_kernel(unsigned *spl, unsigned n)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        spl[i] = i + 1; // bucket(seg) = #splitters <= seg = seg
}
__global__ void gather_kv_kernel(unsigned *dk, unsigned *dv, const unsigned *sk, const unsigned *sv,
                                 const unsigned *perm, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const unsigned p = perm[i];
        dk[i]            = sk[p];
        dv[i]            = sv[p];
    }
}

} // namespace b200

using namespace b200;

extern "C" int b200_segmented_sort_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, const uint32_t *d_srcKey,
                                            const uint32_t *d_srcVal, size_t batchSize, size_t arrayLength,
                                            uint32_t sortDir, b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t N      = batchSize * arrayLength;
    if (N == 0)
        return 0;
    if (N >= (1ull << 30) || arrayLength >= (1ull << 32))
        return (int)cudaErrorInvalidValue;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    if (arrayLength == 1) { // every array is already sorted (and ceil(2^32 / 1) does not fit the kernel's 32-bit divisor)
        if (d_dstKey != d_srcKey)
            B200_TRY(cudaMemcpyAsync(d_dstKey, d_srcKey, N * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream));
        if (d_dstVal != d_srcVal)
            B200_TRY(cudaMemcpyAsync(d_dstVal, d_srcVal, N * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream));
        return 0;
    }
    if (arrayLength <= kSegTile) { // short arrays are sorted inside one CTA's shared memory: 16 B of traffic per pair
        unsigned apc = (unsigned)(kSegTile / arrayLength);
        apc          = apc > kSegMaxArrays ? kSegMaxArrays : apc;
        const size_t tiles = (batchSize + apc - 1) / apc;
        if (tiles < (1ull << 31)) {
            B200_TRY((cudaError_t)opt_in_smem(c, kAttrSegSort, seg_sort_kernel, sizeof(SegSmem)));
            const unsigned magic = (unsigned)(((1ull << 32) + arrayLength - 1) / arrayLength);
            seg_sort_kernel<<<(unsigned)tiles, kSegThreads, sizeof(SegSmem), stream>>>(
                d_dstKey, d_dstVal, d_srcKey, d_srcVal, batchSize, (unsigned)arrayLength, apc, magic,
                sortDir ? 0u : 0xffffffffu);
            count_launch();
            B200_TRY(cudaGetLastError());
            return 0;
        }
    }
    // scratch: idx | permK | perm | bufK | bufV | seg | perm2 | (splitters, counts)
    B200_TRY((cudaError_t)ensure(c->stage[1], (7 * N + 1024) * sizeof(unsigned), stream));
    unsigned *w = reinterpret_cast<unsigned *>(c->stage[1].ptr);
    unsigned *idx = w, *permK = w + N, *perm = w + 2 * N, *bufK = w + 3 * N, *bufV = w + 4 * N, *seg = w + 5 * N,
             *perm2 = w + 6 * N, *spl = w + 7 * N, *counts = w + 7 * N + 512;
    const int grid = (int)((N + 255) / 256 < (size_t)c->num_sms * 8 ? (N + 255) / 256 : (size_t)c->num_sms * 8);
    if (batchSize == 1) { // one array: this IS mergeSort
        return b200_sort_kv_async(d_dstKey, d_dstVal, bufK, bufV, d_srcKey, d_srcVal, N, sortDir, stream_);
    }
    iota_kernel<<<grid, 256, 0, stream>>>(idx, N);
    B200_TRY(cudaGetLastError());
    B200_TRY((cudaError_t)b200_sort_kv_async(permK, perm, bufK, bufV, d_srcKey, idx, N, sortDir, stream_));
    segment_of_kernel<<<grid, 256, 0, stream>>>(seg, perm, N, (unsigned)arrayLength);
    B200_TRY(cudaGetLastError());
    if (batchSize <= 256) {
        iota_splitters_kernel<<<1, 256, 0, stream>>>(spl, (unsigned)batchSize - 1);
        B200_TRY(cudaGetLastError());
        B200_TRY((cudaError_t)b200_partition_kv_async(bufK, perm2, seg, perm, N, spl, (uint32_t)batchSize, 1, counts,
                                                      stream_));
    }
    else {
        B200_TRY((cudaError_t)b200_sort_kv_async(permK, perm2, bufK, bufV, seg, perm, N, 1, stream_));
    }
    gather_kv_kernel<<<grid, 256, 0, stream>>>(d_dstKey, d_dstVal, d_srcKey, d_srcVal, perm2, N);
    B200_TRY(cudaGetLastError());
    count_launch(3);
    return 0;
}

// ---- the sample's entry points (legacy default stream, abort on error) ------------------------------------------
extern "C" unsigned bitonicSort(unsigned *d_DstKey, unsigned *d_DstVal, unsigned *d_SrcKey, unsigned *d_SrcVal,
                                unsigned batchSize, unsigned arrayLength, unsigned dir)
{
    B200_OR_DIE(b200_segmented_sort_kv_async(d_DstKey, d_DstVal, d_SrcKey, d_SrcVal, batchSize, arrayLength, dir, 0),
                "bitonicSort");
    return 256; // CTA size of the pass kernel (the sample returns its thread count; only printed)
}
extern "C" void oddEvenMergeSort(unsigned *d_DstKey, unsigned *d_DstVal, unsigned *d_SrcKey, unsigned *d_SrcVal,
                                 unsigned batchSize, unsigned arrayLength, unsigned dir)
{
    B200_OR_DIE(b200_segmented_sort_kv_async(d_DstKey, d_DstVal, d_SrcKey, d_SrcVal, batchSize, arrayLength, dir, 0),
                "oddEvenMergeSort");
}
