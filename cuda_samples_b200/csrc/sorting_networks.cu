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
#include "common.cuh"
#include "internal.h"

namespace b200 {

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
