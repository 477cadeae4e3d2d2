// dropin.cu -- the samples' own host entry points, same names and signatures, on top of the sm_100a
// kernels.  A caller written against the samples' headers links against libb200prim.so unchanged.
//
//   histogram:  initHistogram64/256, closeHistogram64/256, histogram64, histogram256
//               (Samples/2_Concepts_and_Techniques/histogram/histogram_common.h:75-82)
//   scan:       initScan, closeScan, scanExclusiveShort, scanExclusiveLarge + the five limit constants
//               (Samples/2_Concepts_and_Techniques/scan/scan_common.h:41-55, scan.cu:175-179)
//   mergeSort:  initMergeSort, closeMergeSort, mergeSort
//               (Samples/0_Introduction/mergeSort/mergeSort_common.h:49-54)
//   reduction:  template<class T> void reduce(int,int,int,int,T*,T*) for int/float/double
//               (Samples/2_Concepts_and_Techniques/reduction/reduction.h:32, reduction_kernel.cu:985-989)
//
// Conventions kept from the samples: legacy default stream, no return codes, a CUDA error prints to
// stderr and exits (Common/helper_cuda.h:584-616).  init*() only warms the scratch up; it is optional.
#include "common.cuh"
#include "internal.h"

typedef unsigned int uint;

extern "C" {

// ---- histogram ------------------------------------------------------------------------------------
void initHistogram64(void)
{
    int st = 0;
    if (!b200::ctx(&st))
        B200_OR_DIE(st, "initHistogram64");
}
void initHistogram256(void)
{
    int st = 0;
    if (!b200::ctx(&st))
        B200_OR_DIE(st, "initHistogram256");
}
// close*() of one primitive frees that primitive's scratch only (the samples' close* free their own buffer:
// histogram256.cu:141-143, scan.cu:189, mergeSort.cu:463-469)
void closeHistogram64(void) { B200_OR_DIE(b200_release_scratch_of(0), "closeHistogram64"); }
void closeHistogram256(void) { B200_OR_DIE(b200_release_scratch_of(0), "closeHistogram256"); }

void histogram64(uint *d_Histogram, void *d_Data, uint byteCount)
{
    B200_OR_DIE(b200_histogram64_async(d_Histogram, d_Data, byteCount, 0), "histogram64");
}
void histogram256(uint *d_Histogram, void *d_Data, uint byteCount)
{
    B200_OR_DIE(b200_histogram256_async(d_Histogram, d_Data, byteCount, 0), "histogram256");
}

// ---- scan -------------------------------------------------------------------------------------------
// Same values as the sample so an unmodified scan/main.cpp iterates over the same lengths; the upper
// bounds are not enforced here (the chained scan has no length cap).
extern const uint MAX_BATCH_ELEMENTS   = 64 * 1048576;
extern const uint MIN_SHORT_ARRAY_SIZE = 4;
extern const uint MAX_SHORT_ARRAY_SIZE = 1024;
extern const uint MIN_LARGE_ARRAY_SIZE = 2048;
extern const uint MAX_LARGE_ARRAY_SIZE = 262144;

void initScan(void)
{
    // the sample allocates its scratch here (scan.cu:184-187); pre-size ours for MAX_BATCH_ELEMENTS
    B200_OR_DIE(b200_scan_reserve(64u * 1048576u), "initScan");
}
void closeScan(void) { B200_OR_DIE(b200_release_scratch_of(1), "closeScan"); }

size_t scanExclusiveShort(uint *d_Dst, uint *d_Src, uint batchSize, uint arrayLength)
{
    B200_OR_DIE(b200_scan_exclusive_async(d_Dst, d_Src, batchSize, arrayLength, nullptr, 0), "scanExclusiveShort");
    return 512; // CTA size of the scan kernel (the sample returns its THREADBLOCK_SIZE; only printed)
}
size_t scanExclusiveLarge(uint *d_Dst, uint *d_Src, uint batchSize, uint arrayLength)
{
    B200_OR_DIE(b200_scan_exclusive_async(d_Dst, d_Src, batchSize, arrayLength, nullptr, 0), "scanExclusiveLarge");
    return 512;
}

// ---- mergeSort ----------------------------------------------------------------------------------------
void initMergeSort(void)
{
    // the sample allocates its scratch here (mergeSort.cu:455-461); pre-size ours for its maximum N = 2^22
    B200_OR_DIE(b200_sort_reserve(4u * 1048576u), "initMergeSort");
}
void closeMergeSort(void) { B200_OR_DIE(b200_release_scratch_of(2), "closeMergeSort"); }

void mergeSort(uint *dstKey, uint *dstVal, uint *bufKey, uint *bufVal, uint *srcKey, uint *srcVal, uint N,
               uint sortDir)
{
    B200_OR_DIE(b200_sort_kv_async(dstKey, dstVal, bufKey, bufVal, srcKey, srcVal, N, sortDir, 0), "mergeSort");
}

} // extern "C"

// ---- reduction (C++ linkage: the sample links the mangled instantiations) --------------------------------
template <class T> void reduce(int size, int threads, int blocks, int whichKernel, T *d_idata, T *d_odata);

template <> void reduce<int>(int size, int, int blocks, int, int *d_idata, int *d_odata)
{
    B200_OR_DIE(b200_reduce_i32_async(d_idata, size < 0 ? 0 : size, d_odata, blocks < 1 ? 1 : blocks, 0),
                "reduce<int>");
}
template <> void reduce<float>(int size, int, int blocks, int, float *d_idata, float *d_odata)
{
    B200_OR_DIE(b200_reduce_f32_async(d_idata, size < 0 ? 0 : size, d_odata, blocks < 1 ? 1 : blocks, 0),
                "reduce<float>");
}
template <> void reduce<double>(int size, int, int blocks, int, double *d_idata, double *d_odata)
{
    B200_OR_DIE(b200_reduce_f64_async(d_idata, size < 0 ? 0 : size, d_odata, blocks < 1 ? 1 : blocks, 0),
                "reduce<double>");
}
