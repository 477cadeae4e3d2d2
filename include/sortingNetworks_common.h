/*
 * Drop-in declarations for the sortingNetworks sample's host entry points.
 * Replaces: Samples/2_Concepts_and_Techniques/sortingNetworks/sortingNetworks_common.h:45-55
 * (bitonicSort.cu, oddEvenMergeSort.cu).  batchSize independent arrays of arrayLength (key, value) u32 pairs,
 * dir != 0 ascending / == 0 descending, device pointers, legacy default stream, abort on CUDA error.
 * Relaxed: any arrayLength (the networks need powers of two); the result is additionally STABLE.
 * Implemented as a segmented radix sort on the mergeSort kernels (csrc/sorting_networks.cu).
 */
#ifndef SORTINGNETWORKS_COMMON_H
#define SORTINGNETWORKS_COMMON_H

typedef unsigned int uint;

#ifdef __cplusplus
extern "C" {
#endif

uint bitonicSort(uint *d_DstKey, uint *d_DstVal, uint *d_SrcKey, uint *d_SrcVal, uint batchSize, uint arrayLength,
                 uint dir);
void oddEvenMergeSort(uint *d_DstKey, uint *d_DstVal, uint *d_SrcKey, uint *d_SrcVal, uint batchSize,
                      uint arrayLength, uint dir);

#ifdef __cplusplus
}
#endif
#endif
