/*
 * Drop-in declarations for the mergeSort sample's host entry points.
 * Replaces: Samples/0_Introduction/mergeSort/mergeSort_common.h:49-54.
 *
 * mergeSort() here is a stable onesweep LSD radix sort; the contract of the sample is kept:
 * stable key-value sort of N (key,value) u32 pairs, sortDir != 0 ascending / == 0 descending,
 * result in dst*, buf* clobbered, src* untouched, device pointers, legacy default stream.
 * Relaxed: any N (the sample asserts N % 1024 == 0 and N <= 2^22, mergeSort.cu:500-501);
 * initMergeSort() optional.
 */
#ifndef MERGESORT_COMMON_H
#define MERGESORT_COMMON_H

typedef unsigned int uint;

#ifdef __cplusplus
extern "C" {
#endif

void initMergeSort(void);
void closeMergeSort(void);

void mergeSort(uint *dstKey, uint *dstVal, uint *bufKey, uint *bufVal, uint *srcKey, uint *srcVal, uint N,
               uint sortDir);

#ifdef __cplusplus
}
#endif
#endif
