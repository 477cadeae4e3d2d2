/*
 * Drop-in declarations for the histogram sample's host entry points.
 * Replaces: Samples/2_Concepts_and_Techniques/histogram/histogram_common.h:68-82
 * (prototypes), :34-38 (bin counts, uint/uchar).  Implemented by libb200prim.so
 * (cuda_samples_b200/csrc/dropin.cu -> histogram.cu).
 *
 * Behaviour kept: void return, device pointers, d_Histogram overwritten, launches on the legacy
 * default stream, CUDA errors print to stderr and exit(EXIT_FAILURE).
 * Relaxed: any byteCount (the sample asserts %4 / %16 and caps histogram64 at ~480 MB,
 * histogram64.cu:190-191); init*() is optional (scratch is created on demand).
 */
#ifndef HISTOGRAM_COMMON_H
#define HISTOGRAM_COMMON_H

#define HISTOGRAM64_BIN_COUNT  64
#define HISTOGRAM256_BIN_COUNT 256

typedef unsigned int  uint;
typedef unsigned char uchar;

/* Only printed by the sample's log line ("Workgroup = %u", histogram/main.cpp:126-132,178-184):
 * the CTA size of the sm_100a kernel. */
#define HISTOGRAM64_THREADBLOCK_SIZE  1024
#define HISTOGRAM256_THREADBLOCK_SIZE 1024

#ifdef __cplusplus
extern "C" {
#endif

void initHistogram64(void);
void closeHistogram64(void);
void histogram64(uint *d_Histogram, void *d_Data, uint byteCount);

void initHistogram256(void);
void closeHistogram256(void);
void histogram256(uint *d_Histogram, void *d_Data, uint byteCount);

#ifdef __cplusplus
}
#endif
#endif
