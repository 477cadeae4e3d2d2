/*
 * Drop-in declarations for the scan sample's host entry points.
 * Replaces: Samples/2_Concepts_and_Techniques/scan/scan_common.h:41-55 and the limit constants
 * defined in scan.cu:175-179 (kept with the same values so the unmodified scan/main.cpp loops
 * over the same array lengths; the upper bounds are NOT enforced by this library).
 *
 * Behaviour kept: size_t return (a workgroup size that main.cpp only prints, scan.cu:230,265),
 * device pointers, legacy default stream, abort on CUDA error, batchSize independent arrays.
 * Relaxed: any arrayLength >= 1 (no power-of-two requirement), no size caps, batchSize == 0 is a
 * no-op, initScan() optional.  Both entry points run the same single-pass kernel family.
 */
#ifndef SCAN_COMMON_H
#define SCAN_COMMON_H

#include <stdlib.h>

typedef unsigned int uint;

#ifdef __cplusplus
extern "C" {
#endif

extern const uint MAX_BATCH_ELEMENTS;
extern const uint MIN_SHORT_ARRAY_SIZE;
extern const uint MAX_SHORT_ARRAY_SIZE;
extern const uint MIN_LARGE_ARRAY_SIZE;
extern const uint MAX_LARGE_ARRAY_SIZE;

void initScan(void);
void closeScan(void);

size_t scanExclusiveShort(uint *d_Dst, uint *d_Src, uint batchSize, uint arrayLength);
size_t scanExclusiveLarge(uint *d_Dst, uint *d_Src, uint batchSize, uint arrayLength);

#ifdef __cplusplus
}
#endif
#endif
