/*
 * Drop-in declaration for the reduction sample's launcher.
 * Replaces: Samples/2_Concepts_and_Techniques/reduction/reduction.h:32; libb200prim.so exports the
 * three explicit instantiations the sample links against (reduction_kernel.cu:985-989):
 *   _Z6reduceIiEviiiiPT_S1_   _Z6reduceIfEviiiiPT_S1_   _Z6reduceIdEviiiiPT_S1_
 *
 * Contract kept: after the call  sum_{b<blocks} d_odata[b] == sum_{i<size} d_idata[i]
 * (reduction.cpp:292-318 adds the partials up again, or finishes on the CPU, :279-289).
 * This library reduces in ONE pass and writes d_odata[0] = total, d_odata[1..blocks) = 0.
 * `threads` and `whichKernel` (0..9) are accepted and ignored: all ten teaching kernels of the
 * sample collapse into the one sm_100a kernel.  C twins: b200_reduce_{i32,f32,f64}_async.
 */
#ifndef __REDUCTION_H__
#define __REDUCTION_H__

#ifdef __cplusplus
template <class T> void reduce(int size, int threads, int blocks, int whichKernel, T *d_idata, T *d_odata);
#endif

#endif
