/*
 * Drop-in declarations for the single-pass reduction siblings of the hot path.
 *
 *   reduce / reduceSinglePass      Samples/2_Concepts_and_Techniques/threadFenceReduction/threadFenceReduction.cu:89-93
 *                                  (defined in threadFenceReduction_kernel.cuh:213,308)
 *   reduceSinglePassMultiBlockCG   the host wrapper call_reduceSinglePassMultiBlockCG of
 *                                  Samples/2_Concepts_and_Techniques/reductionMultiBlockCG/reductionMultiBlockCG.cu:113-160
 *   scalarProd                     host form of scalarProdGPU<<<128,256>>>(d_C, d_A, d_B, vectorN, elementN),
 *                                  Samples/2_Concepts_and_Techniques/scalarProd/scalarProd.cu:126; gold scalarProd_cpu.cpp:32-46
 *
 * All three reductions leave d_odata[0] = sum and d_odata[1..blocks) = 0 (the samples' multi-pass callers,
 * which reduce d_odata again, still get the right total).  `threads` is accepted and ignored.
 * Device pointers, legacy default stream, CUDA errors abort like the samples.
 */
#ifndef REDUCTION_SIBLINGS_H
#define REDUCTION_SIBLINGS_H

#ifdef __cplusplus
extern "C" {
#endif

void reduce(int size, int threads, int blocks, float *d_idata, float *d_odata);
void reduceSinglePass(int size, int threads, int blocks, float *d_idata, float *d_odata);
void reduceSinglePassMultiBlockCG(int size, int threads, int numBlocks, float *d_idata, float *d_odata);
/* d_C[v] = sum_i d_A[v*elementN + i] * d_B[v*elementN + i], products rounded to float, summed in double */
void scalarProd(float *d_C, float *d_A, float *d_B, int vectorN, int elementN);

#ifdef __cplusplus
}
#endif
#endif
