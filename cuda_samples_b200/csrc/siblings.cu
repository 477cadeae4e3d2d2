// siblings.cu -- the single-pass reduction siblings of the hot path, on the same kernels:
//   threadFenceReduction   extern "C" reduce / reduceSinglePass (int size, int threads, int blocks, float*, float*)
//                          Samples/2_Concepts_and_Techniques/threadFenceReduction/threadFenceReduction_kernel.cuh:213,308
//   reductionMultiBlockCG  call_reduceSinglePassMultiBlockCG(size, threads, numBlocks, d_idata, d_odata)
//                          Samples/2_Concepts_and_Techniques/reductionMultiBlockCG/reductionMultiBlockCG.cu:113-160
//   scalarProd             scalarProdGPU<<<128,256>>>(d_C, d_A, d_B, vectorN, elementN) -- batched dot products
//                          Samples/2_Concepts_and_Techniques/scalarProd/scalarProd_kernel.cuh, gold scalarProd_cpu.cpp:32-46
// The first two are the sum reduction with a different finish in the samples (last-block ticket / grid.sync);
// here they ARE reduce_kernel<LaneF32>: d_odata[0] = sum, d_odata[1..blocks) = 0.
// scalarProd is a reduce of products: one CTA per (vector, chunk), 16-byte loads of A and B, each product rounded
// to fp32 exactly like the gold code (float * float) and accumulated in fp64 like the gold code (double sum),
// fixed-order finish -> deterministic.  8 algorithmic bytes per element pair, HBM-bound.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kDotThreads = 256;

__device__ __forceinline__ double dot4(uint4 a, uint4 b)
{
    double s = (double)__fmul_rn(__uint_as_float(a.x), __uint_as_float(b.x));
    s += (double)__fmul_rn(__uint_as_float(a.y), __uint_as_float(b.y));
    s += (double)__fmul_rn(__uint_as_float(a.z), __uint_as_float(b.z));
    s += (double)__fmul_rn(__uint_as_float(a.w), __uint_as_float(b.w));
    return s;
}

// grid = vectorN * chunks.  CTA (v, c) handles elements [c*span, min((c+1)*span, elementN)) of vector v.
__global__ void __launch_bounds__(kDotThreads)
    scalar_prod_kernel(float *__restrict__ C, double *__restrict__ partial, const float *__restrict__ A,
                       const float *__restrict__ B, int elementN, int chunks, size_t span)
{
    __shared__ double s_part[kDotThreads / 32];
    const unsigned v = blockIdx.x / chunks, c = blockIdx.x - v * chunks;
    const size_t   lo = (size_t)c * span;
    size_t         hi = lo + span;
    if (hi > (size_t)elementN)
        hi = elementN;
    const float *a = A + (size_t)v * elementN, *b = B + (size_t)v * elementN;
    double       acc = 0.0;
    const bool   vec = ((((uintptr_t)(a + lo)) | ((uintptr_t)(b + lo))) & 15) == 0;
    size_t       i   = lo;
    if (vec && hi > lo) {
        const size_t nvec = (hi - lo) / 4;
        const uint4 *va = reinterpret_cast<const uint4 *>(a + lo), *vb = reinterpret_cast<const uint4 *>(b + lo);
        size_t       k  = threadIdx.x;
        for (; k + 3 * kDotThreads < nvec; k += 4 * kDotThreads) { // 8 independent 16-byte loads in flight
            uint4 ra[4], rb[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                ra[u] = ld_stream_u4(va + k + u * kDotThreads);
                rb[u] = ld_stream_u4(vb + k + u * kDotThreads);
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
                acc += dot4(ra[u], rb[u]);
        }
        for (; k < nvec; k += kDotThreads)
            acc += dot4(ld_stream_u4(va + k), ld_stream_u4(vb + k));
        i = lo + nvec * 4;
    }
    for (size_t j = i + threadIdx.x; j < hi; j += kDotThreads)
        acc += (double)__fmul_rn(a[j], b[j]);
    // fixed-order CTA sum
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0)
        s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < kDotThreads / 32; w++)
            t += s_part[w];
        if (chunks == 1)
            C[v] = (float)t;
        else
            partial[blockIdx.x] = t;
    }
}

__global__ void scalar_prod_finish_kernel(float *__restrict__ C, const double *__restrict__ partial, int vectorN,
                                          int chunks)
{
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= vectorN)
        return;
    double t = 0.0;
    for (int c = 0; c < chunks; c++)
        t += partial[(size_t)v * chunks + c];
    C[v] = (float)t;
}

} // namespace b200

using namespace b200;

extern "C" int b200_scalar_prod_async(float *d_C, const float *d_A, const float *d_B, int vectorN, int elementN,
                                      b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (vectorN <= 0)
        return 0;
    if (elementN < 0)
        return (int)cudaErrorInvalidValue;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    // enough CTAs to fill the chip: split long vectors when there are few of them
    int chunks = 1;
    const int want = c->num_sms * 8;
    if (vectorN < want && elementN > 16384) {
        chunks     = (want + vectorN - 1) / vectorN;
        int maxchk = elementN / 8192;
        if (chunks > maxchk)
            chunks = maxchk;
        if (chunks < 1)
            chunks = 1;
    }
    size_t span = ((size_t)elementN + chunks - 1) / chunks;
    span        = (span + 3) & ~(size_t)3; // keep chunk starts 16-byte aligned relative to the vector
    double *partial = nullptr;
    if (chunks > 1) {
        B200_TRY((cudaError_t)ensure(c->stage[0], (size_t)vectorN * chunks * sizeof(double), stream));
        partial = reinterpret_cast<double *>(c->stage[0].ptr);
    }
    scalar_prod_kernel<<<(unsigned)vectorN * chunks, kDotThreads, 0, stream>>>(d_C, partial, d_A, d_B, elementN, chunks,
                                                                              span);
    B200_TRY(cudaGetLastError());
    count_launch();
    if (chunks > 1) {
        scalar_prod_finish_kernel<<<(vectorN + 255) / 256, 256, 0, stream>>>(d_C, partial, vectorN, chunks);
        B200_TRY(cudaGetLastError());
        count_launch();
    }
    return 0;
}

// ---- drop-in names of the sibling samples (legacy default stream, abort on error) ------------------------------
extern "C" void reduceSinglePass(int size, int threads, int blocks, float *d_idata, float *d_odata)
{
    (void)threads;
    B200_OR_DIE(b200_reduce_f32_async(d_idata, size < 0 ? 0 : size, d_odata, blocks < 1 ? 1 : blocks, 0),
                "reduceSinglePass");
}
// threadFenceReduction's multi-pass entry point has C linkage and the bare name `reduce`
extern "C" void reduce(int size, int threads, int blocks, float *d_idata, float *d_odata)
{
    (void)threads;
    B200_OR_DIE(b200_reduce_f32_async(d_idata, size < 0 ? 0 : size, d_odata, blocks < 1 ? 1 : blocks, 0), "reduce");
}
extern "C" void reduceSinglePassMultiBlockCG(int size, int threads, int numBlocks, float *d_idata, float *d_odata)
{
    (void)threads;
    B200_OR_DIE(b200_reduce_f32_async(d_idata, size < 0 ? 0 : size, d_odata, numBlocks < 1 ? 1 : numBlocks, 0),
                "reduceSinglePassMultiBlockCG");
}
extern "C" void scalarProd(float *d_C, float *d_A, float *d_B, int vectorN, int elementN)
{
    B200_OR_DIE(b200_scalar_prod_async(d_C, d_A, d_B, vectorN, elementN, 0), "scalarProd");
}
