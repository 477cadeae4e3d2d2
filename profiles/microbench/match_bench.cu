// Micro-benchmark: cost of building the "same 8-bit digit" lane mask inside a warp, three ways.
//   ballot : 8 x (bit test + vote.ballot + select)        -- what CUB-style onesweep ranking does
//   match  : one MATCH.ANY instruction (__match_any_sync)
//   atomor : ATOMS.OR of the lane bit into a warp-private shared table + read back + leader reset
// Prints ns per warp-round at full occupancy (148 SMs x 2048 threads).  nvcc -arch=sm_100a.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ROUNDS = 4096;

__device__ __forceinline__ unsigned next_digit(unsigned &x)
{
    x = x * 1664525u + 1013904223u;
    return x >> 24;
}

template <int MODE> __global__ void __launch_bounds__(512) k(unsigned *out)
{
    __shared__ unsigned tab[16][256];
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 16 * 256; i += 512)
        (&tab[0][0])[i] = 0;
    __syncthreads();
    unsigned x = threadIdx.x * 2654435761u + blockIdx.x, acc = 0;
    unsigned lt = (1u << lane) - 1u;
    for (int r = 0; r < ROUNDS; r++) {
        unsigned d = next_digit(x), mask;
        if (MODE == 0) {
            mask = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; b++) {
                bool     bit  = (d >> b) & 1u;
                unsigned vote = __ballot_sync(0xffffffffu, bit);
                mask &= bit ? vote : ~vote;
            }
        }
        else if (MODE == 1) {
            mask = __match_any_sync(0xffffffffu, d);
        }
        else {
            atomicOr(&tab[warp][d], 1u << lane);
            __syncwarp();
            mask = tab[warp][d];
            __syncwarp();
            if ((mask & lt) == 0)
                tab[warp][d] = 0;
            __syncwarp();
        }
        acc += __popc(mask & lt);
    }
    out[blockIdx.x * 512 + threadIdx.x] = acc;
}

template <int MODE> float run(unsigned *out, int grid)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k<MODE><<<grid, 512>>>(out);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    k<MODE><<<grid, 512>>>(out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int main()
{
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * 4;
    unsigned *out;
    cudaMalloc(&out, (size_t)grid * 512 * 4);
    const char *names[3] = {"ballot x8", "MATCH.ANY", "ATOMS.OR+LDS"};
    float       ms[3]    = {run<0>(out, grid), run<1>(out, grid), run<2>(out, grid)};
    unsigned    chk[3];
    for (int m = 0; m < 3; m++) {
        // all three must produce identical rank sums
        if (m == 0) k<0><<<grid, 512>>>(out);
        if (m == 1) k<1><<<grid, 512>>>(out);
        if (m == 2) k<2><<<grid, 512>>>(out);
        cudaMemcpy(&chk[m], out + 12345, 4, cudaMemcpyDeviceToHost);
    }
    const double warp_rounds = (double)grid * 16 * ROUNDS;
    for (int m = 0; m < 3; m++)
        printf("%-14s %8.3f ms  %7.3f ns per warp-round per SM-slot  (%.2f Gitems/s chip)  check %u\n", names[m], ms[m],
               ms[m] * 1e6 / (warp_rounds / sms), warp_rounds * 32 / ms[m] / 1e6, chk[m]);
    return cudaGetLastError() != cudaSuccess;
}
