// generate.cu -- counter-based synthetic inputs, bit-identical to oracle_generate() in oracle/oracle.c.
// element i of the stream = f(seed, first + i): every GPU fills its own shard in place and the CPU can
// regenerate any chunk for the gold path.  Not on the timed path.
#include "common.cuh"
#include "internal.h"

namespace b200 {

__host__ __device__ __forceinline__ unsigned long long splitmix64(unsigned long long x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ __forceinline__ unsigned bijection32(unsigned x)
{
    x *= 2654435761u;
    x ^= x >> 15;
    x *= 2246822519u;
    x ^= x >> 13;
    x *= 3266489917u;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ unsigned gen_half(unsigned long long seed, unsigned long long idx32)
{
    unsigned long long w = splitmix64(seed + (idx32 >> 1));
    return (unsigned)((idx32 & 1) ? (w >> 32) : w);
}

__global__ void generate_kernel(void *out, size_t first, size_t n, unsigned long long seed, int kind, unsigned param)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned long long g = first + i;
        switch (kind) {
        case 0: {
            unsigned long long w            = splitmix64(seed + (g >> 3));
            ((unsigned char *)out)[i] = (unsigned char)(w >> (8 * (g & 7)));
            break;
        }
        case 1: ((unsigned *)out)[i] = gen_half(seed, g) >> 1; break;
        case 2: ((int *)out)[i] = (int)(gen_half(seed, g) & 0xFF); break;
        case 3: ((float *)out)[i] = __fdiv_rn((float)(gen_half(seed, g) & 0xFF), 2147483647.0f); break;
        case 4: ((unsigned *)out)[i] = bijection32((unsigned)g ^ (unsigned)seed); break;
        case 5: ((unsigned *)out)[i] = bijection32((unsigned)(g % param)); break;
        case 6: ((unsigned *)out)[i] = gen_half(seed, g) % param; break;
        case 7: ((double *)out)[i] = __ddiv_rn((double)(gen_half(seed, g) & 0xFF), 2147483647.0); break;
        }
    }
}

} // namespace b200

using namespace b200;

extern "C" int b200_generate_async(void *d_out, size_t first, size_t n, uint64_t seed, int kind, uint32_t param,
                                   b200_stream_t stream)
{
    if (kind < 0 || kind > 7 || ((kind == 5 || kind == 6) && param == 0))
        return (int)cudaErrorInvalidValue;
    if (n == 0)
        return 0;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    size_t want = (n + 255) / 256, cap = (size_t)c->num_sms * 16;
    int    grid = (int)(want < cap ? want : cap);
    generate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_out, first, n, seed, kind, param);
    B200_TRY(cudaGetLastError());
    return 0;
}
