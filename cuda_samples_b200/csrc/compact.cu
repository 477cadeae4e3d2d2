// compact.cu -- stable stream compaction (keep src[i] > 0) for sm_100a: the consumer either side of the scan.
//
// Replaces filter_arr + atomicAggInc (Samples/3_CUDA_Features/warpAggregatedAtomicsCG/
// warpAggregatedAtomicsCG.cu:43-68: one warp-aggregated global atomic per warp per iteration, output order
// arbitrary, grid-strided 4-byte loads) with a single-pass chained-scan compaction:
//   * tiles of 8192 ints in ticket order; warp w of the tile owns 512 consecutive ints, read as four coalesced
//     16-byte vectors per lane;
//   * flags -> one packed warp scan (the four per-vector counts of a lane ride in the bytes of one u32);
//   * the tile count goes to a look-back descriptor (2 status bits + 30-bit count), warp 0 walks back;
//   * kept values are compacted in shared memory and leave as one contiguous run;
//   * the output keeps the input order, which makes it bit-exact against the sample's host loop
//     (main(): `if (data[i] > 0) host_filtered[count++] = data[i]`, :287-292); the sample itself only compares
//     the counts because its GPU order is arbitrary.
// Traffic: 4 B read per element + 4 B written per kept element; HBM-bound.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int      kCompThreads = 512;
constexpr int      kCompWarps   = kCompThreads / 32;
constexpr unsigned kCompTile    = kCompThreads * 16;
constexpr unsigned kCompA = 1u << 30, kCompP = 2u << 30, kCompValue = (1u << 30) - 1;

__global__ void __launch_bounds__(kCompThreads, 3)
    compact_kernel(const int *__restrict__ src, size_t n, int *__restrict__ dst, int *__restrict__ count_out,
                   unsigned *__restrict__ ticket, unsigned *__restrict__ desc, unsigned tiles)
{
    __shared__ int      s_out[kCompTile];
    __shared__ unsigned s_warp[kCompWarps];
    __shared__ unsigned s_tile, s_base;
    const unsigned      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0)
        s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile  = s_tile;
    const size_t   first = (size_t)tile * kCompTile;
    const size_t   left  = n - first;
    const unsigned valid = left < kCompTile ? (unsigned)left : kCompTile;
    const unsigned wbase = warp * 512 + lane * 4; // + k * 128: element (inside the tile) of component x of vector k

    int4       v[4];
    const bool vec_ok = valid == kCompTile && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    if (vec_ok) {
#pragma unroll
        for (int k = 0; k < 4; k++)
            v[k] = __ldg(reinterpret_cast<const int4 *>(src + first + wbase + k * 128));
    }
    else {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const unsigned e = wbase + k * 128;
            v[k].x = e + 0 < valid ? __ldg(src + first + e + 0) : 0; // 0 is never kept
            v[k].y = e + 1 < valid ? __ldg(src + first + e + 1) : 0;
            v[k].z = e + 2 < valid ? __ldg(src + first + e + 2) : 0;
            v[k].w = e + 3 < valid ? __ldg(src + first + e + 3) : 0;
        }
    }
    // per-vector keep counts of this lane, byte k of `packed` (each <= 4, a warp sums to <= 128 per byte)
    unsigned packed = 0;
#pragma unroll
    for (int k = 0; k < 4; k++)
        packed |= (unsigned)((v[k].x > 0) + (v[k].y > 0) + (v[k].z > 0) + (v[k].w > 0)) << (8 * k);
    unsigned inc = packed;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d)
            inc += t;
    }
    const unsigned tot  = __shfl_sync(0xffffffffu, inc, 31); // byte k: kept in vector row k of the warp
    const unsigned excl = inc - packed;
    // order inside the warp is (k, lane, component): rows before k, then lanes before this one in row k
    unsigned row_start[4];
    row_start[0] = 0;
    row_start[1] = tot & 255u;
    row_start[2] = row_start[1] + ((tot >> 8) & 255u);
    row_start[3] = row_start[2] + ((tot >> 16) & 255u);
    const unsigned warp_total = row_start[3] + (tot >> 24);
    if (lane == 0)
        s_warp[warp] = warp_total;
    __syncthreads();
    unsigned warp_off = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < kCompWarps; w++) {
        const unsigned c = s_warp[w];
        warp_off += (w < (int)warp) ? c : 0u;
        tile_total += c;
    }
    // ---- publish the tile count, look back (warp 0), while everyone compacts into shared memory -----------
    if (warp == 0) {
        if (lane == 0)
            st_relaxed_u32(desc + tile, (tile == 0 ? kCompP : kCompA) | tile_total);
        unsigned excl_tiles = 0;
        if (tile > 0) {
            long long t    = (long long)tile - 1;
            bool      done = false;
            while (!done) { // 32 descriptors per poll, nearest first
                const long long mine = t - lane;
                unsigned        d    = mine >= 0 ? ld_relaxed_u32(desc + mine) : kCompP;
                const unsigned  ready = __ballot_sync(0xffffffffu, (d >> 30) != 0);
                const unsigned  isp   = __ballot_sync(0xffffffffu, (d & kCompP) != 0);
                // usable lanes: the run of published descriptors from lane 0 up to the first prefix (inclusive)
                const unsigned run  = ~ready ? (__ffs(~ready) - 1) : 32;          // lanes 0..run-1 are ready
                const unsigned stop = isp ? (unsigned)(__ffs(isp) - 1) : 32;      // first prefix lane
                const unsigned take = stop < run ? stop + 1 : run;                // lanes 0..take-1 are summed
                unsigned       part = lane < take ? (d & kCompValue) : 0u;
                part                = __reduce_add_sync(0xffffffffu, part);
                excl_tiles += part;
                done = stop < run;
                t -= take;
            }
            if (lane == 0)
                st_relaxed_u32(desc + tile, kCompP | (excl_tiles + tile_total));
        }
        if (lane == 0) {
            s_base = excl_tiles;
            if (tile == tiles - 1)
                *count_out = (int)(excl_tiles + tile_total);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        unsigned pos = warp_off + row_start[k] + ((excl >> (8 * k)) & 255u);
        if (v[k].x > 0)
            s_out[pos++] = v[k].x;
        if (v[k].y > 0)
            s_out[pos++] = v[k].y;
        if (v[k].z > 0)
            s_out[pos++] = v[k].z;
        if (v[k].w > 0)
            s_out[pos++] = v[k].w;
    }
    __syncthreads();
    int *out = dst + s_base;
    for (unsigned i = tid; i < tile_total; i += kCompThreads)
        out[i] = s_out[i];
}

} // namespace b200

using namespace b200;

// dst[0 .. *d_count) = the elements of src[0 .. n) that are > 0, in input order; *d_count = how many.
// dst must hold n ints in the worst case; n < 2^30.  Scratch: the sort's (ticket + one u32 per tile).
extern "C" int b200_compact_positive_async(int32_t *d_dst, int32_t *d_count, const int32_t *d_src, size_t n,
                                           b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n >= (1ull << 30))
        return (int)cudaErrorInvalidValue;
    if (n == 0) {
        B200_TRY(cudaMemsetAsync(d_count, 0, 4, stream));
        return 0;
    }
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const unsigned tiles = (unsigned)((n + kCompTile - 1) / kCompTile);
    const size_t   need  = 16 + 4 * (size_t)tiles;
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    B200_TRY(cudaMemsetAsync(c->sort.ptr, 0, need, stream));
    unsigned *ticket = reinterpret_cast<unsigned *>(c->sort.ptr);
    compact_kernel<<<tiles, kCompThreads, 0, stream>>>(d_src, n, d_dst, d_count, ticket, ticket + 4, tiles);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}
