// compact.cu -- stable stream compaction (keep src[i] > 0) for sm_100a: the consumer either side of the scan.
//
// Replaces filter_arr + atomicAggInc (Samples/3_CUDA_Features/warpAggregatedAtomicsCG/
// warpAggregatedAtomicsCG.cu:43-68: one warp-aggregated global atomic per warp per iteration, output order
// arbitrary, grid-strided 4-byte loads) with a single-pass chained-scan compaction:
//   * tiles of 8192 ints in ticket order; warp w of the tile owns 512 consecutive ints, read as four coalesced
//     16-byte vectors per lane;
//   * flags -> one packed warp scan (the four per-vector counts of a lane ride in the bytes of one u32);
//   * the tile count goes to a look-back descriptor (2 status bits + 30-bit count); warp 0 sums the earlier tiles
//     of its 32-tile chunk and walks back over chunk descriptors (two-level look-back);
//   * kept values are compacted in shared memory and leave as one contiguous run;
//   * the output keeps the input order, which makes it bit-exact against the sample's host loop
//     (main(): `if (data[i] > 0) host_filtered[count++] = data[i]`, :287-292); the sample itself only compares
//     the counts because its GPU order is arbitrary.
// Long aligned inputs run on the chained scan's TMA-staged pipeline instead (scan.cu, scan_tma_kernel<.., kCompact>).
// Traffic: 4 B read per element + 4 B written per kept element; HBM-bound.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int      kCompThreads = 512;
constexpr int      kCompWarps   = kCompThreads / 32;
constexpr unsigned kCompTile    = kCompThreads * 16;
constexpr unsigned kCompA = 1u << 30, kCompP = 2u << 30, kCompValue = (1u << 30) - 1;

// v[0..4) = the four vectors of this lane in tile `tile` (zeros beyond the data: 0 is never kept)
__device__ __forceinline__ void compact_load(int4 (&v)[4], const int *__restrict__ src, size_t n, unsigned tile,
                                             unsigned wbase)
{
    const size_t   first = (size_t)tile * kCompTile;
    const size_t   left  = n - first;
    const unsigned valid = left < kCompTile ? (unsigned)left : kCompTile;
    if (valid == kCompTile && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
#pragma unroll
        for (int k = 0; k < 4; k++)
            v[k] = __ldg(reinterpret_cast<const int4 *>(src + first + wbase + k * 128));
    }
    else {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const unsigned e = wbase + k * 128;
            v[k].x = e + 0 < valid ? __ldg(src + first + e + 0) : 0;
            v[k].y = e + 1 < valid ? __ldg(src + first + e + 1) : 0;
            v[k].z = e + 2 < valid ? __ldg(src + first + e + 2) : 0;
            v[k].w = e + 3 < valid ? __ldg(src + first + e + 3) : 0;
        }
    }
}

// Register-staged fallback (short, unaligned or n % 4 != 0 inputs): one tile per CTA in ticket order.  Long aligned
// inputs take the TMA-staged role-split pipeline of the chained scan (scan.cu, kCompact: 4.4 TB/s at 2^28, 80 % kept);
// this kernel reaches 2.1-2.4 TB/s there (a tile keeps its bytes in flight for only a quarter of its lifetime; a
// persistent variant that prefetched the next tile into registers was slower still: 1.9 TB/s).
__global__ void __launch_bounds__(kCompThreads, 3)
    compact_kernel(const int *__restrict__ src, size_t n, int *__restrict__ dst, int *__restrict__ count_out,
                   unsigned *__restrict__ ticket, unsigned *__restrict__ desc, unsigned tiles)
{
    __shared__ int      s_out[kCompTile];
    __shared__ unsigned s_warp[kCompWarps];
    __shared__ unsigned s_tile, s_base;
    const unsigned      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned      wbase = warp * 512 + lane * 4; // + k * 128: element (inside the tile) of vector k, component x
    if (tid == 0)
        s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned tile = s_tile;
    int4           v[4];
    compact_load(v, src, n, tile, wbase);
    {
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
        if (lane == 0)
            s_warp[warp] = row_start[3] + (tot >> 24);
        __syncthreads();
        unsigned warp_off = 0, tile_total = 0;
#pragma unroll
        for (int w = 0; w < kCompWarps; w++) {
            const unsigned c = s_warp[w];
            warp_off += (w < (int)warp) ? c : 0u;
            tile_total += c;
        }
        if (tid == 0) // publish early: successors only need the count
            st_relaxed_u32(desc + tile, kCompA | tile_total);
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

        // ---- two-level look-back (warp 0) ------------------------------------------------------------------
        // Level 1: the <= 31 earlier tiles of the own 32-tile chunk (one poll, every lane its own descriptor).
        // Level 2: chunk descriptors (aggregate published by the chunk's last tile, prefix once known), 32 per poll.
        if (warp == 0) {
            const unsigned chunk = tile >> 5, j = tile & 31;
            unsigned       mine  = 0;
            if (lane < j) {
                const unsigned *dp = desc + (chunk << 5) + lane;
                unsigned        d;
                do {
                    d = ld_relaxed_u32(dp);
                } while ((d >> 30) == 0);
                mine = d & kCompValue;
            }
            const unsigned local = __reduce_add_sync(0xffffffffu, mine); // kept in the chunk's tiles before this one
            unsigned      *cdesc = desc + tiles;
            const bool     closes = j == 31; // this tile completes its chunk
            if (closes && lane == 0)
                st_relaxed_u32(cdesc + chunk, (chunk == 0 ? kCompP : kCompA) | (local + tile_total));
            unsigned cprefix = 0;
            if (chunk > 0) {
                long long t    = (long long)chunk - 1;
                bool      done = false;
                while (!done) { // nearest chunk first
                    const long long at    = t - lane;
                    const unsigned  d     = at >= 0 ? ld_relaxed_u32(cdesc + at) : kCompP;
                    const unsigned  ready = __ballot_sync(0xffffffffu, (d >> 30) != 0);
                    const unsigned  isp   = __ballot_sync(0xffffffffu, (d & kCompP) != 0);
                    const unsigned  run   = ~ready ? (__ffs(~ready) - 1) : 32;     // lanes 0..run-1 are published
                    const unsigned  stop  = isp ? (unsigned)(__ffs(isp) - 1) : 32; // first lane holding a prefix
                    const unsigned  take  = stop < run ? stop + 1 : run;
                    cprefix += __reduce_add_sync(0xffffffffu, lane < take ? (d & kCompValue) : 0u);
                    done = stop < run;
                    t -= take;
                }
                if (closes && lane == 0)
                    st_relaxed_u32(cdesc + chunk, kCompP | (cprefix + local + tile_total));
            }
            if (lane == 0) {
                s_base = cprefix + local;
                if (tile == tiles - 1)
                    *count_out = (int)(cprefix + local + tile_total);
            }
        }
        __syncthreads();
        int *out = dst + s_base;
        for (unsigned i = tid; i < tile_total; i += kCompThreads)
            out[i] = s_out[i];
    }
}

} // namespace b200

using namespace b200;

int compact_tma(int32_t *d_dst, int32_t *d_count, const int32_t *d_src, size_t n, cudaStream_t stream); // scan.cu

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
    { // long aligned inputs: the TMA-staged, role-split pipeline of the chained scan (scan.cu, kCompact)
        const int r = compact_tma(d_dst, d_count, d_src, n, stream);
        if (r != (int)cudaErrorNotSupported)
            return r;
    }
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const unsigned tiles = (unsigned)((n + kCompTile - 1) / kCompTile);
    const size_t   need  = 16 + 4 * ((size_t)tiles + (tiles + 31) / 32); // ticket | tile | chunk descriptors
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    B200_TRY(cudaMemsetAsync(c->sort.ptr, 0, need, stream));
    unsigned *ticket = reinterpret_cast<unsigned *>(c->sort.ptr);
    compact_kernel<<<tiles, kCompThreads, 0, stream>>>(d_src, n, d_dst, d_count, ticket, ticket + 4, tiles);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}
