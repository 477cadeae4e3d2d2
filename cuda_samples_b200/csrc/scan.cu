// scan.cu -- single-pass batched exclusive scan (u32, wrap mod 2^32) for sm_100a.
//
// Replaces scanExclusiveShort / scanExclusiveLarge and their three kernels (Samples/
// 2_Concepts_and_Techniques/scan/scan.cu: scanExclusiveShared :102, scanExclusiveShared2 :121,
// uniformUpdate :149 -- 16 B of traffic per element) with ONE launch that reads and writes every
// element once (8 B per element):
//   * a tile is 8192 elements; a CTA of 512 threads takes tiles in ticket order (atomic counter),
//     so a tile's predecessors are always already running -> the look-back cannot deadlock;
//   * in-tile: 16-byte loads in a warp-striped arrangement (warp = 4 rows of 128 elements, lane
//     = 4 consecutive elements of a row), serial scan of the 4, shuffle scan across lanes, running
//     carry across rows, shuffle scan of the 16 warp totals;
//   * across tiles: decoupled look-back over 64-bit {status|epoch, value} descriptors written with
//     one relaxed 64-bit store.  The epoch lives in device memory and is bumped by the CTA that
//     draws the last ticket, so descriptors never need clearing and the launch is graph-replayable;
//   * batch semantics of scanExclusiveHost (scan_gold.cpp:30-38): the carry restarts at every
//     arrayLength boundary.  Power-of-two arrays shorter than a tile are scanned as segments of the
//     lane / row / warp hierarchy with no look-back at all (the sample's "short" case); any other
//     length gets tile-aligned arrays and, if it breaks 16-byte alignment, guarded scalar accesses.
// Roofline: HBM-bound, 8 algorithmic bytes per element.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int      kScanThreads = 512;
constexpr int      kScanWarps   = kScanThreads / 32;
constexpr int      kScanRows    = 4;                                  // 16-byte vectors per thread
constexpr unsigned kScanTile    = kScanThreads * kScanRows * 4;       // 8192 elements
constexpr unsigned kStA = 1, kStP = 2;                                // descriptor states

struct ScanParams {
    size_t          total;      // batch * L
    size_t          L;          // array length
    unsigned        num_tiles;
    unsigned        tpa;        // tiles per array (arrays >= one tile), else 1
    unsigned        lane_seg;   // lanes per array inside a row      (32 when L >= 128)
    unsigned        rows_seg;   // rows per array inside a warp      (4 when L >= 512)
    unsigned        warps_seg;  // warps per array inside a tile     (16 when L >= 8192)
    unsigned        packed;     // 1: power-of-two L < tile, tiles cover whole arrays back to back
    const unsigned *carry;      // nullable: seed of element 0 of every array
};

__device__ __forceinline__ unsigned long long make_desc(unsigned epoch, unsigned state, unsigned value)
{
    return ((unsigned long long)((epoch << 2) | state) << 32) | value;
}

// Warp 0: sum of the aggregates of the tiles before `tile` back to the nearest published prefix.
__device__ __forceinline__ unsigned look_back(const unsigned long long *desc, unsigned tile, unsigned first_tile,
                                              unsigned epoch)
{
    const unsigned lane   = lane_id();
    const unsigned want_a = (epoch << 2) | kStA, want_p = (epoch << 2) | kStP;
    unsigned       prefix = 0;
    long long      base   = (long long)tile - 1;
    while (true) {
        const long long idx   = base - lane;
        const bool      valid = idx >= (long long)first_tile;
        unsigned        st = want_p, val = 0, pmask, firstp = 0;
        bool            ok;
        do {
            if (valid) {
                unsigned long long d = ld_relaxed_u64(desc + idx);
                st                   = (unsigned)(d >> 32);
                val                  = (unsigned)d;
            }
            pmask          = __ballot_sync(0xffffffffu, st == want_p);
            unsigned rmask = __ballot_sync(0xffffffffu, st == want_p || st == want_a);
            if (pmask) {
                firstp        = __ffs(pmask) - 1;
                unsigned need = (1u << firstp) - 1u;
                ok            = (rmask & need) == need;
            }
            else
                ok = (rmask == 0xffffffffu);
        } while (!ok);
        if (pmask) {
            prefix += __reduce_add_sync(0xffffffffu, lane <= firstp ? val : 0u);
            break;
        }
        prefix += __reduce_add_sync(0xffffffffu, val);
        base -= 32;
    }
    return prefix;
}

// kLarge: arrays of at least one tile (lane_seg=32, rows_seg=4, warps_seg=16 known at compile time)
template <bool kLarge>
__global__ void __launch_bounds__(kScanThreads)
    scan_kernel(unsigned *dst, const unsigned *src, const ScanParams p, // may alias (in place)
                unsigned *__restrict__ state, unsigned long long *__restrict__ desc)
{
    __shared__ unsigned s_tile, s_epoch, s_prefix;
    __shared__ unsigned s_warp_tot[kScanWarps];

    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // Only arrays spanning several tiles need ordered tickets and the look-back descriptors.
    const bool chained = kLarge && p.tpa > 1;
    unsigned   tile = blockIdx.x, epoch = 0;
    if (chained) {
        if (threadIdx.x == 0) {
            // epoch first (acquire), then the ticket: whoever holds a ticket has already read the
            // epoch, so the CTA that draws the last ticket may advance it for the next launch.
            unsigned ep = ld_acquire_u32(&state[1]);
            unsigned t  = atomicAdd(&state[0], 1u);
            if (t == p.num_tiles - 1) {
                st_relaxed_u32(&state[0], 0u);
                st_relaxed_u32(&state[1], (ep + 1u) & 0x3fffffffu);
            }
            s_tile  = t;
            s_epoch = ep;
        }
        __syncthreads();
        tile  = s_tile;
        epoch = s_epoch;
    }

    // ---- where is this tile ---------------------------------------------------------------------
    size_t   off;        // global element index of the tile's first element
    unsigned count;      // valid elements in the tile
    unsigned first_tile; // first tile of the array this tile belongs to (look-back never crosses it)
    if (!kLarge && p.packed) {
        off           = (size_t)tile * kScanTile;
        size_t remain = p.total - off;
        count         = remain < kScanTile ? (unsigned)remain : kScanTile;
        first_tile    = tile;
    }
    else {
        const unsigned a = tile / p.tpa, t = tile - a * p.tpa;
        off              = (size_t)a * p.L + (size_t)t * kScanTile;
        size_t remain    = p.L - (size_t)t * kScanTile;
        count            = remain < kScanTile ? (unsigned)remain : kScanTile;
        first_tile       = a * p.tpa;
    }
    const unsigned lane_seg  = kLarge ? 32u : p.lane_seg;
    const unsigned rows_seg  = kLarge ? (unsigned)kScanRows : p.rows_seg;
    const unsigned warps_seg = kLarge ? (unsigned)kScanWarps : p.warps_seg;

    // ---- load: warp-striped 16-byte vectors --------------------------------------------------------
    const bool fast = count == kScanTile && (((uintptr_t)(src + off) | (uintptr_t)(dst + off)) & 15) == 0;
    uint4      v[kScanRows];
    unsigned   ebase[kScanRows]; // element index inside the tile of v[k].x
#pragma unroll
    for (int k = 0; k < kScanRows; k++)
        ebase[k] = warp * (kScanRows * 128) + k * 128 + lane * 4;
    if (fast) {
#pragma unroll
        for (int k = 0; k < kScanRows; k++)
            v[k] = ld_once_u4(src + off + ebase[k]);
    }
    else {
#pragma unroll
        for (int k = 0; k < kScanRows; k++) {
            const unsigned e = ebase[k];
            v[k].x           = (e + 0 < count) ? src[off + e + 0] : 0u;
            v[k].y           = (e + 1 < count) ? src[off + e + 1] : 0u;
            v[k].z           = (e + 2 < count) ? src[off + e + 2] : 0u;
            v[k].w           = (e + 3 < count) ? src[off + e + 3] : 0u;
        }
    }

    // ---- lane level: totals of the 4-element vectors, segmented inclusive scan across lanes ----------
    unsigned tot[kScanRows], inc[kScanRows];
#pragma unroll
    for (int k = 0; k < kScanRows; k++) {
        tot[k] = (v[k].x + v[k].y) + (v[k].z + v[k].w);
        inc[k] = tot[k];
    }
    const unsigned seg_lane = lane & (lane_seg - 1);
    if (kLarge) {
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
#pragma unroll
            for (int k = 0; k < kScanRows; k++) {
                unsigned t = __shfl_up_sync(0xffffffffu, inc[k], d);
                if (lane >= d)
                    inc[k] += t;
            }
    }
    else {
        for (unsigned d = 1; d < lane_seg; d <<= 1)
#pragma unroll
            for (int k = 0; k < kScanRows; k++) {
                unsigned t = __shfl_up_sync(0xffffffffu, inc[k], d);
                if (seg_lane >= d)
                    inc[k] += t;
            }
    }
    // ---- row level: carry across the rows of one array inside the warp ------------------------------
    unsigned rowoff[kScanRows];
    unsigned run = 0;
#pragma unroll
    for (int k = 0; k < kScanRows; k++) {
        if (!kLarge && (k & (rows_seg - 1)) == 0)
            run = 0;
        rowoff[k] = run;
        run += __shfl_sync(0xffffffffu, inc[k], 31); // row total (only meaningful when lane_seg == 32)
    }
    // ---- warp level: segmented scan of the 16 warp totals ------------------------------------------
    unsigned warpoff = 0, aggregate = 0;
    if (kLarge || warps_seg > 1) {
        if (lane == 0)
            s_warp_tot[warp] = run;
        __syncthreads();
        unsigned x  = (lane < kScanWarps) ? s_warp_tot[lane] : 0u;
        unsigned xi = x;
        const unsigned seg_w = lane & (warps_seg - 1);
#pragma unroll
        for (int d = 1; d < kScanWarps; d <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, xi, d);
            if (seg_w >= (unsigned)d && (unsigned)d < warps_seg)
                xi += t;
        }
        warpoff   = __shfl_sync(0xffffffffu, xi - x, warp);
        aggregate = __shfl_sync(0xffffffffu, xi, kScanWarps - 1); // whole tile (when warps_seg == 16)
    }

    // ---- tile level: decoupled look-back -----------------------------------------------------------
    unsigned prefix = p.carry ? __ldg(p.carry) : 0u;
    if (chained) {
        if (warp == 0) {
            if (tile == first_tile) {
                if (lane == 0)
                    st_relaxed_u64(desc + tile, make_desc(epoch, kStP, prefix + aggregate));
            }
            else {
                if (lane == 0)
                    st_relaxed_u64(desc + tile, make_desc(epoch, kStA, aggregate));
                prefix = look_back(desc, tile, first_tile, epoch); // the first tile's prefix already holds the carry
                if (lane == 0)
                    st_relaxed_u64(desc + tile, make_desc(epoch, kStP, prefix + aggregate));
            }
            if (lane == 0)
                s_prefix = prefix;
        }
        __syncthreads();
        prefix = s_prefix;
    }

    // ---- write -------------------------------------------------------------------------------------
#pragma unroll
    for (int k = 0; k < kScanRows; k++) {
        const unsigned e0 = prefix + warpoff + rowoff[k] + (inc[k] - tot[k]);
        uint4          o;
        o.x = e0;
        o.y = e0 + v[k].x;
        o.z = o.y + v[k].y;
        o.w = o.z + v[k].z;
        if (fast) {
            st_stream_u4(dst + off + ebase[k], o);
        }
        else {
            const unsigned e = ebase[k];
            if (e + 0 < count)
                dst[off + e + 0] = o.x;
            if (e + 1 < count)
                dst[off + e + 1] = o.y;
            if (e + 2 < count)
                dst[off + e + 2] = o.z;
            if (e + 3 < count)
                dst[off + e + 3] = o.w;
        }
    }
}

static bool is_pow2(size_t x) { return x && !(x & (x - 1)); }

} // namespace b200

using namespace b200;

extern "C" int b200_scan_exclusive_async(uint32_t *d_dst, const uint32_t *d_src, size_t batch, size_t L,
                                         const uint32_t *d_carry, b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (batch == 0 || L == 0)
        return 0;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    ScanParams p;
    p.total = batch * L;
    p.L     = L;
    p.carry = d_carry;
    bool large;
    if (L >= kScanTile || !is_pow2(L) || L < 4) {
        // tile-aligned arrays (look-back inside an array when it spans several tiles)
        size_t tpa = (L + kScanTile - 1) / kScanTile;
        if (tpa * batch > 0xfffffff0ull || tpa > 0xffffffffull)
            return (int)cudaErrorInvalidValue;
        p.tpa       = (unsigned)tpa;
        p.num_tiles = (unsigned)(tpa * batch);
        p.lane_seg  = 32;
        p.rows_seg  = kScanRows;
        p.warps_seg = kScanWarps;
        p.packed    = 0;
        large       = true;
    }
    else {
        // power-of-two arrays shorter than a tile: tiles hold whole arrays back to back
        size_t tiles = (p.total + kScanTile - 1) / kScanTile;
        if (tiles > 0xfffffff0ull)
            return (int)cudaErrorInvalidValue;
        p.tpa       = 1;
        p.num_tiles = (unsigned)tiles;
        p.lane_seg  = (unsigned)(L / 4 < 32 ? L / 4 : 32);
        p.rows_seg  = (unsigned)(L / 128 < 1 ? 1 : (L / 128 > kScanRows ? kScanRows : L / 128));
        p.warps_seg = (unsigned)(L / 512 < 1 ? 1 : L / 512);
        p.packed    = 1;
        large       = false;
    }
    // scratch: [0] ticket, [1] epoch, 64-bit descriptors from byte 16
    size_t ndesc = (large && p.tpa > 1) ? p.num_tiles : 0;
    B200_TRY((cudaError_t)ensure(c->scan, 16 + 8 * ndesc, stream));
    unsigned           *state = reinterpret_cast<unsigned *>(c->scan.ptr);
    unsigned long long *desc  = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(c->scan.ptr) + 16);
    if (large)
        scan_kernel<true><<<p.num_tiles, kScanThreads, 0, stream>>>(d_dst, d_src, p, state, desc);
    else
        scan_kernel<false><<<p.num_tiles, kScanThreads, 0, stream>>>(d_dst, d_src, p, state, desc);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}

extern "C" int b200_scan_total_async(uint32_t *d_total, const uint32_t *d_src, size_t n, b200_stream_t stream)
{
    // wrapping u32 sum == the int reduction
    return b200_reduce_i32_async(reinterpret_cast<const int32_t *>(d_src), n, reinterpret_cast<int32_t *>(d_total), 1,
                                 stream);
}
