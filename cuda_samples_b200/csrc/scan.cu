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

#include <cstdio>
#include <mutex>

#ifndef B200_SCAN_LOAD_EVICT_FIRST
#define B200_SCAN_LOAD_EVICT_FIRST 0
#endif
#ifndef B200_SCAN_RELOAD_HINTS
#define B200_SCAN_RELOAD_HINTS 0
#endif
#ifndef B200_SCAN_STORE_EVICT_FIRST
#define B200_SCAN_STORE_EVICT_FIRST 0
#endif
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
    unsigned        backoff_ns; // pause between unsuccessful look-back polls
    unsigned        incl;       // 0: exclusive; 0xffffffff: inclusive (every element also receives itself)
    int            *count_out;  // compaction mode of the TMA kernel: number of kept elements
    const unsigned *tile_prefix; // nullable: exclusive prefix of every tile, produced by the kernel in front (sharded
                                 // scan, reduce.cu: scan_tile_sums_kernel) -- no look-back, no tickets
};

__device__ __forceinline__ unsigned long long make_desc(unsigned epoch, unsigned state, unsigned value)
{
    return ((unsigned long long)((epoch << 2) | state) << 32) | value;
}

#ifdef B200_SCAN_TRACE
// Debug build only (cuda_samples_b200/build.py --trace): per-tile timestamps of the pipeline events.
__device__ unsigned long long *g_scan_trace = nullptr; // [num_tiles][8]
__device__ __forceinline__ unsigned long long trace_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define SCAN_TRACE(tile, slot) \
    do { if (g_scan_trace) g_scan_trace[(size_t)(tile) * 8 + (slot)] = trace_now(); } while (0)
#define SCAN_TRACE_VAL(tile, slot, v) \
    do { if (g_scan_trace) g_scan_trace[(size_t)(tile) * 8 + (slot)] = (v); } while (0)
#else
#define SCAN_TRACE(tile, slot) do { } while (0)
#define SCAN_TRACE_VAL(tile, slot, v) do { } while (0)
#endif

// Warp 0: sum of the aggregates of the tiles before `tile` back to the nearest published prefix.
// One poll inspects kLookWindow * 32 descriptors (all loads issued together, so a poll costs one L2
// round trip).  Measured on B200 (profiles/r01_scan_window_sweep.log): wider is NOT better -- every
// in-flight tile polls the same few L2 lines and a 256-wide window slows all memory traffic down
// (tile loads 5 -> 8 us); 32 is the best flat window.  The TMA kernel uses the two-level scheme below.
template <int kLookWindow>
__device__ __forceinline__ unsigned look_back(const unsigned long long *desc, unsigned tile, unsigned first_tile,
                                              unsigned epoch, unsigned backoff_ns = 0)
{
    const unsigned lane   = lane_id();
    const unsigned want_a = (epoch << 2) | kStA, want_p = (epoch << 2) | kStP;
    unsigned       prefix = 0;
    long long      base   = (long long)tile - 1; // lane l of group j inspects tile base - 32 j - l
    unsigned       polls  = 0;
    while (true) {
        polls++;
        unsigned st[kLookWindow], val[kLookWindow];
#pragma unroll
        for (int j = 0; j < kLookWindow; j++) {
            const long long idx = base - 32 * j - lane;
            st[j]               = want_p; // before the array's first tile: a prefix of 0
            val[j]              = 0;
            if (idx >= (long long)first_tile) {
                unsigned long long d = ld_relaxed_u64(desc + idx);
                st[j]                = (unsigned)(d >> 32);
                val[j]               = (unsigned)d;
            }
        }
        bool done = false;
#pragma unroll
        for (int j = 0; j < kLookWindow; j++) {
            const unsigned pmask = __ballot_sync(0xffffffffu, st[j] == want_p);
            const unsigned rmask = __ballot_sync(0xffffffffu, st[j] == want_p || st[j] == want_a);
            if (pmask) {
                const unsigned firstp = __ffs(pmask) - 1;
                const unsigned need   = (1u << firstp) - 1u;
                if ((rmask & need) == need) {
                    prefix += __reduce_add_sync(0xffffffffu, lane <= firstp ? val[j] : 0u);
                    done = true;
                }
                break; // either finished, or an aggregate before the prefix is still missing: poll again
            }
            if (rmask != 0xffffffffu)
                break; // someone in this group has not published yet: poll again from here
            prefix += __reduce_add_sync(0xffffffffu, val[j]);
            base -= 32;
        }
        if (done) {
            if (lane == 0)
                SCAN_TRACE_VAL(tile, 4, polls);
            return prefix;
        }
        if (backoff_ns)
            __nanosleep(backoff_ns);
    }
}

// kLarge: arrays of at least one tile (lane_seg=32, rows_seg=4, warps_seg=16 known at compile time)
template <bool kLarge>
__global__ void __launch_bounds__(kScanThreads, 3)
    scan_kernel(unsigned *dst, const unsigned *src, const ScanParams p, // may alias (in place)
                unsigned *__restrict__ state, unsigned long long *__restrict__ desc)
{
    __shared__ unsigned s_tile, s_epoch, s_prefix;
    __shared__ unsigned s_warp_tot[kScanWarps];

    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // Only arrays spanning several tiles need ordered tickets and the look-back descriptors.
    const bool chained = kLarge && p.tpa > 1 && !p.tile_prefix;
    unsigned   tile = blockIdx.x, epoch = 0;
    if (chained) {
        if (threadIdx.x == 0) {
            // one 64-bit atomic returns {epoch (high) | ticket (low)}; every CTA draws exactly once, so
            // the CTA that draws the last ticket can re-arm {epoch+1, 0} for the next launch.
            const unsigned long long drawn = atomicAdd(reinterpret_cast<unsigned long long *>(state), 1ull);
            const unsigned           t = (unsigned)drawn, ep = (unsigned)(drawn >> 32);
            if (t == p.num_tiles - 1)
                *reinterpret_cast<volatile unsigned long long *>(state) =
                    (unsigned long long)((ep + 1u) & 0x3fffffffu) << 32;
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
    if (kLarge && p.tile_prefix) {
        pdl_wait(); // the tile is loaded and scanned; its prefix is the first thing needed from the kernel in front
        prefix = ld_relaxed_u32(p.tile_prefix + tile);
    }
    if (chained) {
        if (warp == 0) {
            if (tile == first_tile) {
                if (lane == 0)
                    st_relaxed_u64(desc + tile, make_desc(epoch, kStP, prefix + aggregate));
            }
            else {
                if (lane == 0)
                    st_relaxed_u64(desc + tile, make_desc(epoch, kStA, aggregate));
                prefix = look_back<1>(desc, tile, first_tile, epoch); // the first tile's prefix already holds the carry
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
        o.x += v[k].x & p.incl; // inclusive scan (shfl_scan sample): element i also receives src[i]
        o.y += v[k].y & p.incl;
        o.z += v[k].z & p.incl;
        o.w += v[k].w & p.incl;
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


// ---------------------------------------------------------------------------------------------------
// Two-level look-back (used by the TMA kernel).  Polling a flat descriptor array has two costs on this
// part: the front of published prefixes advances by at most one window per L2 round trip, and a wide
// window makes every in-flight tile hammer the same few L2 lines, which measurably slows ALL traffic
// (profiles/r01_scan_trace*.log: 256-wide window -> tile loads take 8 us, 32-wide -> 5 us).
// So tiles are grouped in chunks of 32: level 1 = the <= 31 earlier tiles of the own chunk (one 256-byte
// line, shared by 32 tiles only), level 2 = one descriptor per chunk {chunk total | prefix through the
// chunk}, published by the chunk's last tile.  A hop on level 2 covers 32 chunks = 1024 tiles.
// ---------------------------------------------------------------------------------------------------
// One 32-wide window ending at index `last` (lane l inspects last - l, indices < first count as a
// published prefix of 0).  Returns true when the window is decided: *found_p tells whether a prefix was
// reached, *sum is the total of the aggregates after it (+ the prefix).  Returns false: poll again.
__device__ __forceinline__ bool window32(const unsigned long long *desc, long long last, long long first,
                                         unsigned want_a, unsigned want_p, bool *found_p, unsigned *sum)
{
    const unsigned  lane = lane_id();
    const long long idx  = last - lane;
    unsigned        st = want_p, val = 0;
    if (idx >= first) {
        unsigned long long d = ld_relaxed_u64(desc + idx);
        st                   = (unsigned)(d >> 32);
        val                  = (unsigned)d;
    }
    const unsigned pmask = __ballot_sync(0xffffffffu, st == want_p);
    const unsigned rmask = __ballot_sync(0xffffffffu, st == want_p || st == want_a);
    // lanes below `first` only exist to terminate the walk: a real prefix always comes first
    const unsigned firstp = __ffs(pmask) - 1; // pmask != 0 whenever the window reaches `first`
    if (pmask) {
        const unsigned need = (1u << firstp) - 1u;
        if ((rmask & need) != need)
            return false;
        *sum     = __reduce_add_sync(0xffffffffu, lane <= firstp ? val : 0u);
        *found_p = (last - firstp) >= first; // a lane below `first` is the virtual "0" prefix
        return true;
    }
    if (rmask != 0xffffffffu)
        return false;
    *sum     = __reduce_add_sync(0xffffffffu, val);
    *found_p = false;
    return true;
}

// ---------------------------------------------------------------------------------------------------
// Chained scan, persistent + TMA-staged (the hot case: arrays spanning several tiles, 16-byte aligned).
// One CTA per SM, kTmaStages tiles (32 or 40 KiB) in flight in shared memory, four warp roles that only
// meet through mbarriers (no CTA-wide barrier anywhere in the steady state):
//   producer  (1 lane)   draws {epoch|ticket} with ONE 64-bit atomicAdd, streams the tile in with
//                        cp.async.bulk (1-D TMA, SASS UBLKCP) and, once the scan warps have overwritten
//                        the stage with results, streams it out with cp.async.bulk.global.shared;
//   scan warps (16)      phase A on tile k+kAhead: per-warp total -> shared; the warp that completes the
//                        tile's total publishes the look-back aggregate at once, long before the tile is
//                        scanned.  Phase B on tile k: warp-striped scan in shared memory, seeded with the
//                        tile prefix, conflict-free 16-byte shared accesses;
//   look-back warps (2)  alternate over the CTA's tiles: wait for the tile's total, two-level look-back
//                        (own chunk of 32 tiles, then chunk descriptors), publish the inclusive prefix, hand the exclusive prefix to
//                        the scan warps.  Look-back latency is off the scan warps' critical path.
// Every CTA draws exactly one out-of-range ticket, so the draw numbered tiles+grid-1 is the last one and
// its owner re-arms {epoch+1, ticket 0} for the next launch.
// ---------------------------------------------------------------------------------------------------
// Tile geometry is a template parameter <kTmaRows 16-byte vectors per thread, kTmaStages, kTmaAhead>; measured for
// one 2^28-element array (profiles/r01_scan_geometry_sweep.log): 24 KiB x 9 stages 520 us, 32 KiB x 6 434 us,
// 40 KiB x 5 (phase A 3 ahead) 395 us, 48 KiB x 4 413 us, 64 KiB x 3 477 us.  Short arrays (a few tiles each)
// prefer 32 KiB tiles that divide them evenly (16384 arrays of 16384: 344 us vs 396 us).
#ifndef B200_COMPACT_INPLACE
#define B200_COMPACT_INPLACE 0 // compaction: the warp's rows of the stage double as its packing slab (see phase B);
                               // measured equal (profiles/r02x_compact_inplace.log: 399 vs 401 us at 80 % kept, 357 vs
                               // 362 at 50 %): the pipeline's tile period is set by stage residency, not by this loop
#endif
#ifndef B200_SCAN_EAGER_B
#define B200_SCAN_EAGER_B 0 // measured: 410 vs 396 us for one 2^28 array (phase A falls behind, look-backs get longer)
#endif
constexpr int      kTmaLookWarps = 2;
constexpr int      kTmaThreads   = kScanThreads + 32 * (1 + kTmaLookWarps); // 16 scan + 1 producer + 2 look-back
constexpr unsigned kTileEnd      = 0xffffffffu;

// kRing: slots of per-tile bookkeeping.  kRing == kTmaStages: a tile owns its stage from landing to the scan (the tile
// waits for its prefix in shared memory).  kRing > kTmaStages ("reload"): the stage is handed back as soon as phase A
// has summed the tile; phase B reads the tile a second time with ordinary 16-byte loads -- an L2 hit, the first read
// was a few microseconds ago -- so a stage is busy for the landing time only and the look-ahead is bounded by the
// bookkeeping ring, not by shared memory.
template <int kTmaRows, int kTmaStages, int kRing = kTmaStages>
struct ScanTmaSmem {
    static constexpr unsigned kTmaTile = kScanThreads * kTmaRows * 4;
    unsigned buf[kTmaStages][kTmaTile];     // tiles in flight / being scanned
    uint64_t full[kTmaStages];              // producer  -> all     : tile landed (tx bytes) / END marker
    uint64_t scanned[kTmaStages];           // scan warps -> producer : the stage's tile has been read
    uint64_t agg_ready[kRing];              // scan warps -> look-back: the 16 warp totals are in
    uint64_t prefix_ready[kRing];           // look-back -> scan warps: exclusive tile prefix is in
    uint64_t done[kRing];                   // scan warps -> producer : the slot's tile is finished (reload only)
    unsigned tile[kRing], count[kRing], epoch[kRing];
    unsigned wtot[kRing][kScanWarps];       // per-warp totals of the tile
    unsigned agg[kRing], agg_cnt[kRing];
    unsigned prefix[kRing];
    unsigned end_index;                     // CTA-local index of the END marker (look-back warps)
};

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// kCompact: the same pipeline as a stable stream compaction (compact.cu): the scanned quantity is the flag
// (int)x > 0, and instead of the prefixes the kept elements themselves are stored at their prefix positions.
template <int kTmaRows, int kTmaStages, int kTmaAhead, bool kCompact = false, int kRing = kTmaStages>
__global__ void __launch_bounds__(kTmaThreads, 1)
    scan_tma_kernel(unsigned *dst, const unsigned *src, const ScanParams p, unsigned long long *state,
                    unsigned long long *desc)
{
    // the end marker is passed on through agg_ready of the stages of indices a, a+1: their previous phases must
    // have been consumed by then, which holds when the ring is at least two stages longer than the look-ahead
    static_assert(kRing >= kTmaAhead + 2, "bookkeeping ring too short for the phase-A look-ahead");
    constexpr bool kReload = kRing != kTmaStages;
    static_assert(!kReload || !kCompact, "the compaction keeps its tiles in shared memory");
    typedef ScanTmaSmem<kTmaRows, kTmaStages, kRing> Smem;
    constexpr unsigned kTmaTile = Smem::kTmaTile;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem          &S    = *reinterpret_cast<Smem *>(smem_raw);
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kTmaStages; s++) {
            mbar_init(&S.full[s], 1);
            mbar_init(&S.scanned[s], kScanWarps);
        }
        for (int m = 0; m < kRing; m++) {
            mbar_init(&S.agg_ready[m], kScanWarps);
            mbar_init(&S.prefix_ready[m], 1);
            mbar_init(&S.done[m], kScanWarps);
            S.agg[m]     = 0;
            S.agg_cnt[m] = 0;
        }
        S.end_index = 0xffffffffu;
        fence_mbar_init();
    }
    __syncthreads();
    // The carry (seed of element 0: the sum of the lower shards of a sharded scan) is produced by the kernel in front
    // of this one.  With programmatic dependent launch this kernel is already loading tiles and publishing aggregates
    // while that kernel still exchanges shard totals with its peers; the carry is fetched -- behind
    // griddepcontrol.wait -- only where a first tile's prefix is formed.  Nothing is written to dst before that: every
    // tile's prefix descends from a first tile's.
    auto carry_now = [&]() -> unsigned {
        if (!p.carry)
            return 0u;
        pdl_wait();
        return ld_relaxed_u32(p.carry);
    };
    unsigned long long *const cdesc = desc + p.num_tiles; // chunk-level descriptors follow the tile-level ones

    if (warp == kScanWarps) {
        // ================================ producer (one lane) ==========================================
        if (lane != 0)
            return;
        // Only the LOADS go through the bulk-copy engine.  Measured on B200: one SM's engine moves about
        // 37 GB/s of 1-D bulk traffic, so 32 KiB in + 32 KiB out per tile caps the whole scan near 5.5 TB/s
        // (and an extra L2 prefetch per tile lowers it to 4.3 TB/s); the results therefore leave through
        // ordinary 16-byte stores from the scan warps' registers, which also frees the stage as soon as the
        // tile has been read.
        unsigned k = 0;
        while (true) {
            const unsigned           s = k % kTmaStages, m = k % kRing;
            const unsigned long long drawn  = atomicAdd(state, 1ull);
            const unsigned           ticket = (unsigned)drawn, epoch = (unsigned)(drawn >> 32);
            if (k >= (unsigned)kTmaStages) // the scan warps have pulled the stage's previous tile into registers
                mbar_wait(&S.scanned[s], (k / kTmaStages - 1) & 1);
            if (kReload && k >= (unsigned)kRing) // ... and the slot's previous tile has been scanned and stored
                mbar_wait(&S.done[m], (k / kRing - 1) & 1);
            if (ticket >= p.num_tiles) {
                if (ticket == p.num_tiles + gridDim.x - 1) // the very last draw of this launch
                    *reinterpret_cast<volatile unsigned long long *>(state) =
                        (unsigned long long)((epoch + 1u) & 0x3fffffffu) << 32;
                S.end_index = k;
                S.tile[m]   = kTileEnd;
                mbar_arrive(&S.full[s]);
                // (the scan warps pass the end on to the look-back warps through agg_ready)
                break;
            }
            const unsigned a = ticket / p.tpa, t = ticket - a * p.tpa;
            const size_t   off    = (size_t)a * p.L + (size_t)t * kTmaTile;
            const size_t   remain = p.L - (size_t)t * kTmaTile;
            const unsigned count  = remain < kTmaTile ? (unsigned)remain : kTmaTile;
            S.tile[m]  = ticket;
            S.count[m] = count;
            S.epoch[m] = epoch;
            SCAN_TRACE(ticket, 0);
            mbar_expect_tx(&S.full[s], count * 4u);
#if B200_SCAN_LOAD_EVICT_FIRST
            bulk_g2s_hint(S.buf[s], src + off, count * 4u, &S.full[s], kL2EvictFirst);
#elif B200_SCAN_RELOAD_HINTS
            if (kReload) // the tile is read again in a few microseconds: keep it
                bulk_g2s_hint(S.buf[s], src + off, count * 4u, &S.full[s], kL2EvictLast);
            else
                bulk_g2s(S.buf[s], src + off, count * 4u, &S.full[s]);
#else
            bulk_g2s(S.buf[s], src + off, count * 4u, &S.full[s]);
#endif
            k++;
        }
        return;
    }

    if (warp > kScanWarps) {
        // ================================ look-back warps ==============================================
        for (unsigned k = warp - kScanWarps - 1;; k += kTmaLookWarps) {
            const unsigned s = k % kRing, par = (k / kRing) & 1; // (bookkeeping slot)
            mbar_wait(&S.agg_ready[s], par); // all 16 warp totals are in -- or the scan warps reached the end
            if (k >= *reinterpret_cast<volatile unsigned *>(&S.end_index))
                break;
            const unsigned tile  = S.tile[s];
            const unsigned epoch = S.epoch[s], first_tile = (tile / p.tpa) * p.tpa;
            if (lane == 0)
                SCAN_TRACE(tile, 2);
            const unsigned aggregate = S.agg[s];
            unsigned       prefix    = 0;
            if (tile == first_tile)
                prefix = carry_now();
            if (tile != first_tile) { // (a first tile's inclusive prefix was published by the scan warps)
                const unsigned want_a = (epoch << 2) | kStA, want_p = (epoch << 2) | kStP;
                const unsigned t_in   = tile - first_tile, chunk_in = t_in >> 5;
                const unsigned cpa    = (p.tpa + 31u) >> 5; // chunks per array
                const unsigned chunk  = (tile / p.tpa) * cpa + chunk_in;
                const bool     closes = (t_in & 31u) == 31u || t_in == p.tpa - 1; // last tile of its chunk
                unsigned       sum1 = 0, polls = 0;
                bool           found = false;
                // level 1: earlier tiles of the own chunk
                while (!window32(desc, (long long)tile - 1, (long long)first_tile + (chunk_in << 5), want_a, want_p,
                                 &found, &sum1)) {
                    polls++;
                    if (p.backoff_ns)
                        __nanosleep(p.backoff_ns);
                }
                if (found) {
                    prefix = sum1; // reached a tile that already knows its inclusive prefix (carry included)
                }
                else {
                    if (closes && lane == 0) // the chunk total is known before the chunk's prefix is
                        st_relaxed_u64(cdesc + chunk, make_desc(epoch, kStA, sum1 + aggregate));
                    // level 2: earlier chunks of the same array (chunk 0 always publishes a prefix)
                    long long      last   = (long long)chunk - 1;
                    const long long firstc = (long long)(tile / p.tpa) * cpa;
                    unsigned       sum2   = 0;
                    while (true) {
                        unsigned part = 0;
                        bool     fp   = false;
                        if (!window32(cdesc, last, firstc, want_a, want_p, &fp, &part)) {
                            polls++;
                            if (p.backoff_ns)
                                __nanosleep(p.backoff_ns);
                            continue;
                        }
                        sum2 += part;
                        if (fp)
                            break;
                        if (last - 31 < firstc) { // walked past the array's first chunk (cannot happen: chunk 0
                            sum2 += carry_now();  // always publishes a prefix) -- the seed is the carry
                            break;
                        }
                        last -= 32;
                    }
                    prefix = sum2 + sum1;
                }
                if (lane == 0) {
                    SCAN_TRACE_VAL(tile, 4, (unsigned long long)(polls + 1) | ((unsigned long long)blockIdx.x << 32));
                    st_relaxed_u64(desc + tile, make_desc(epoch, kStP, prefix + aggregate));
                    if (closes)
                        st_relaxed_u64(cdesc + chunk, make_desc(epoch, kStP, prefix + aggregate));
                }
            }
            __syncwarp();
            if (lane == 0) {
                SCAN_TRACE(tile, 3);
                S.prefix[s]  = prefix;
                S.agg[s]     = 0;
                S.agg_cnt[s] = 0;
                mbar_arrive(&S.prefix_ready[s]);
            }
        }
        return;
    }

    // ==================================== scan warps (16) ==============================================
    unsigned a = 0, end_index = 0xffffffffu;
    for (unsigned b = 0;; b++) {
        uint4 v[kTmaRows];
        // reload: the second read of tile b is issued here (its phase A ran kTmaAhead tiles ago, the lines are in L2)
        // and lands under phase A of tile b + kTmaAhead
        auto reload_tile = [&]() {
            const unsigned mb = b % kRing, tile_b = S.tile[mb], count = S.count[mb], ab = tile_b / p.tpa;
            const unsigned *in = src + (size_t)ab * p.L + (size_t)(tile_b - ab * p.tpa) * kTmaTile;
#pragma unroll
            for (int r = 0; r < kTmaRows; r++) {
                const unsigned e = warp * (kTmaRows * 128) + r * 128 + lane * 4;
#if B200_SCAN_RELOAD_HINTS
                v[r] = e < count ? ld_once_hint_u4(in + e, kL2EvictFirst) : make_uint4(0, 0, 0, 0); // last use
#else
                v[r] = e < count ? ld_once_u4(in + e) : make_uint4(0, 0, 0, 0);
#endif
            }
        };
        const bool early = kReload && b < a;
        if (early)
            reload_tile();
        // ---- phase A: warp totals of tiles up to kTmaAhead ahead; the completing warp publishes -------
        while (end_index == 0xffffffffu && a <= b + kTmaAhead) {
            const unsigned s = a % kTmaStages, m = a % kRing;
#if B200_SCAN_EAGER_B
            if (a > b) {
                // look-ahead is opportunistic: if the prefix of tile b arrives before tile a has landed, scan tile b
                // first (its stage goes back to the producer sooner) and come back for tile a afterwards
                bool landed, ready = false;
                do {
                    landed = __any_sync(0xffffffffu, mbar_test(&S.full[s], (a / kTmaStages) & 1));
                    if (!landed)
                        ready = __any_sync(0xffffffffu,
                                           mbar_test(&S.prefix_ready[b % kTmaStages], (b / kTmaStages) & 1));
                } while (!landed && !ready);
                if (!landed)
                    break;
            }
#endif
            mbar_wait(&S.full[s], (a / kTmaStages) & 1);
            const unsigned tile = S.tile[m];
            if (tile == kTileEnd) {
                end_index = a;
                // tell the look-back warps: the owners of indices a .. a+kTmaLookWarps-1 wait on agg_ready of those
                // stages.  (Their previous phases completed and were consumed long ago: phase A is at most
                // kTmaAhead tiles ahead of the prefixes the look-back warps have already delivered.)
                if (lane == 0)
                    for (unsigned jj = 0; jj < (unsigned)kTmaLookWarps; jj++)
                        mbar_arrive(&S.agg_ready[(a + jj) % kRing]);
                break;
            }
            const unsigned count = S.count[m];
            unsigned       sum   = 0;
            if (threadIdx.x == 0)
                SCAN_TRACE(tile, 1);
#pragma unroll
            for (int r = 0; r < kTmaRows; r++) {
                const unsigned e = warp * (kTmaRows * 128) + r * 128 + lane * 4;
                uint4          v = *reinterpret_cast<const uint4 *>(&S.buf[s][e]);
                if (e < count) { // ragged last tile of an array (count is a multiple of 4 here)
                    if (kCompact)
                        sum += ((int)v.x > 0) + ((int)v.y > 0) + ((int)v.z > 0) + ((int)v.w > 0);
                    else
                        sum += (v.x + v.y) + (v.z + v.w);
                }
            }
            sum = __reduce_add_sync(0xffffffffu, sum);
            if (lane == 0) {
                if (kReload) // every lane's shared loads have been consumed by the reduction: the stage can be refilled
                    mbar_arrive(&S.scanned[s]);
                S.wtot[m][warp] = sum;
                atomicAdd(&S.agg[m], sum);
                __threadfence_block();
                if (atomicAdd(&S.agg_cnt[m], 1u) == kScanWarps - 1) {
                    // all 16 warp totals are in: publish for the tiles behind us right away
                    __threadfence_block();
                    const unsigned agg        = *reinterpret_cast<volatile unsigned *>(&S.agg[m]);
                    const unsigned first_tile = (tile / p.tpa) * p.tpa;
                    st_relaxed_u64(desc + tile, tile == first_tile ? make_desc(S.epoch[m], kStP, carry_now() + agg)
                                                                   : make_desc(S.epoch[m], kStA, agg));
                }
                mbar_arrive(&S.agg_ready[m]);
            }
            a++;
        }
        if (b == end_index)
            break;
        // ---- phase B: scan tile b (from its stage, or from the registers the reload filled) ---------------------
        const unsigned s = b % kTmaStages, m = b % kRing;
        if (kReload && !early)
            reload_tile();
        if (threadIdx.x == 0)
            SCAN_TRACE(S.tile[m], 7);
        mbar_wait(&S.prefix_ready[m], (b / kRing) & 1);
        const unsigned count = S.count[m];
        if (threadIdx.x == 0)
            SCAN_TRACE(S.tile[m], 5);
        unsigned       ebase[kTmaRows], tot[kTmaRows], inc[kTmaRows];
#pragma unroll
        for (int r = 0; r < kTmaRows; r++) {
            ebase[r] = warp * (kTmaRows * 128) + r * 128 + lane * 4;
            if (!kReload) {
                v[r] = *reinterpret_cast<const uint4 *>(&S.buf[s][ebase[r]]);
                if (ebase[r] >= count)
                    v[r] = make_uint4(0, 0, 0, 0);
            }
            if (kCompact)
                tot[r] = ((int)v[r].x > 0) + ((int)v[r].y > 0) + ((int)v[r].z > 0) + ((int)v[r].w > 0);
            else
                tot[r] = (v[r].x + v[r].y) + (v[r].z + v[r].w);
            inc[r] = tot[r];
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
#pragma unroll
            for (int r = 0; r < kTmaRows; r++) {
                unsigned t = __shfl_up_sync(0xffffffffu, inc[r], d);
                if (lane >= d)
                    inc[r] += t;
            }
        unsigned x = (lane < kScanWarps) ? S.wtot[m][lane] : 0u, xi = x;
#pragma unroll
        for (int d = 1; d < kScanWarps; d <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, xi, d);
            if (lane >= d)
                xi += t;
        }
        unsigned run = S.prefix[m] + __shfl_sync(0xffffffffu, xi - x, warp);
        const unsigned tile_b = S.tile[m];
        // everything this warp needs from the stage / the slot is in registers now: hand it back to the producer
        __syncwarp();
        if (threadIdx.x == 0)
            SCAN_TRACE(tile_b, 6);
#if B200_COMPACT_INPLACE
        if (kCompact) {
            // The warp's own rows of the stage (all in registers now) are the packing slab for the WHOLE tile slice:
            // one pass of shared stores, one __syncwarp, then the kept elements leave as full 128-byte store
            // instructions -- two warp barriers per tile instead of two per row, and the loads of all the output
            // instructions in flight together.  The stage goes back to the producer when the slab has been read.
            unsigned *slab    = &S.buf[s][warp * (kTmaRows * 128)];
            unsigned  rowbase = 0;
#pragma unroll
            for (int r = 0; r < kTmaRows; r++) {
                unsigned pos = rowbase + inc[r] - tot[r];
                if ((int)v[r].x > 0)
                    slab[pos++] = v[r].x;
                if ((int)v[r].y > 0)
                    slab[pos++] = v[r].y;
                if ((int)v[r].z > 0)
                    slab[pos++] = v[r].z;
                if ((int)v[r].w > 0)
                    slab[pos++] = v[r].w;
                rowbase += __shfl_sync(0xffffffffu, inc[r], 31);
            }
            __syncwarp();
            unsigned t[kTmaRows * 4];
#pragma unroll
            for (int q = 0; q < kTmaRows * 4; q++)
                t[q] = slab[q * 32 + lane];
#pragma unroll
            for (int q = 0; q < kTmaRows * 4; q++)
                if (q * 32 + lane < rowbase)
                    dst[run + q * 32 + lane] = t[q];
            __syncwarp(); // (every lane's slab loads have been consumed by its stores)
            if (lane == 0)
                mbar_arrive(&S.scanned[s]);
            if (tile_b == p.num_tiles - 1 && warp == kScanWarps - 1 && lane == 0)
                *p.count_out = (int)(run + rowbase); // prefix + everything kept in the last tile
            continue;
        }
#endif
        if (lane == 0)
            mbar_arrive(kReload ? &S.done[m] : &S.scanned[s]);
        const unsigned ab = tile_b / p.tpa;
        unsigned      *out = dst + (size_t)ab * p.L + (size_t)(tile_b - ab * p.tpa) * kTmaTile;
#pragma unroll
        for (int r = 0; r < kTmaRows; r++) {
            const unsigned e0 = run + (inc[r] - tot[r]);
            run += __shfl_sync(0xffffffffu, inc[r], 31);
            if (kCompact) { // zeros stand in for elements beyond `count`: never kept
                // the row's kept elements are packed in the warp's 512-byte slab first, so that they leave as full
                // 128-byte store instructions (lane-private stores straight from the registers touch 4x the sectors)
                unsigned *slab = reinterpret_cast<unsigned *>(smem_raw + sizeof(Smem)) + (warp & (kScanWarps - 1)) * 128;
                unsigned  pos  = inc[r] - tot[r];
                if ((int)v[r].x > 0)
                    slab[pos++] = v[r].x;
                if ((int)v[r].y > 0)
                    slab[pos++] = v[r].y;
                if ((int)v[r].z > 0)
                    slab[pos++] = v[r].z;
                if ((int)v[r].w > 0)
                    slab[pos++] = v[r].w;
                const unsigned kept  = __shfl_sync(0xffffffffu, inc[r], 31);
                const unsigned start = __shfl_sync(0xffffffffu, e0, 0);
                __syncwarp();
#pragma unroll
                for (int q = 0; q < 4; q++)
                    if (q * 32 + lane < kept)
                        dst[start + q * 32 + lane] = slab[q * 32 + lane];
                __syncwarp();
                continue;
            }
            uint4 o;
            o.x = e0;
            o.y = e0 + v[r].x;
            o.z = o.y + v[r].y;
            o.w = o.z + v[r].z;
            o.x += v[r].x & p.incl;
            o.y += v[r].y & p.incl;
            o.z += v[r].z & p.incl;
            o.w += v[r].w & p.incl;
            if (ebase[r] < count) {
#if B200_SCAN_STORE_EVICT_FIRST || B200_SCAN_RELOAD_HINTS
                st_stream_hint_u4(out + ebase[r], o, kL2EvictFirst);
#else
                st_stream_u4(out + ebase[r], o);
#endif
            }
        }
        if (kCompact && tile_b == p.num_tiles - 1 && warp == kScanWarps - 1 && lane == 0)
            *p.count_out = (int)run; // prefix + everything kept in the last tile
    }
}

static bool is_pow2(size_t x) { return x && !(x & (x - 1)); }
static bool g_scan_use_tma = true;
static int  g_scan_backoff_ns = 100;

} // namespace b200

using namespace b200;

#ifndef B200_SCAN_RELOAD_AHEAD
#define B200_SCAN_RELOAD_AHEAD 6
#endif
// Reload pipeline (stage freed after phase A, phase B re-reads the tile with 16-byte loads): measured and NOT used
// (profiles/r02u_scan_reload_*.log, one 2^28 array): 391 us for tiles that wait in shared memory against 467 / 427 /
// 484 us with 2 / 3 / 6 tiles of look-ahead; ncu at 3: the second read hits L2 for 73 % of its sectors only (DRAM
// reads 1.27 GB instead of 1.07 GB) and eviction hints change nothing.  Kept behind the knob for the record.
static bool g_scan_reload = false;
extern "C" int b200_tune_scan_reload(int on)
{
    g_scan_reload = on != 0;
    return 0;
}
static bool g_scan_pdl = true; // a scan with a carry overlaps the tail of the kernel that produces the carry

template <typename K>
static int launch_tma(K kernel, unsigned grid, size_t smem, cudaStream_t stream, bool pdl, uint32_t *d_dst,
                      const uint32_t *d_src, ScanParams p, unsigned long long *st64, unsigned long long *desc)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(grid);
    cfg.blockDim           = dim3(kTmaThreads);
    cfg.dynamicSmemBytes   = smem;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    // measured (profiles/r02n_scan_shard_l2_pdl.log): the overlap pays for shards up to ~128 MiB (80.1 -> 76.5 us);
    // behind a longer total pass the early tiles only compete with it for HBM (2^27 elements: 286 -> 304 us)
    int dev = -1;
    B200_TRY(cudaGetDevice(&dev));
    // (tiles are bulk-loaded from d_src before griddepcontrol.wait: see the programmatic-launch guard in internal.h)
    cfg.numAttrs = (pdl && g_scan_pdl && (size_t)p.num_tiles * 8192u <= (40u << 20) &&
                    pdl_early_read_ok(dev, d_src, p.total * sizeof(uint32_t)))
                       ? 1
                       : 0;
    B200_TRY(cudaLaunchKernelEx(&cfg, kernel, d_dst, d_src, p, st64, desc));
    return 0;
}

static int scan_impl(uint32_t *d_dst, const uint32_t *d_src, size_t batch, size_t L, const uint32_t *d_carry,
                     bool inclusive, b200_stream_t stream_)
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
    p.backoff_ns = (unsigned)g_scan_backoff_ns;
    p.incl       = inclusive ? 0xffffffffu : 0u;
    p.count_out  = nullptr;
    p.tile_prefix = nullptr;
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
    // descriptors: one per tile, then one per chunk of 32 tiles of each array
    size_t ndesc = (large && p.tpa > 1) ? (size_t)p.num_tiles + batch * (((size_t)p.tpa + 31) / 32) : 0;
    B200_TRY((cudaError_t)ensure(c->scan, 16 + 8 * ndesc, stream));
    unsigned           *state = reinterpret_cast<unsigned *>(c->scan.ptr);
    unsigned long long *desc  = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(c->scan.ptr) + 16);
    const bool tma_ok = large && L > kScanTile && (L % 4) == 0 && g_scan_use_tma &&
                        ((((uintptr_t)d_dst) | ((uintptr_t)d_src)) & 15) == 0;
    if (tma_ok) {
        // long arrays: 40 KiB tiles x 5 stages; arrays of a few tiles: 32 KiB x 6 (divides powers of two evenly)
        const bool     big  = L >= (1u << 20);
        const unsigned tile = big ? kScanThreads * 5 * 4 : kScanThreads * 4 * 4;
        const size_t   tpa_t = (L + tile - 1) / tile;
        if (tpa_t * batch > 0xfffffff0ull)
            return (int)cudaErrorInvalidValue;
        p.tpa       = (unsigned)tpa_t;
        p.num_tiles = (unsigned)(tpa_t * batch);
        const size_t nd = (size_t)p.num_tiles + batch * ((tpa_t + 31) / 32);
        B200_TRY((cudaError_t)ensure(c->scan, 16 + 8 * nd, stream));
        state = reinterpret_cast<unsigned *>(c->scan.ptr);
        desc  = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(c->scan.ptr) + 16);
        unsigned grid = (unsigned)c->num_sms; // one persistent CTA per SM
        if (grid > p.num_tiles)
            grid = p.num_tiles;
        auto *st64 = reinterpret_cast<unsigned long long *>(state);
        if (big && g_scan_reload) {
            typedef ScanTmaSmem<5, 5, 16> SmemR;
            B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma3, scan_tma_kernel<5, 5, B200_SCAN_RELOAD_AHEAD, false, 16>,
                                              sizeof(SmemR)));
            B200_TRY((cudaError_t)launch_tma(scan_tma_kernel<5, 5, B200_SCAN_RELOAD_AHEAD, false, 16>, grid, sizeof(SmemR),
                                             stream, d_carry != nullptr, d_dst, d_src, p, st64, desc));
        }
        else if (big) {
            B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma0, scan_tma_kernel<5, 5, 3>, sizeof(ScanTmaSmem<5, 5>)));
            B200_TRY((cudaError_t)launch_tma(scan_tma_kernel<5, 5, 3>, grid, sizeof(ScanTmaSmem<5, 5>), stream,
                                             d_carry != nullptr, d_dst, d_src, p, st64, desc));
        }
        else {
            B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma1, scan_tma_kernel<4, 6, 3>, sizeof(ScanTmaSmem<4, 6>)));
            B200_TRY((cudaError_t)launch_tma(scan_tma_kernel<4, 6, 3>, grid, sizeof(ScanTmaSmem<4, 6>), stream,
                                             d_carry != nullptr, d_dst, d_src, p, st64, desc));
        }
    }
    else if (large)
        scan_kernel<true><<<p.num_tiles, kScanThreads, 0, stream>>>(d_dst, d_src, p, state, desc);
    else
        scan_kernel<false><<<p.num_tiles, kScanThreads, 0, stream>>>(d_dst, d_src, p, state, desc);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}

extern "C" int b200_scan_exclusive_async(uint32_t *d_dst, const uint32_t *d_src, size_t batch, size_t L,
                                         const uint32_t *d_carry, b200_stream_t stream)
{
    return scan_impl(d_dst, d_src, batch, L, d_carry, false, stream);
}

// Inclusive +scan (dst[j] = sum of src[0..j], wrapping): what the shfl_scan sample computes with three launches
// (shfl_scan_test x2 + uniform_add, Samples/2_Concepts_and_Techniques/shfl_scan/shfl_scan.cu:55-147, host
// sequence :275-277, gold :157-161) -- here the same single-pass kernels with one extra add per element.
extern "C" int b200_scan_inclusive_async(uint32_t *d_dst, const uint32_t *d_src, size_t batch, size_t L,
                                         const uint32_t *d_carry, b200_stream_t stream)
{
    return scan_impl(d_dst, d_src, batch, L, d_carry, true, stream);
}

// Stream compaction on the TMA-staged chained-scan pipeline (called by compact.cu for long, 16-byte aligned
// inputs with n % 4 == 0; everything else takes compact_kernel).  Returns cudaErrorNotSupported when the shape
// does not qualify.
int compact_tma(int32_t *d_dst, int32_t *d_count, const int32_t *d_src, size_t n, cudaStream_t stream)
{
    constexpr unsigned tile = kScanThreads * 5 * 4;
    if (n < 4 * (size_t)tile || (n & 3) || (((uintptr_t)d_src) & 15) || !g_scan_use_tma)
        return (int)cudaErrorNotSupported;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    ScanParams p;
    p.total      = n;
    p.L          = n;
    p.carry      = nullptr;
    p.backoff_ns = (unsigned)g_scan_backoff_ns;
    p.incl       = 0;
    p.count_out  = d_count;
    p.tile_prefix = nullptr;
    p.lane_seg   = 32;
    p.rows_seg   = kScanRows;
    p.warps_seg  = kScanWarps;
    p.packed     = 0;
    const size_t tpa = (n + tile - 1) / tile;
    if (tpa > 0xfffffff0ull)
        return (int)cudaErrorInvalidValue;
    p.tpa       = (unsigned)tpa;
    p.num_tiles = (unsigned)tpa;
    const size_t nd = tpa + (tpa + 31) / 32;
    B200_TRY((cudaError_t)ensure(c->scan, 16 + 8 * nd, stream));
    auto *st64 = reinterpret_cast<unsigned long long *>(c->scan.ptr);
    auto *desc = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(c->scan.ptr) + 16);
    unsigned grid = (unsigned)c->num_sms;
    if (grid > p.num_tiles)
        grid = p.num_tiles;
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma2, scan_tma_kernel<5, 5, 3, true>,
                                      sizeof(ScanTmaSmem<5, 5>) + kScanWarps * 512));
    // + one 512-byte packing slab per scan warp behind the pipeline's shared memory
    scan_tma_kernel<5, 5, 3, true><<<grid, kTmaThreads, sizeof(ScanTmaSmem<5, 5>) + kScanWarps * 512, stream>>>(
        reinterpret_cast<unsigned *>(d_dst), reinterpret_cast<const unsigned *>(d_src), p, st64, desc);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}

// Second pass of the sharded scan (internal.h): every tile's exclusive prefix is known, tiles are independent.
namespace b200 {
int scan_tiled(uint32_t *d_dst, const uint32_t *d_src, size_t n, const uint32_t *d_tile_prefix, cudaStream_t stream)
{
    ScanParams p;
    p.total       = n;
    p.L           = n;
    p.carry       = nullptr;
    p.backoff_ns  = 0;
    p.incl        = 0;
    p.count_out   = nullptr;
    p.tile_prefix = d_tile_prefix;
    p.lane_seg    = 32;
    p.rows_seg    = kScanRows;
    p.warps_seg   = kScanWarps;
    p.packed      = 0;
    const size_t tiles = (n + kScanTile - 1) / kScanTile;
    if (tiles == 0)
        return 0;
    if (tiles > 0x7fffffffull)
        return (int)cudaErrorInvalidValue;
    p.tpa       = (unsigned)tiles;
    p.num_tiles = (unsigned)tiles;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3((unsigned)tiles);
    cfg.blockDim           = dim3(kScanThreads);
    cfg.dynamicSmemBytes   = 0;
    cfg.stream             = stream;
    cudaLaunchAttribute at[1];
    at[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                        = at;
    int dev = -1;
    B200_TRY(cudaGetDevice(&dev));
    cfg.numAttrs = (g_scan_pdl && pdl_early_read_ok(dev, d_src, n * sizeof(uint32_t))) ? 1 : 0;
    unsigned           *no_state = nullptr;
    unsigned long long *no_desc  = nullptr;
    B200_TRY(cudaLaunchKernelEx(&cfg, scan_kernel<true>, d_dst, d_src, p, no_state, no_desc));
    count_launch();
    return 0;
}
} // namespace b200

extern "C" int b200_tune_scan_pdl(int on)
{
    g_scan_pdl = on != 0;
    return 0;
}
// tuning knob for tests / bench (not part of the stable ABI): 0 = register-staged chained scan only
extern "C" int b200_tune_scan_tma(int enable)
{
    g_scan_use_tma = enable != 0;
    return 0;
}
extern "C" int b200_tune_scan_backoff(int ns)
{
    g_scan_backoff_ns = ns;
    return 0;
}

extern "C" int b200_scan_total_async(uint32_t *d_total, const uint32_t *d_src, size_t n, b200_stream_t stream)
{
    // wrapping u32 sum == the int reduction
    return b200_reduce_i32_async(reinterpret_cast<const int32_t *>(d_src), n, reinterpret_cast<int32_t *>(d_total), 1,
                                 stream);
}

// Pre-size the descriptor scratch for scans of up to n elements in one array (initScan()).
extern "C" int b200_scan_reserve(size_t n)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const size_t tiles = (n + kScanTile - 1) / kScanTile; // (40 KiB tiles need fewer)
    B200_TRY((cudaError_t)ensure(c->scan, 16 + 8 * (tiles + tiles / 32 + 64), 0));
    B200_TRY((cudaError_t)ensure(c->scan_tiles, 16 + 4 * (tiles + 64), 0)); // tile sums / prefixes of a sharded scan
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Scan for callers that must not allocate.
//
// Some users of this library are not free to call cudaMalloc: a plug-in inside a host application that handed the
// whole device to its own memory pool at start-up (a later cudaMalloc fails or steals from the pool), a service
// running under a fixed memory budget that is audited once and then frozen, a real-time loop in which the implicit
// device-wide synchronisation of cudaMalloc / cudaFree would stall every other stream.  The other scan entry points
// keep their tile descriptors in grow-on-demand cudaMalloc'ed scratch (context.cu: ensure), so such a caller could
// not use them the first time, or the first time with a longer array.  This one keeps the descriptors in a fixed
// array that is part of the library image instead: loading the library on a device costs that device
// sizeof(g_scan_noalloc) bytes (about 64 KiB) until the process exits, and nothing is ever allocated or freed here.
//
// The price is a size limit (kNoAllocMaxTiles tiles of 8192 elements per call) and ONE array per device shared by every
// call: see the comment on the entry point.
// ---------------------------------------------------------------------------------------------------------------
constexpr unsigned kNoAllocMaxTiles = 8192; // == B200_SCAN_NOALLOC_MAX_TILES (b200prim.h): one array of up to 2^26 elements
// [0] {epoch | ticket} of scan_kernel, [1] unused (keeps the descriptors 16-byte aligned like the cudaMalloc'ed
// layout), [2 ..) one 64-bit look-back descriptor per tile.  Zero at load time == epoch 0, no ticket drawn, nothing
// published: exactly what a freshly zero-filled scratch block is for the other entry points.
static __device__ __align__(256) unsigned long long g_scan_noalloc[2 + kNoAllocMaxTiles];

// In host code `g_scan_noalloc` is only a handle the runtime knows the array by, NOT its device address: the address
// is looked up once per device (with that device current) and remembered here.
static std::mutex          g_noalloc_mutex;
static unsigned long long *g_noalloc_addr[64] = {};

static int noalloc_scratch(unsigned long long **out)
{
    int dev = -1;
    B200_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64)
        return (int)cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_addr[dev]) {
        void *p = nullptr;
        B200_TRY(cudaGetSymbolAddress(&p, g_scan_noalloc)); // this device's copy of the array
        if (!p)
            return (int)cudaErrorInvalidSymbol;
        g_noalloc_addr[dev] = static_cast<unsigned long long *>(p);
    }
    *out = g_noalloc_addr[dev];
    return 0;
}

// Same result as b200_scan_exclusive_async (batchSize independent exclusive +scans of arrayLength u32, wrapping,
// optional d_carry seed, in place allowed), computed by the register-staged chained-scan kernel with its descriptors
// in the library's built-in array.  Makes no allocation of any kind, on any call.
//   * Limit: the launch may need at most 8192 tiles of 8192 elements -- batchSize * ceil(arrayLength / 8192) for
//     arrays of 8192 elements or more and for lengths that are not a power of two, ceil(batchSize * arrayLength /
//     8192) for shorter power-of-two lengths.  One array can be 2^26 elements (256 MiB) long.  Anything larger is
//     refused with cudaErrorInvalidValue and a line on stderr that names the limit; nothing is launched.
//   * The first call on each device looks the array's address up (cudaGetSymbolAddress, which may also make the
//     runtime load the library's device code).  Make that first call outside any stream capture.
//   * All calls on a device use the same array and find in it what the previous call left.  Nothing has to be
//     cleared by the caller: the one word that must be zero when a launch starts, the ticket counter, is put back to
//     zero by the kernel itself (the CTA that draws the last ticket re-arms it and advances the epoch that tags the
//     descriptors, so stale descriptors are never mistaken for fresh ones).  That only works when the calls do not
//     overlap on the GPU: two of them running at once -- different streams with no event between them -- draw from the
//     same ticket counter and overwrite each other's descriptors, which gives wrong results or a hang.  Keep them on one
//     stream or order them with events.
extern "C" int b200_scan_exclusive_noalloc_async(uint32_t *d_dst, const uint32_t *d_src, size_t batch, size_t L,
                                                 const uint32_t *d_carry, b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (batch == 0 || L == 0)
        return 0;
    int st = 0;
    if (!ctx(&st)) // sm_100a check only; the context object holds no memory that this path uses
        return st;
    ScanParams p;
    p.total       = batch * L;
    p.L           = L;
    p.carry       = d_carry;
    p.backoff_ns  = 0;
    p.incl        = 0;
    p.count_out   = nullptr;
    p.tile_prefix = nullptr;
    const bool large = L >= kScanTile || !is_pow2(L) || L < 4;
    size_t     tiles;
    if (large) {
        const size_t tpa = (L + kScanTile - 1) / kScanTile;
        // (no overflow: tpa <= 8192 in the product, batch < 2^32 is checked below)
        tiles = (tpa > kNoAllocMaxTiles || batch > 0xffffffffull) ? (size_t)kNoAllocMaxTiles + 1 : tpa * batch;
        p.tpa = (unsigned)(tpa > kNoAllocMaxTiles ? 0 : tpa);
        p.lane_seg       = 32;
        p.rows_seg       = kScanRows;
        p.warps_seg      = kScanWarps;
        p.packed         = 0;
    }
    else {
        tiles       = (p.total + kScanTile - 1) / kScanTile;
        p.tpa       = 1;
        p.lane_seg  = (unsigned)(L / 4 < 32 ? L / 4 : 32);
        p.rows_seg  = (unsigned)(L / 128 < 1 ? 1 : (L / 128 > kScanRows ? kScanRows : L / 128));
        p.warps_seg = (unsigned)(L / 512 < 1 ? 1 : L / 512);
        p.packed    = 1;
    }
    if (tiles > kNoAllocMaxTiles || batch > 0xffffffffull) {
        fprintf(stderr,
                "b200_scan_exclusive_noalloc_async: %zu arrays of %zu elements need more than the %u tiles of 8192 "
                "elements that the built-in scratch holds (one array: at most %zu elements); use "
                "b200_scan_exclusive_async, which allocates\n",
                batch, L, kNoAllocMaxTiles, (size_t)kNoAllocMaxTiles * kScanTile);
        return (int)cudaErrorInvalidValue;
    }
    p.num_tiles = (unsigned)tiles;
    unsigned long long *scratch = nullptr;
    B200_TRY((cudaError_t)noalloc_scratch(&scratch));
    unsigned           *state = reinterpret_cast<unsigned *>(scratch);
    unsigned long long *desc  = scratch + 2;
    if (large)
        scan_kernel<true><<<p.num_tiles, kScanThreads, 0, stream>>>(d_dst, d_src, p, state, desc);
    else
        scan_kernel<false><<<p.num_tiles, kScanThreads, 0, stream>>>(d_dst, d_src, p, state, desc);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}

namespace b200 {
int scan_preload()
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma0, scan_tma_kernel<5, 5, 3>, sizeof(ScanTmaSmem<5, 5>)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma1, scan_tma_kernel<4, 6, 3>, sizeof(ScanTmaSmem<4, 6>)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrScanTma3, scan_tma_kernel<5, 5, B200_SCAN_RELOAD_AHEAD, false, 16>,
                                      sizeof(ScanTmaSmem<5, 5, 16>)));
    cudaFuncAttributes fa;
    B200_TRY(cudaFuncGetAttributes(&fa, scan_kernel<true>));
    B200_TRY(cudaFuncGetAttributes(&fa, scan_kernel<false>));
    return ensure(c->scan, 4096, 0);
}
} // namespace b200

#ifdef B200_SCAN_TRACE
extern "C" int b200_debug_set_scan_trace(void *d_trace)
{
    unsigned long long *p = reinterpret_cast<unsigned long long *>(d_trace);
    B200_TRY(cudaMemcpyToSymbol(g_scan_trace, &p, sizeof(p)));
    return 0;
}
#endif
