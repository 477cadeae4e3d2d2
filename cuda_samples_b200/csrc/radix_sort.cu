// radix_sort.cu -- stable key-value u32 sort (onesweep LSD radix, 8-bit digits) for sm_100a.
//
// Replaces mergeSort() and its four kernels (Samples/0_Introduction/mergeSort/mergeSort.cu:
// mergeSortSharedKernel :108, generateSampleRanksKernel :182, mergeRanksAndIndicesKernel :237,
// mergeElementaryIntervalsKernel :326, host loop :471-537: 1 + 4*log2(N/1024) launches and
// ~16 B/pair of traffic per merge level, N <= 2^22) with
//   1 memset + 1 digit-histogram kernel (reads the keys once for all four digits and leaves the
//   exclusive digit offsets behind) + 4 scatter passes, each a single chained-scan kernel:
//     * tiles of 8192 pairs are taken in ticket order; keys are loaded warp-striped;
//     * ranking is stable by construction: inside a warp, items are ranked in index order with a
//       ballot-built match mask per 8-bit digit (rank = running warp-private digit counter + number
//       of lower lanes holding the same digit), warps are ordered by a per-digit scan over warps;
//     * the tile's 256 digit counts go to a look-back descriptor row (2 status bits + 30-bit
//       value per digit); thread d walks back over earlier tiles for digit d;
//     * keys, then values, are staged through shared memory in digit order so that each digit's
//       run leaves the SM as one contiguous store.
//   Passes ping-pong src -> buf -> dst -> buf -> dst: src is never written, the result lands in dst.
//   sortDir == 0 (descending) sorts on ~key; ties keep source order in both directions.
// The same pass kernel with a splitter-based "digit" is the local partition step of the multi-GPU
// sample sort (b200_partition_kv_async).
// Traffic model: (4 + 16*4) = 68 B per pair; HBM-bound in bytes, issue-bound in practice.
#include "common.cuh"
#include "internal.h"
#include "p2p.cuh"

namespace b200 {

#ifndef B200_SORT_THREADS // tile geometry: overridable for sweeps (profiles/sort_variants.sh)
#define B200_SORT_THREADS 512
#endif
#ifndef B200_SORT_ITEMS
#define B200_SORT_ITEMS 16
#endif
#ifndef B200_SORT_CTAS
#define B200_SORT_CTAS 2
#endif
constexpr int      kSortThreads = B200_SORT_THREADS;
constexpr int      kSortWarps   = kSortThreads / 32;
constexpr int      kSortItems   = B200_SORT_ITEMS;     // pairs per thread
constexpr int      kSortCtasPerSm = B200_SORT_CTAS;
constexpr unsigned kSortTile    = kSortThreads * kSortItems; // pairs per tile (8192)
static_assert(kSortThreads >= 256 && kSortTile <= 65536, "digit owners are threads 0..255; 16-bit slots");
constexpr int      kRadix       = 256;
#ifndef B200_SORT_PACKED
#define B200_SORT_PACKED 1 // 1: (key, value) staged as ONE 64-bit word per pair (one STS.64 from inside the ranking
                           //    loop, one LDS.64 in the store phase); 0: keys and values staged separately
#endif
#ifndef B200_SORT_VALS_EARLY
#define B200_SORT_VALS_EARLY (B200_SORT_PACKED ? B200_SORT_ITEMS : 8) // values whose loads precede the ranking loop
#endif
#ifndef B200_SORT_LB_FIRST
#define B200_SORT_LB_FIRST 0 // 1: the look-back runs BEFORE the ranking loop (prefix published ~5 us earlier: shorter
                             //    chains for every successor, but this tile waits for it up front)
#endif
#ifndef B200_SORT_VALS_BEFORE_OWNERS
#define B200_SORT_VALS_BEFORE_OWNERS 0 // 1: the value loads are issued before the digit owners' phase
#endif
#ifndef B200_SORT_LEADER_HIGH
#define B200_SORT_LEADER_HIGH 0 // 1: the group's HIGHEST lane bumps the counter (bfind, no brev)
#endif
#ifndef B200_SORT_LB_STEPS
#define B200_SORT_LB_STEPS 4 // look-back round trips issued from INSIDE the ranking loop (0: look-back after it)
#endif
constexpr int      kLbSteps   = B200_SORT_LB_STEPS;
constexpr int      kValsEarly = B200_SORT_VALS_EARLY;
static_assert(!B200_SORT_PACKED || kValsEarly == kSortItems, "packed staging needs every value before the ranking loop");
constexpr unsigned kDescA = 1u << 30, kDescP = 2u << 30, kDescValue = (1u << 30) - 1;

#ifdef B200_SCAN_TRACE
// Debug build only: per-tile timestamps of the pass kernel's phases (profiles/sort_trace.py)
__device__ unsigned long long *g_sort_trace = nullptr; // [tiles][8]
__device__ __forceinline__ unsigned long long sort_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define SORT_TRACE(tile, slot) \
    do { if (g_sort_trace && threadIdx.x == 0) g_sort_trace[(size_t)(tile) * 8 + (slot)] = sort_now(); } while (0)
#else
#define SORT_TRACE(tile, slot) do { } while (0)
#endif

// shared memory of the pass kernel (dynamic)
struct SortSmem {
#if B200_SORT_PACKED
    uint2    pair[kSortTile + kP2PMaxRanks]; // (key, value) in digit order (+ room to re-space the runs of a peer pass)
#else
    unsigned buf[kSortTile];              // keys in digit order
    unsigned vbuf[kSortTile];             // values in digit order
#endif
    unsigned wcnt[kSortWarps][kRadix];    // per warp and digit: count, then the warp's running scatter position
    unsigned gofs[kRadix];                // global index of a digit's first slot minus its tile-local start
    unsigned cnt[kRadix];                 // parking space (tile count / tile-local start per digit) while the
    unsigned tstart[kRadix];              //   ranking loop needs every register
    unsigned wsum[kRadix / 32];
    unsigned respace[kP2PMaxRanks];       // peer pass: slots inserted in front of a bucket's run (parity match for bulk copies)
    unsigned splitters[kRadix];
    unsigned tile;
};

// Peer variant of the partition (multi-GPU sample sort): bucket b IS rank b, and its run is stored straight into
// rank b's receive buffer over NVLink instead of into a local array that an all-to-all would copy again.
struct SortPeer {
    unsigned long long k[kP2PMaxRanks], v[kP2PMaxRanks]; // receive buffers of every rank (peer-mapped); v[0] == 0:
                                                         // k[r] holds (key, value) PAIRS (one 8-byte peer store per pair)
    const unsigned    *info;   // [0] n_out, [1] != 0: a receive buffer would overflow -> nothing is sent
    unsigned           bulk;   // pair format only: runs leave through the bulk-copy engine instead of lane stores
};
struct NoPeer {};

template <bool kPeer> struct PeerParam { typedef NoPeer type; };
template <> struct PeerParam<true> { typedef SortPeer type; };

// "digit" of a key: radix byte of (key ^ flip), or the bucket index among ascending splitters
template <bool kBucket> struct DigitOp {
    unsigned        shift, flip, nspl;
    const unsigned *spl; // shared memory
    __device__ __forceinline__ unsigned operator()(unsigned key) const
    {
        if (!kBucket)
            return ((key >> shift) ^ flip) & 255u; // shift + one LOP3
        const unsigned o = key ^ flip;
        // number of splitters <= o  (upper bound), nspl <= 255
        unsigned lo = 0, hi = nspl;
        while (lo < hi) {
            unsigned mid = (lo + hi) >> 1;
            if (spl[mid] <= o)
                lo = mid + 1;
            else
                hi = mid;
        }
        return lo;
    }
};

// lanes of the warp holding the same 8-bit digit (one ballot per bit, registers only)
// per bit: test -> predicate, ballot, predicate -> 0/-1, mask &= ~(vote ^ that): 4 issue slots (the compiler's
// own sequence for the C expression spends 6: it shifts and re-compares to form the predicate)
template <int B> __device__ __forceinline__ void match_bit(unsigned &mask, unsigned d)
{
    asm volatile("{\n\t"
                 ".reg .pred p;\n\t"
                 ".reg .b32 t, v;\n\t"
                 "and.b32 t, %1, %2;\n\t"
                 "setp.ne.u32 p, t, 0;\n\t"
                 "vote.sync.ballot.b32 v, p, 0xffffffff;\n\t"
                 "selp.b32 t, -1, 0, p;\n\t"
                 "lop3.b32 %0, %0, v, t, 0x90;\n\t"
                 "}"
                 : "+r"(mask)
                 : "r"(d), "n"(1 << B));
}
__device__ __forceinline__ unsigned match_digit(unsigned d)
{
    unsigned mask = 0xffffffffu;
    match_bit<0>(mask, d);
    match_bit<1>(mask, d);
    match_bit<2>(mask, d);
    match_bit<3>(mask, d);
    match_bit<4>(mask, d);
    match_bit<5>(mask, d);
    match_bit<6>(mask, d);
    match_bit<7>(mask, d);
    return mask;
}

__device__ __forceinline__ uint4 ld_relaxed_v4(const unsigned *p)
{
    uint4 r;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_relaxed_v4(unsigned *p, uint4 v)
{
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// Last CTA of a histogram kernel: warp p turns histogram p (256 counts) into exclusive offsets in place
// (lane l owns entries [8l, 8l+8)); optionally copies the first nspl+1 raw counts out first.
__device__ __forceinline__ void finish_digit_offsets(unsigned *ghist, int passes, unsigned *counts_out, unsigned nspl)
{
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if ((int)warp < passes) {
        unsigned *g = ghist + warp * kRadix + lane * 8;
        unsigned  c[8], sum = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            c[i] = __ldcg(g + i);
            sum += c[i];
        }
        if (counts_out)
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (lane * 8 + i <= nspl)
                    counts_out[lane * 8 + i] = c[i];
        unsigned inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d)
                inc += t;
        }
        unsigned run = inc - sum;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            g[i] = run;
            run += c[i];
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// Digit histograms of all passes in one read of the keys; the last CTA turns each 256-bin histogram
// into exclusive offsets in place.  ghist: [passes][256], zero on entry.
// ---------------------------------------------------------------------------------------------------
// What the last CTA of the bucket histogram does in the peer variant: the G bucket counts go to every rank, the
// G x G matrix comes back, and ghist[b] becomes the position of this rank's run inside rank b's receive buffer.
struct HistPeer {
    P2PArgs   p2p;
    unsigned *matrix;   // [world][world] out: matrix[r][b] = pairs rank r sends to rank b
    unsigned *info;     // [0] pairs this rank receives, [1] overflow flag
    unsigned  capacity; // pairs a receive buffer can hold
};
template <bool kPeer> struct HistPeerParam { typedef NoPeer type; };
template <> struct HistPeerParam<true> { typedef HistPeer type; };

template <bool kBucket, bool kPeer = false>
__global__ void __launch_bounds__(kSortThreads)
    sort_hist_kernel(const unsigned *__restrict__ keys, size_t n, unsigned flip, const unsigned *__restrict__ splitters,
                     unsigned nspl, unsigned *__restrict__ ghist, unsigned *__restrict__ ticket,
                     unsigned *__restrict__ counts_out, typename HistPeerParam<kPeer>::type peer)
{
    constexpr int       P = kBucket ? 1 : 4;
    __shared__ unsigned h[P][kRadix];
    __shared__ unsigned s_spl[kRadix];
    __shared__ bool     s_last;
    for (int i = threadIdx.x; i < P * kRadix; i += kSortThreads)
        (&h[0][0])[i] = 0;
    if (kBucket && threadIdx.x < nspl)
        s_spl[threadIdx.x] = splitters[threadIdx.x];
    __syncthreads();
    DigitOp<kBucket> op{0, flip, nspl, s_spl};

    auto count = [&](unsigned key) {
        if (kBucket) {
            atomicAdd(&h[0][op(key)], 1u);
        }
        else {
            const unsigned o = key ^ flip;
            atomicAdd(&h[0][o & 255u], 1u);
            atomicAdd(&h[1 % P][(o >> 8) & 255u], 1u);
            atomicAdd(&h[2 % P][(o >> 16) & 255u], 1u);
            atomicAdd(&h[3 % P][o >> 24], 1u);
        }
    };
    const uintptr_t addr = (uintptr_t)keys;
    size_t          head = ((16 - (addr & 15)) & 15) / 4;
    if (head > n)
        head = n;
    const size_t nvec   = (n - head) / 4;
    const size_t tail   = head + nvec * 4;
    const uint4 *vin    = reinterpret_cast<const uint4 *>(keys + head);
    const size_t stride = (size_t)gridDim.x * kSortThreads;
    for (size_t v = (size_t)blockIdx.x * kSortThreads + threadIdx.x; v < nvec; v += stride) {
        uint4 r = ld_stream_u4(vin + v);
        count(r.x);
        count(r.y);
        count(r.z);
        count(r.w);
    }
    if (blockIdx.x == 0) {
        for (size_t i = threadIdx.x; i < head; i += kSortThreads)
            count(keys[i]);
        for (size_t i = tail + threadIdx.x; i < n; i += kSortThreads)
            count(keys[i]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < P * kRadix; i += kSortThreads) {
        unsigned c = (&h[0][0])[i];
        if (c)
            atomicAdd(&ghist[i], c);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        s_last     = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    finish_digit_offsets(ghist, P, kBucket ? counts_out : nullptr, nspl);
    if (threadIdx.x == 0)
        *ticket = 0;
    if constexpr (kPeer) {
        __shared__ unsigned long long s_seq;
        __shared__ unsigned           s_mat[kP2PMaxRanks * kP2PMaxRanks];
        __shared__ unsigned           s_over;
        const unsigned                world = peer.p2p.world, rank = peer.p2p.rank;
        if (threadIdx.x == 0)
            s_over = 0;
        __syncthreads(); // counts_out (written by warp 0 above) is visible to the CTA
        const unsigned seq = p2p_exchange(peer.p2p, counts_out, world, &s_seq);
        if (threadIdx.x < world * world) {
            const unsigned m = p2p_peer_word(peer.p2p, seq, threadIdx.x / world, threadIdx.x % world);
            s_mat[threadIdx.x]       = m;
            peer.matrix[threadIdx.x] = m;
        }
        __syncthreads();
        if (threadIdx.x < world) {
            unsigned off = 0, col = 0;
            for (unsigned r = 0; r < world; r++) {
                const unsigned m = s_mat[r * world + threadIdx.x];
                off += (r < rank) ? m : 0u;
                col += m;
            }
            ghist[threadIdx.x] = off; // sources land in rank order = original index order (global stability)
            if (col > peer.capacity)
                atomicOr(&s_over, 1u);
            if (threadIdx.x == rank)
                peer.info[0] = col;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            peer.info[1] = s_over;
            if (s_over) // nothing will be sent: the local sort behind this exchange must find nothing to sort either
                peer.info[0] = 0; // (the sizes that would be needed are in the matrix)
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// Radix case of the above with the byte-histogram technique of histogram.cu: four [bin][lane] tables
// (one per key byte, 128 KiB of shared memory), lane l only touches column l -> every shared atomic
// is bank-conflict-free whatever the key distribution.  One CTA of 1024 threads per SM.
// ---------------------------------------------------------------------------------------------------
constexpr int kHist4Threads = 1024;

// Pass plan of the four-pass sort, decided ON THE DEVICE by the last CTA of the histogram kernel: a pass whose digit
// is the same for every key (one bin holds all n) is the identity permutation -- a stable partition of one class --
// and is not run at all.  The mergeSort sample's own keys are rand() % 65536 (mergeSort/main.cpp:48,68): two of its
// four passes are of that kind.  The passes that do run are chained on the fly: the first reads the caller's arrays,
// the last writes the caller's dst, the ones in between alternate over the two pair buffers.
// plan[p] = {skip, in_fmt, out_fmt, in_sel, out_sel, -, -, -}; fmt 0: split key / value arrays, 1: (key, value) pairs;
// sel 0: caller's source, 1: pair buffer P0, 2: pair buffer P1, 3: caller's dst.
struct SortPlanBufs {
    const void *src_k, *src_v; // src_v == nullptr: the source is a pair buffer
    void       *p0, *p1;
    void       *dst_k, *dst_v;
};
__device__ __forceinline__ void make_sort_plan(unsigned *plan, const unsigned *trivial, bool src_pairs)
{
    int  first = -1, last = -1;
    bool run[4];
    for (int p = 0; p < 4; p++) {
        run[p] = !trivial[p];
        if (run[p]) {
            if (first < 0)
                first = p;
            last = p;
        }
    }
    if (first < 0) { // nothing to sort by: the last pass still moves the data from the source to dst (identity ranking)
        run[3] = true;
        first = last = 3;
    }
    unsigned cur = 0; // where the data is now (sel)
    for (int p = 0; p < 4; p++) {
        unsigned *q = plan + p * 8;
        if (!run[p]) {
            q[0] = 1;
            continue;
        }
        const unsigned out = p == last ? 3u : (cur == 1u ? 2u : 1u);
        q[0] = 0;
        q[1] = cur == 0 ? (src_pairs ? 1u : 0u) : 1u;
        q[2] = out == 3 ? 0u : 1u;
        q[3] = cur;
        q[4] = out;
        cur  = out;
    }
}

__global__ void __launch_bounds__(kHist4Threads, 1)
    sort_hist4_kernel(const unsigned *__restrict__ keys, size_t n, unsigned flip, unsigned *__restrict__ ghist,
                      unsigned *__restrict__ ticket, const unsigned *__restrict__ n_dev, uint4 *__restrict__ zero,
                      size_t zero_vecs, bool pairs /* keys[2 i] is key i: the input is a (key, value) pair buffer */,
                      unsigned *__restrict__ plan /* nullable: [4][8], see make_sort_plan */)
{
    // the look-back descriptors of the four passes start at zero: stored from here (fire and forget, under the key
    // reads) instead of by a memset in front of the kernel
    for (size_t i = (size_t)blockIdx.x * kHist4Threads + threadIdx.x; i < zero_vecs; i += (size_t)gridDim.x * kHist4Threads)
        zero[i] = make_uint4(0u, 0u, 0u, 0u);
    if (n_dev) // pair count produced on the device (receive buffer of the multi-GPU sort); n is its capacity
        n = *n_dev < n ? *n_dev : n;
    extern __shared__ __align__(16) unsigned tab[]; // [4][256][32]
    __shared__ bool s_last;
    const unsigned  lane = threadIdx.x & 31, warp = threadIdx.x >> 5, lane4 = lane * 4;
    for (int i = threadIdx.x; i < 4 * kRadix * 32; i += kHist4Threads)
        tab[i] = 0;
    __syncthreads();
    char *base = reinterpret_cast<char *>(tab);
    auto  count = [&](unsigned key) {
        const unsigned o = key ^ flip;
        atomicAdd(reinterpret_cast<unsigned *>(base + (((o << 7) & 0x7F80u) | lane4)), 1u);
        atomicAdd(reinterpret_cast<unsigned *>(base + 32768 + (((o >> 1) & 0x7F80u) | lane4)), 1u);
        atomicAdd(reinterpret_cast<unsigned *>(base + 65536 + (((o >> 9) & 0x7F80u) | lane4)), 1u);
        atomicAdd(reinterpret_cast<unsigned *>(base + 98304 + (((o >> 17) & 0x7F80u) | lane4)), 1u);
    };
    const uintptr_t addr = (uintptr_t)keys;
    size_t          head = pairs ? 0 : ((16 - (addr & 15)) & 15) / 4; // (a pair buffer is 16-byte aligned)
    if (head > n)
        head = n;
    const size_t per    = pairs ? 2 : 4; // keys per 16-byte vector
    const size_t nvec   = (n - head) / per;
    const size_t tail   = head + nvec * per;
    const uint4 *vin    = reinterpret_cast<const uint4 *>(keys + head);
    const size_t stride = (size_t)gridDim.x * kHist4Threads;
    size_t       v      = (size_t)blockIdx.x * kHist4Threads + threadIdx.x;
    for (; v + 3 * stride < nvec; v += 4 * stride) {
        uint4 r[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            r[k] = ld_stream_u4(vin + v + k * stride);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            count(r[k].x);
            if (!pairs)
                count(r[k].y);
            count(r[k].z);
            if (!pairs)
                count(r[k].w);
        }
    }
    for (; v < nvec; v += stride) {
        uint4 r = ld_stream_u4(vin + v);
        count(r.x);
        if (!pairs)
            count(r.y);
        count(r.z);
        if (!pairs)
            count(r.w);
    }
    if (blockIdx.x == 0) {
        for (size_t i = threadIdx.x; i < head; i += kHist4Threads)
            count(keys[i]);
        for (size_t i = tail + threadIdx.x; i < n; i += kHist4Threads)
            count(keys[pairs ? 2 * i : i]);
    }
    __syncthreads();
    // fold the lane columns: 1024 rows, 32 per warp; lane r keeps row warp*32 + r
    unsigned mine = 0;
#pragma unroll 8
    for (int r = 0; r < 32; r++) {
        unsigned t = __reduce_add_sync(0xffffffffu, tab[(warp * 32 + r) * 32 + lane]);
        if ((int)lane == r)
            mine = t;
    }
    if (mine)
        atomicAdd(&ghist[warp * 32 + lane], mine);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = atomicAdd(ticket, 1u);
        s_last     = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    if (plan) { // (before the counts become offsets) which passes have a single digit?
        __shared__ unsigned s_trivial[4];
        if (warp < 4) {
            unsigned any = 0;
#pragma unroll
            for (int i = 0; i < 8; i++)
                any |= (__ldcg(ghist + warp * kRadix + lane * 8 + i) == (unsigned)n) ? 1u : 0u;
            any = __reduce_or_sync(0xffffffffu, any);
            if (lane == 0)
                s_trivial[warp] = any;
        }
        __syncthreads();
        if (threadIdx.x == 0)
            make_sort_plan(plan, s_trivial, pairs);
    }
    finish_digit_offsets(ghist, 4, nullptr, 0);
    if (threadIdx.x == 0)
        *ticket = 0;
}

// ---------------------------------------------------------------------------------------------------
// One scatter pass (chained scan over tiles).  Measured limits (ncu + per-tile traces, profiles/r02*): the pass is
// bound by SM-side work per pair and by phases that wait -- not by DRAM (36 %).  Round 1 ranked first and published
// the tile's digit counts afterwards; 24 % of all warp stall samples sat at barriers, 16 % at the one that waits for
// the look-back.  Order of a tile now:
//   1. keys in (warp-striped), then ONE shared RED per key on the warp's private digit row ("early counts");
//   2. the digit owners (threads 0..255) add the warps' counts and PUBLISH the tile aggregate right away (~4 us
//      earlier than before: every successor's look-back finds it already there), scan the 256 counts and turn the
//      rows into running scatter positions;
//   3. ranking: eight vote.ballot per key find the lanes with the same digit (registers only), the group's lowest
//      lane bumps the warp's scatter position with one shared atomic, and the (key, value) pair goes straight to its
//      tile-local slot as one 64-bit store -- no rank array, no second staging loop;
//   4. look-back by the digit owners (kLB descriptor rows per round trip), inclusive prefix published;
//   5. every digit's run leaves as contiguous stores.
// Index order inside the tile = (warp, item, lane): ranks are stable by construction.
// ---------------------------------------------------------------------------------------------------
// kIO: layout of the pass's input / output.  0: keys and values in separate arrays on both sides (the caller's
// layout); 1: separate in, interleaved (key, value) pairs out; 2: pairs in, pairs out; 3: pairs in, separate out.
// The three intermediate results of a four-pass sort live in library-owned pair buffers: one 8-byte load and one 8-byte
// store per pair instead of two of each, and a digit run is one contiguous 8 B x count segment.
template <bool kBucket, int kLB, bool kPeer, int kIO, bool kProlog = true>
__device__ __forceinline__ void
sort_pass_body(const unsigned *__restrict__ kin, const unsigned *__restrict__ vin, unsigned *__restrict__ kout,
               unsigned *__restrict__ vout, unsigned n, unsigned shift, unsigned flip,
               const unsigned *__restrict__ splitters, unsigned nspl, const unsigned *__restrict__ gbase,
               unsigned *__restrict__ ticket, unsigned *__restrict__ desc, typename PeerParam<kPeer>::type peer,
               const unsigned *__restrict__ n_dev)
{
    if (kProlog && n_dev) { // device-side pair count: the grid covers the capacity, CTAs beyond the data leave at once
        n = *n_dev < n ? *n_dev : n;
        if ((size_t)blockIdx.x * kSortTile >= n)
            return; // (tickets are only drawn by CTAs that have a tile: tiles stay dense in ticket order)
    }
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SortSmem      &S    = *reinterpret_cast<SortSmem *>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = lanemask_lt();
    if constexpr (kPeer) {
        if (peer.info[1] != 0) // some receive buffer is too small: send nothing (the caller falls back)
            return;
    }

    if (kProlog) { // (the planned kernel has done this already, under the load of its plan entry)
        if (tid == 0)
            S.tile = atomicAdd(ticket, 1u);
        uint4 *z = reinterpret_cast<uint4 *>(&S.wcnt[0][0]);
        for (int i = tid; i < kSortWarps * kRadix / 4; i += kSortThreads)
            z[i] = make_uint4(0u, 0u, 0u, 0u);
        if (kBucket && tid < nspl)
            S.splitters[tid] = splitters[tid];
        __syncthreads();
    }
    const unsigned   tile  = S.tile;
    SORT_TRACE(tile, 0);
    const unsigned   base  = tile * kSortTile;
    const unsigned   valid = (n - base < kSortTile) ? (n - base) : kSortTile;
    const bool       full  = valid == kSortTile;
    DigitOp<kBucket> op{shift, flip, nspl, S.splitters};

    // ---- 1. keys (warp-striped: item i of lane l is element warp*(items*32) + i*32 + l of the tile), early counts --
    unsigned       key[kSortItems];
    const unsigned wbase = warp * (kSortItems * 32) + lane;
    const unsigned pad   = ~flip; // (pad ^ flip) = 0xffffffff: last digit in every pass, ranks after all valid keys
    unsigned      *mycnt = S.wcnt[warp];
    constexpr bool kInPk = (kIO == 2 || kIO == 3), kOutPk = (kIO == 1 || kIO == 2);
    unsigned       val[kSortItems];
    if (kInPk) { // one 8-byte load brings the pair
        const uint2 *pin = reinterpret_cast<const uint2 *>(kin);
#pragma unroll
        for (int i = 0; i < kSortItems; i++) {
            uint2 kv = make_uint2(pad, 0u);
            if (full || wbase + i * 32 < valid)
                kv = __ldg(pin + base + wbase + i * 32);
            key[i] = kv.x;
            val[i] = kv.y;
        }
    }
    else if (full) {
#pragma unroll
        for (int i = 0; i < kSortItems; i++)
            key[i] = __ldg(kin + base + wbase + i * 32);
    }
    else {
#pragma unroll
        for (int i = 0; i < kSortItems; i++)
            key[i] = (wbase + i * 32 < valid) ? __ldg(kin + base + wbase + i * 32) : pad;
    }
#ifdef B200_SCAN_TRACE
    if (key[kSortItems - 1] == 0x12345678u)
        SORT_TRACE(tile, 7); // (forces the last key load to complete before slot 1 is stamped)
    SORT_TRACE(tile, 1);
#endif
    unsigned dig[kBucket ? kSortItems / 4 : 1]; // bucket digits (a binary search each) are computed once and kept
    if (kBucket) {
#pragma unroll
        for (int i = 0; i < kSortItems / 4; i++)
            dig[i] = 0;
    }
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const unsigned d = op(key[i]);
        if (kBucket)
            dig[i / 4] |= d << (8 * (i & 3));
        atomicAdd(mycnt + d, 1u);
    }
    SORT_TRACE(tile, 2);
    __syncthreads();

    auto     load_vals = [&](int lo, int hi) {
        if (kInPk)
            return; // the values came with the keys
        if (full) {
#pragma unroll
            for (int i = lo; i < hi; i++)
                val[i] = __ldg(vin + base + wbase + i * 32);
        }
        else {
#pragma unroll
            for (int i = lo; i < hi; i++)
                val[i] = (wbase + i * 32 < valid) ? __ldg(vin + base + wbase + i * 32) : 0u;
        }
    };
#if B200_SORT_VALS_BEFORE_OWNERS
    load_vals(0, kValsEarly);
#endif
    // ---- 2. per digit: add the warps' counts, publish the tile aggregate, scan, running scatter positions --------
    {
        unsigned count = 0, tile_start = 0;
        if (tid < kRadix) {
#pragma unroll
            for (int w = 0; w < kSortWarps; w++)
                count += S.wcnt[w][tid];
            st_relaxed_u32(desc + (size_t)tile * kRadix + tid, (tile == 0 ? kDescP : kDescA) | count);
            unsigned inc = count;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned t = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d)
                    inc += t;
            }
            if (lane == 31)
                S.wsum[warp] = inc;
            tile_start = inc - count;
        }
        SORT_TRACE(tile, 3);
        __syncthreads();
        if (tid < kRadix) {
#pragma unroll
            for (int w = 0; w < kRadix / 32; w++)
                tile_start += (w < (int)warp) ? S.wsum[w] : 0u;
            unsigned run = tile_start;
#pragma unroll
            for (int w = 0; w < kSortWarps; w++) { // (re-read rather than 16 registers kept across the barrier)
                const unsigned cw = S.wcnt[w][tid];
                S.wcnt[w][tid]    = run;
                run += cw;
            }
            S.cnt[tid]    = count; // parked: the ranking loop needs every register
            S.tstart[tid] = tile_start;
        }
    }
    __syncthreads();

    // Look-back state of digit owner `tid`: it walks over the earlier tiles' counts of its digit, kLB rows per round
    // trip.  The round trips are issued from INSIDE the ranking loop (every kSortItems / kLbSteps keys: the loads of one
    // step are consumed by the next, ~1500 cycles later), so the walk costs the tile no time of its own and the
    // inclusive prefix is published several microseconds earlier -- which shortens every successor's walk.
    unsigned  lb_excl = 0, lb_d[kLB];
    int       lb_back = (int)tile - 1; // next row to consume
    unsigned  lb_state = (tid < kRadix && tile > 0) ? 0u : 2u; // 0: nothing in flight, 1: lb_d in flight, 2: done
#ifdef B200_SCAN_TRACE
    unsigned lb_polls = 0;
#endif
    auto lb_issue = [&]() {
#pragma unroll
        for (int j = 0; j < kLB; j++)
            lb_d[j] = (lb_back - j >= 0) ? ld_relaxed_u32(desc + tid + (size_t)(lb_back - j) * kRadix) : kDescP;
        lb_state = 1;
#ifdef B200_SCAN_TRACE
        lb_polls++;
#endif
    };
    auto lb_consume = [&]() {
        bool done = false;
#pragma unroll
        for (int j = 0; j < kLB; j++) {
            if (done || (lb_d[j] >> 30) == 0)
                break; // prefix found, or a row that is not published yet: poll again from there
            lb_excl += lb_d[j] & kDescValue;
            lb_back--;
            done = (lb_d[j] & kDescP) != 0;
        }
        lb_state = 0;
        if (done) {
            st_relaxed_u32(desc + (size_t)tile * kRadix + tid, kDescP | (lb_excl + S.cnt[tid]));
            lb_state = 2;
        }
    };
    auto lb_step = [&]() { // warp-uniform for warps 0..7, a no-op for the others
        if (lb_state == 1)
            lb_consume();
        if (lb_state == 0)
            lb_issue();
    };
    auto look_back_finish = [&]() {
        if (tid < kRadix) {
            while (lb_state != 2)
                lb_step();
#ifdef B200_SCAN_TRACE
            if (g_sort_trace && threadIdx.x == 0)
                g_sort_trace[(size_t)tile * 8 + 7] = lb_polls;
#endif
            S.gofs[tid] = __ldg(gbase + tid) + lb_excl - S.tstart[tid];
            SORT_TRACE(tile, 4);
        }
    };
#if B200_SORT_LB_FIRST
    look_back_finish();
#endif

    // ---- 3. values, then ranking: the pair goes straight to its tile-local slot ---------------------------------
#if !B200_SORT_VALS_BEFORE_OWNERS
    load_vals(0, kValsEarly);
#endif
#if !B200_SORT_PACKED
    unsigned slot2[kSortItems / 2]; // two 16-bit slots per register
#endif
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        unsigned kk = key[i];
        asm volatile("" : "+r"(kk)); // recompute the digit here: keeping 16 digits/addresses alive from the counting
                                     // loop costs more registers than the two instructions that rebuild them
        const unsigned d     = kBucket ? ((dig[i / 4] >> (8 * (i & 3))) & 255u) : op(kk);
        const unsigned mask  = match_digit(d);
        const unsigned below = mask & lt;
        unsigned       pre   = 0;
#if B200_SORT_LEADER_HIGH
        if ((mask & ~lt & ~(1u << lane)) == 0) // no higher lane in the group
            pre = atomicAdd(mycnt + d, (unsigned)__popc(mask));
        pre = __shfl_sync(0xffffffffu, pre, 31 - __clz(mask));
#else
        if (below == 0)
            pre = atomicAdd(mycnt + d, (unsigned)__popc(mask));
        pre = __shfl_sync(0xffffffffu, pre, __ffs(mask) - 1);
#endif
        const unsigned slot = pre + __popc(below);
        if (kLbSteps > 0 && (i + 1) % (kSortItems / (kLbSteps > 0 ? kLbSteps : 1)) == 0)
            lb_step();
#if B200_SORT_PACKED
        S.pair[slot] = make_uint2(kk, val[i]);
#else
        S.buf[slot] = kk;
        if (i & 1)
            slot2[i / 2] |= slot << 16;
        else
            slot2[i / 2] = slot;
#endif
    }
    SORT_TRACE(tile, 6);
#if !B200_SORT_PACKED
    load_vals(kValsEarly, kSortItems);
#endif
    // ---- 4. look-back ----------------------------------------------------------------------------------------
#if !B200_SORT_LB_FIRST
    look_back_finish();
#endif
#if !B200_SORT_PACKED
#pragma unroll
    for (int i = 0; i < kSortItems; i++)
        S.vbuf[(i & 1) ? slot2[i / 2] >> 16 : slot2[i / 2] & 0xffffu] = val[i];
#endif
    __syncthreads();
    SORT_TRACE(tile, 5);
    // ---- 5. out: every digit's run leaves as contiguous stores -------------------------------------------------
#if B200_SORT_PACKED
#define SORT_KEY_AT(j) S.pair[j].x
#define SORT_PAIR_AT(j) const uint2 pr = S.pair[j]; const unsigned x = pr.x, y = pr.y
#else
#define SORT_PAIR_AT(j) const unsigned x = S.buf[j], y = S.vbuf[j]
#endif
    if constexpr (kPeer) {
        if (peer.v[0] == 0ull && peer.bulk) {
            // Pair-format receive buffers, runs sent by the bulk-copy engine (cp.async.bulk shared -> peer global): the
            // stores of a tile no longer sit in the LSU queue in front of the other CTA's shared-memory traffic while
            // NVLink pushes back, and the CTA's warps are free the moment the copies are issued.  A bulk copy wants
            // 16-byte alignment on both sides and a pair has 8 bytes: the runs are re-spaced in shared memory (at most
            // one slot in front of each bucket) so that a run's slot has the parity of its index in the receive buffer;
            // an odd first / last pair travels as an ordinary 8-byte store.
            const unsigned nb = nspl + 1;
            if (tid == 0) {
                unsigned p = 0;
                for (unsigned d = 0; d < nb; d++) {
                    p += (p + S.gofs[d]) & 1u; // slot tstart + p and index gofs + tstart: same parity
                    S.respace[d] = p;
                }
            }
            __syncthreads();
            uint2    pr[kSortItems];
            unsigned ns[kSortItems];
#pragma unroll
            for (int k = 0; k < kSortItems; k++) {
                const unsigned j = tid + k * kSortThreads;
                pr[k]            = S.pair[j];
                ns[k]            = j + S.respace[op(pr[k].x)];
            }
            __syncthreads(); // every slot has been read: the runs may move
#pragma unroll
            for (int k = 0; k < kSortItems; k++)
                S.pair[ns[k]] = pr[k];
            fence_proxy_async_smem(); // the copy engine reads what the threads wrote
            __syncthreads();
            if (tid < nb) {
                const unsigned d = tid;
                unsigned       c = S.cnt[d];
                if (d == nb - 1)
                    c -= kSortTile - valid; // the pads of a ragged last tile close the last bucket's run
                unsigned s = S.tstart[d] + S.respace[d];
                unsigned g = S.gofs[d] + S.tstart[d];
                uint2   *q = reinterpret_cast<uint2 *>(peer.k[d]);
                if (c && (g & 1u)) {
                    q[g] = S.pair[s];
                    s++, g++, c--;
                }
                unsigned body = c & ~1u;
                if (c & 1u)
                    q[g + body] = S.pair[s + body];
                while (body) { // 16 KiB per copy
                    const unsigned m = body < 2048u ? body : 2048u;
                    bulk_s2g(q + g, &S.pair[s], m * 8u);
                    s += m, g += m, body -= m;
                }
                bulk_commit();
                bulk_wait_all<0>(); // written, not just read: the kernel's end is what the barrier kernel relies on
            }
        }
        else if (peer.v[0] == 0ull) { // pair-format receive buffers: half the store instructions, runs twice as long
#pragma unroll
            for (int k = 0; k < kSortItems; k++) {
                const unsigned j = tid + k * kSortThreads;
                SORT_PAIR_AT(j);
                const unsigned d = op(x); // bucket == destination rank
                const unsigned g = S.gofs[d] + j;
                if (j < valid)
                    reinterpret_cast<uint2 *>(peer.k[d])[g] = make_uint2(x, y);
            }
        }
        else {
#pragma unroll
            for (int k = 0; k < kSortItems; k++) {
                const unsigned j = tid + k * kSortThreads;
                SORT_PAIR_AT(j);
                const unsigned d = op(x);
                const unsigned g = S.gofs[d] + j;
                if (j < valid) {
                    reinterpret_cast<unsigned *>(peer.k[d])[g] = x;
                    reinterpret_cast<unsigned *>(peer.v[d])[g] = y;
                }
            }
        }
    }
    else if (kOutPk) {
        uint2 *pout = reinterpret_cast<uint2 *>(kout);
#pragma unroll
        for (int k = 0; k < kSortItems; k++) {
            const unsigned j = tid + k * kSortThreads;
            SORT_PAIR_AT(j);
            const unsigned g = S.gofs[op(x)] + j;
            if (full || j < valid)
                pout[g] = make_uint2(x, y);
        }
    }
    else if (full) {
#pragma unroll
        for (int k = 0; k < kSortItems; k++) {
            const unsigned j = tid + k * kSortThreads;
            SORT_PAIR_AT(j);
            const unsigned g = S.gofs[op(x)] + j;
            kout[g]          = x;
            vout[g]          = y;
        }
    }
    else {
#pragma unroll
        for (int k = 0; k < kSortItems; k++) {
            const unsigned j = tid + k * kSortThreads;
            SORT_PAIR_AT(j);
            const unsigned g = S.gofs[op(x)] + j;
            if (j < valid) {
                kout[g] = x;
                vout[g] = y;
            }
        }
    }
}

template <bool kBucket, int kLB, bool kPeer = false, int kIO = 0>
__global__ void __launch_bounds__(kSortThreads, kSortCtasPerSm)
    sort_pass_kernel(const unsigned *__restrict__ kin, const unsigned *__restrict__ vin, unsigned *__restrict__ kout,
                     unsigned *__restrict__ vout, unsigned n, unsigned shift, unsigned flip,
                     const unsigned *__restrict__ splitters, unsigned nspl, const unsigned *__restrict__ gbase,
                     unsigned *__restrict__ ticket, unsigned *__restrict__ desc, typename PeerParam<kPeer>::type peer,
                     const unsigned *__restrict__ n_dev)
{
    sort_pass_body<kBucket, kLB, kPeer, kIO>(kin, vin, kout, vout, n, shift, flip, splitters, nspl, gbase, ticket, desc,
                                             peer, n_dev);
}

// A pass of the PLANNED four-pass sort: formats and buffers come from the device-side plan entry (make_sort_plan); a
// skipped pass leaves at once; the others run one of the four statically specialised bodies (a CTA-uniform switch, so
// each body keeps the register allocation and the code of its own kernel).
__global__ void __launch_bounds__(kSortThreads, kSortCtasPerSm)
    sort_pass_plan_kernel(const unsigned *__restrict__ plan_entry, const SortPlanBufs B, unsigned n, unsigned shift,
                          unsigned flip, const unsigned *__restrict__ gbase, unsigned *__restrict__ ticket,
                          unsigned *__restrict__ desc, const unsigned *__restrict__ n_dev)
{
    // the plan entry is fetched UNDER the tile's prologue (ticket draw, counters zeroed): one L2 round trip, not two
    const uint4    pl   = __ldg(reinterpret_cast<const uint4 *>(plan_entry));
    const unsigned osel = __ldg(plan_entry + 4);
    if (n_dev) {
        n = *n_dev < n ? *n_dev : n;
        if ((size_t)blockIdx.x * kSortTile >= n)
            return;
    }
    {
        extern __shared__ __align__(16) unsigned char smem_raw[];
        SortSmem &S = *reinterpret_cast<SortSmem *>(smem_raw);
        if (threadIdx.x == 0)
            S.tile = atomicAdd(ticket, 1u); // (a skipped pass draws tickets nobody reads)
        uint4 *z = reinterpret_cast<uint4 *>(&S.wcnt[0][0]);
        for (int i = threadIdx.x; i < kSortWarps * kRadix / 4; i += kSortThreads)
            z[i] = make_uint4(0u, 0u, 0u, 0u);
        __syncthreads();
    }
    if (pl.x)
        return; // every key has the same digit in this pass: nothing moves
    const unsigned *kin  = reinterpret_cast<const unsigned *>(pl.w == 0 ? B.src_k : (pl.w == 1 ? B.p0 : B.p1));
    const unsigned *vin  = reinterpret_cast<const unsigned *>(B.src_v);
    unsigned       *kout = reinterpret_cast<unsigned *>(osel == 3 ? B.dst_k : (osel == 1 ? B.p0 : B.p1));
    unsigned       *vout = reinterpret_cast<unsigned *>(B.dst_v);
    if (pl.y == 0 && pl.z != 0)
        sort_pass_body<false, 4, false, 1, false>(kin, vin, kout, vout, n, shift, flip, nullptr, 0u, gbase, ticket, desc,
                                                  NoPeer(), n_dev);
    else if (pl.y != 0 && pl.z != 0)
        sort_pass_body<false, 4, false, 2, false>(kin, vin, kout, vout, n, shift, flip, nullptr, 0u, gbase, ticket, desc,
                                                  NoPeer(), n_dev);
    else if (pl.y != 0)
        sort_pass_body<false, 4, false, 3, false>(kin, vin, kout, vout, n, shift, flip, nullptr, 0u, gbase, ticket, desc,
                                                  NoPeer(), n_dev);
    else
        sort_pass_body<false, 4, false, 0, false>(kin, vin, kout, vout, n, shift, flip, nullptr, 0u, gbase, ticket, desc,
                                                  NoPeer(), n_dev);
}

// ---------------------------------------------------------------------------------------------------
// Splitters of the multi-GPU sample sort, chosen on the device with no collective call: every rank takes 512
// regular samples of its keys (in ordered space: key ^ flip), the samples are exchanged through the peer windows
// (512 words = one slot), every rank sorts the same world x 512 samples with a bitonic network in shared memory
// and keeps samples [b * 512], b = 1 .. world-1 -- the same choice on every rank, with no host round trip.
// ---------------------------------------------------------------------------------------------------
constexpr unsigned kSplitSamples = kP2PSlotWords; // 512
constexpr int      kSplitThreads = 1024;
__global__ void __launch_bounds__(kSplitThreads, 1)
    sort_splitters_kernel(P2PArgs a, const unsigned *__restrict__ keys, unsigned n, unsigned flip,
                          unsigned *__restrict__ splitters /* world-1, ascending, ordered space */)
{
    __shared__ unsigned           s_own[kSplitSamples];
    __shared__ unsigned           s_all[kP2PMaxRanks * kSplitSamples]; // 32 KiB
    __shared__ unsigned long long s_seq;
    const unsigned                step = n / kSplitSamples > 0 ? n / kSplitSamples : 1;
    for (unsigned i = threadIdx.x; i < kSplitSamples; i += kSplitThreads) {
        // stratified, jittered positions: regular positions alias with periodic key patterns (keys = f(i mod 4096)
        // sampled every 2048 elements show two values) and would pick useless splitters
        unsigned h = i * 2654435761u + 0x9E3779B9u;
        h ^= h >> 15;
        h *= 2246822519u;
        h ^= h >> 13;
        const unsigned long long idx = (unsigned long long)i * step + h % step;
        s_own[i] = n ? (keys[idx < n ? idx : n - 1] ^ flip) : 0xffffffffu; // an empty shard votes for "high"
    }
    __syncthreads();
    const unsigned seq   = p2p_exchange(a, s_own, kSplitSamples, &s_seq);
    const unsigned total = a.world * kSplitSamples;
    unsigned       pow2  = 1;
    while (pow2 < total)
        pow2 <<= 1;
    for (unsigned i = threadIdx.x; i < pow2; i += kSplitThreads)
        s_all[i] = i < total ? p2p_peer_word(a, seq, i / kSplitSamples, i % kSplitSamples) : 0xffffffffu;
    __syncthreads();
    for (unsigned k = 2; k <= pow2; k <<= 1)
        for (unsigned j = k >> 1; j > 0; j >>= 1) {
            for (unsigned i = threadIdx.x; i < pow2; i += kSplitThreads) {
                const unsigned p = i ^ j;
                if (p > i) {
                    const unsigned x = s_all[i], y = s_all[p];
                    if (((i & k) == 0) == (x > y)) {
                        s_all[i] = y;
                        s_all[p] = x;
                    }
                }
            }
            __syncthreads();
        }
    if (threadIdx.x + 1 < a.world)
        splitters[threadIdx.x] = s_all[(threadIdx.x + 1) * kSplitSamples];
}

// scratch layout (bytes): [0,16) hist ticket | [16, 16+4 KiB) digit offsets [4][256]
//                         | [4112, 4128) pass tickets | [4128, 4256) pass plan | descriptors [passes][tiles][256] from 4256
constexpr size_t kSortOffHist = 16, kSortOffTickets = 16 + 4 * kRadix * 4, kSortOffPlan = kSortOffTickets + 16,
                 kSortOffDesc = kSortOffPlan + 4 * 32; // (+ the pass plan: 4 x 8 words, see SortPlanBufs)
static_assert(kSortOffDesc % 16 == 0, "descriptor rows are zeroed with 16-byte stores");

static int g_sort_lb = 4; // look-back descriptors fetched per round trip (1, 4 or 8)
static bool g_sort_peer_bulk = true; // peer partition pass (pair format): cp.async.bulk shared -> peer global
static bool g_sort_plan = true;  // passes whose digit is the same for every key are skipped (device-side plan)
static bool g_sort_pairs = true; // intermediate results as (key, value) pairs in library-owned buffers (2 x 8N bytes)
constexpr size_t kSortPairsMin = 1u << 16;

template <bool kBucket, int kLB>
static int launch_pass(unsigned tiles, cudaStream_t stream, const unsigned *kin, const unsigned *vin, unsigned *kout,
                       unsigned *vout, unsigned n, unsigned shift, unsigned flip, const unsigned *spl, unsigned nspl,
                       const unsigned *gbase, unsigned *ticket, unsigned *desc, const unsigned *n_dev)
{
    {
        int        st = 0;
        DeviceCtx *c  = ctx(&st);
        if (!c)
            return st;
        const unsigned bit = (kBucket ? kAttrSortPass3 : kAttrSortPass0) << (kLB <= 1 ? 0 : (kLB <= 4 ? 1 : 2));
        B200_TRY((cudaError_t)opt_in_smem(c, bit, sort_pass_kernel<kBucket, kLB>, sizeof(SortSmem)));
    }
    sort_pass_kernel<kBucket, kLB><<<tiles, kSortThreads, sizeof(SortSmem), stream>>>(
        kin, vin, kout, vout, n, shift, flip, spl, nspl, gbase, ticket, desc, NoPeer(), n_dev);
    return (int)cudaGetLastError();
}

// passes of the pair-buffer pipeline (kLB = 4 only): kIO 1 / 2 / 3
template <int kIO>
static int launch_pass_io(unsigned tiles, cudaStream_t stream, const void *in_a, const void *in_b, void *out_a,
                          void *out_b, unsigned n, unsigned shift, unsigned flip, const unsigned *gbase,
                          unsigned *ticket, unsigned *desc, const unsigned *n_dev)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc0 << (kIO - 1), sort_pass_kernel<false, 4, false, kIO>,
                                      sizeof(SortSmem)));
    sort_pass_kernel<false, 4, false, kIO><<<tiles, kSortThreads, sizeof(SortSmem), stream>>>(
        (const unsigned *)in_a, (const unsigned *)in_b, (unsigned *)out_a, (unsigned *)out_b, n, shift, flip, nullptr, 0u,
        gbase, ticket, desc, NoPeer(), n_dev);
    return (int)cudaGetLastError();
}

// a pass of the planned pipeline: formats and buffers are read from the device-side plan entry
static int launch_pass_dyn(unsigned tiles, cudaStream_t stream, const unsigned *plan_entry, const SortPlanBufs &bufs,
                           unsigned n, unsigned shift, unsigned flip, const unsigned *gbase, unsigned *ticket,
                           unsigned *desc, const unsigned *n_dev)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPassDyn, sort_pass_plan_kernel, sizeof(SortSmem)));
    sort_pass_plan_kernel<<<tiles, kSortThreads, sizeof(SortSmem), stream>>>(plan_entry, bufs, n, shift, flip, gbase, ticket,
                                                                           desc, n_dev);
    return (int)cudaGetLastError();
}

template <bool kBucket, typename... Args> static int launch_pass_lb(Args... args)
{
    if (g_sort_lb <= 1)
        return launch_pass<kBucket, 1>(args...);
    if (g_sort_lb <= 4)
        return launch_pass<kBucket, 4>(args...);
    return launch_pass<kBucket, 8>(args...);
}

} // namespace b200

using namespace b200;

static int sort_kv(uint32_t *d_dstKey, uint32_t *d_dstVal, uint32_t *d_bufKey, uint32_t *d_bufVal,
                   const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t N, uint32_t sortDir,
                   const uint32_t *d_n, b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (N == 0)
        return 0;
    if (N >= (1ull << 30))
        return (int)cudaErrorInvalidValue; // 30-bit look-back values; shard larger inputs
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const unsigned tiles      = (unsigned)((N + kSortTile - 1) / kSortTile);
    const size_t   desc_bytes = (size_t)tiles * kRadix * 4;
    const size_t   need       = kSortOffDesc + 4 * desc_bytes;
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    char *base = reinterpret_cast<char *>(c->sort.ptr);
    B200_TRY(cudaMemsetAsync(base, 0, kSortOffDesc, stream)); // ticket, digit histograms, pass tickets (4 KiB)
    unsigned *hticket = reinterpret_cast<unsigned *>(base);
    unsigned *ghist   = reinterpret_cast<unsigned *>(base + kSortOffHist);
    unsigned *tickets = reinterpret_cast<unsigned *>(base + kSortOffTickets);
    const unsigned flip = sortDir ? 0u : 0xffffffffu;
    unsigned *plan = reinterpret_cast<unsigned *>(base + kSortOffPlan);
    // planned pipeline (passes with a single digit are skipped, see make_sort_plan): the pair-buffer pipeline only
    const bool pair_pipeline = d_srcVal == nullptr || (g_sort_pairs && g_sort_lb == 4 && N >= kSortPairsMin);
    bool       planned       = g_sort_plan && pair_pipeline;
    if (planned && ensure(c->sort_pairs, 2 * N * sizeof(uint2), stream) != 0) {
        cudaGetLastError();
        planned = false;
    }

    {
        B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortHist4, sort_hist4_kernel, 4 * kRadix * 32 * 4));
        size_t want = (N / 4 + kHist4Threads - 1) / kHist4Threads;
        int    grid = (int)(want < (size_t)c->num_sms ? want : (size_t)c->num_sms);
        if (grid < 1)
            grid = 1;
        sort_hist4_kernel<<<grid, kHist4Threads, 4 * kRadix * 32 * 4, stream>>>(
            d_srcKey, N, flip, ghist, hticket, d_n, reinterpret_cast<uint4 *>(base + kSortOffDesc), 4 * desc_bytes / 16,
            d_srcVal == nullptr, planned ? plan : nullptr);
        B200_TRY(cudaGetLastError());
    }

    if (planned) {
        SortPlanBufs bufs;
        bufs.src_k = d_srcKey;
        bufs.src_v = d_srcVal;
        bufs.p0    = c->sort_pairs.ptr;
        bufs.p1    = reinterpret_cast<uint2 *>(c->sort_pairs.ptr) + N;
        bufs.dst_k = d_dstKey;
        bufs.dst_v = d_dstVal;
        for (int p = 0; p < 4; p++)
            B200_TRY((cudaError_t)launch_pass_dyn(tiles, stream, plan + p * 8, bufs, (unsigned)N, 8u * p, flip,
                                                  ghist + p * kRadix, tickets + p,
                                                  reinterpret_cast<unsigned *>(base + kSortOffDesc + p * desc_bytes), d_n));
        count_launch(5);
        return 0;
    }
    if (d_srcVal == nullptr) {
        // the input IS a pair buffer (the receive buffer of the sharded sort): every pass reads pairs
        B200_TRY((cudaError_t)ensure(c->sort_pairs, 2 * N * sizeof(uint2), stream));
        uint2 *P0 = reinterpret_cast<uint2 *>(c->sort_pairs.ptr), *P1 = P0 + N;
        auto   gb = [&](int p) { return (const unsigned *)(ghist + p * kRadix); };
        auto   ds = [&](int p) { return reinterpret_cast<unsigned *>(base + kSortOffDesc + p * desc_bytes); };
        B200_TRY((cudaError_t)launch_pass_io<2>(tiles, stream, d_srcKey, nullptr, P0, nullptr, (unsigned)N, 0u, flip,
                                                gb(0), tickets + 0, ds(0), d_n));
        B200_TRY((cudaError_t)launch_pass_io<2>(tiles, stream, P0, nullptr, P1, nullptr, (unsigned)N, 8u, flip, gb(1),
                                                tickets + 1, ds(1), d_n));
        B200_TRY((cudaError_t)launch_pass_io<2>(tiles, stream, P1, nullptr, P0, nullptr, (unsigned)N, 16u, flip, gb(2),
                                                tickets + 2, ds(2), d_n));
        B200_TRY((cudaError_t)launch_pass_io<3>(tiles, stream, P0, nullptr, d_dstKey, d_dstVal, (unsigned)N, 24u, flip,
                                                gb(3), tickets + 3, ds(3), d_n));
        count_launch(5);
        return 0;
    }
    if (g_sort_pairs && g_sort_lb == 4 && N >= kSortPairsMin) {
        // src -> P0 -> P1 -> P0 -> dst with the three intermediate results as (key, value) pairs in library scratch
        if (ensure(c->sort_pairs, 2 * N * sizeof(uint2), stream) == 0) {
            uint2    *P0 = reinterpret_cast<uint2 *>(c->sort_pairs.ptr), *P1 = P0 + N;
            auto      gb = [&](int p) { return (const unsigned *)(ghist + p * kRadix); };
            auto      ds = [&](int p) { return reinterpret_cast<unsigned *>(base + kSortOffDesc + p * desc_bytes); };
            B200_TRY((cudaError_t)launch_pass_io<1>(tiles, stream, d_srcKey, d_srcVal, P0, nullptr, (unsigned)N, 0u, flip,
                                                    gb(0), tickets + 0, ds(0), d_n));
            B200_TRY((cudaError_t)launch_pass_io<2>(tiles, stream, P0, nullptr, P1, nullptr, (unsigned)N, 8u, flip, gb(1),
                                                    tickets + 1, ds(1), d_n));
            B200_TRY((cudaError_t)launch_pass_io<2>(tiles, stream, P1, nullptr, P0, nullptr, (unsigned)N, 16u, flip,
                                                    gb(2), tickets + 2, ds(2), d_n));
            B200_TRY((cudaError_t)launch_pass_io<3>(tiles, stream, P0, nullptr, d_dstKey, d_dstVal, (unsigned)N, 24u,
                                                    flip, gb(3), tickets + 3, ds(3), d_n));
            count_launch(5);
            return 0;
        }
        cudaGetLastError(); // no room for the pair buffers: the caller's buf arrays carry the intermediate results
    }
    const uint32_t *kin = d_srcKey, *vin = d_srcVal;
    for (int p = 0; p < 4; p++) {
        uint32_t *kout = (p & 1) ? d_dstKey : d_bufKey;
        uint32_t *vout = (p & 1) ? d_dstVal : d_bufVal;
        unsigned *desc = reinterpret_cast<unsigned *>(base + kSortOffDesc + p * desc_bytes);
        B200_TRY((cudaError_t)launch_pass_lb<false>(tiles, stream, kin, vin, kout, vout, (unsigned)N, 8u * p, flip,
                                                    (const unsigned *)nullptr, 0u, (const unsigned *)(ghist + p * kRadix),
                                                    tickets + p, desc, (const unsigned *)d_n));
        kin = kout;
        vin = vout;
    }
    count_launch(5);
    return 0;
}

extern "C" int b200_sort_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, uint32_t *d_bufKey, uint32_t *d_bufVal,
                                  const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t N, uint32_t sortDir,
                                  b200_stream_t stream)
{
    return sort_kv(d_dstKey, d_dstVal, d_bufKey, d_bufVal, d_srcKey, d_srcVal, N, sortDir, nullptr, stream);
}

// The same with the pair count read on the device: sorts the first min(*d_n, capacity) pairs; the launch geometry
// and the scratch are sized for `capacity`.  Lets the multi-GPU sort run on a receive buffer whose fill level only
// the GPU knows yet (no host round trip between the exchange and the local sort).
extern "C" int b200_sort_kv_dyn_async(uint32_t *d_dstKey, uint32_t *d_dstVal, uint32_t *d_bufKey, uint32_t *d_bufVal,
                                      const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t capacity,
                                      const uint32_t *d_n, uint32_t sortDir, b200_stream_t stream)
{
    if (!d_n)
        return (int)cudaErrorInvalidValue;
    return sort_kv(d_dstKey, d_dstVal, d_bufKey, d_bufVal, d_srcKey, d_srcVal, capacity, sortDir, d_n, stream);
}

extern "C" int b200_partition_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, const uint32_t *d_srcKey,
                                       const uint32_t *d_srcVal, size_t N, const uint32_t *d_splitters,
                                       uint32_t numBuckets, uint32_t sortDir, uint32_t *d_bucketCounts,
                                       b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (numBuckets < 1 || numBuckets > kRadix || N >= (1ull << 30))
        return (int)cudaErrorInvalidValue;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    if (N == 0) {
        B200_TRY(cudaMemsetAsync(d_bucketCounts, 0, numBuckets * 4, stream));
        return 0;
    }
    const unsigned tiles      = (unsigned)((N + kSortTile - 1) / kSortTile);
    const size_t   desc_bytes = (size_t)tiles * kRadix * 4;
    const size_t   need       = kSortOffDesc + desc_bytes;
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    char *base = reinterpret_cast<char *>(c->sort.ptr);
    B200_TRY(cudaMemsetAsync(base, 0, need, stream));
    unsigned *hticket = reinterpret_cast<unsigned *>(base);
    unsigned *ghist   = reinterpret_cast<unsigned *>(base + kSortOffHist);
    unsigned *tickets = reinterpret_cast<unsigned *>(base + kSortOffTickets);
    unsigned *desc    = reinterpret_cast<unsigned *>(base + kSortOffDesc);
    const unsigned flip = sortDir ? 0u : 0xffffffffu;
    const unsigned nspl = numBuckets - 1;

    size_t want = (N / 4 + kSortThreads - 1) / kSortThreads;
    size_t cap  = (size_t)c->num_sms * 4;
    int    grid = (int)(want < cap ? want : cap);
    if (grid < 1)
        grid = 1;
    sort_hist_kernel<true><<<grid, kSortThreads, 0, stream>>>(d_srcKey, N, flip, d_splitters, nspl, ghist, hticket,
                                                            d_bucketCounts, NoPeer());
    B200_TRY(cudaGetLastError());
    B200_TRY((cudaError_t)launch_pass_lb<true>(tiles, stream, (const unsigned *)d_srcKey, (const unsigned *)d_srcVal,
                                               d_dstKey, d_dstVal, (unsigned)N, 0u, flip, (const unsigned *)d_splitters,
                                               nspl, (const unsigned *)ghist, tickets, desc, (const unsigned *)nullptr));
    count_launch(2);
    return 0;
}

// Partition + all-to-all of the multi-GPU sample sort in two kernels and no collective call (b200_p2p_connect
// first; numBuckets == world, bucket b == rank b).  recvKeys / recvVals: HOST arrays of `world` device pointers,
// entry r = rank r's receive buffer (capacity pairs each) as mapped into THIS process (b200_p2p_buffer_*).
// Kernel 1 counts the buckets, exchanges the counts over the peer windows and derives, per destination, where
// this rank's run starts (sources in rank order => global stability); kernel 2 is the stable partition pass whose
// digit runs are stored straight into the destination ranks' receive buffers over NVLink; a node-wide barrier kernel
// follows it on the stream ("all my stores are out").  The local sort then runs with the received count read on the device
// (b200_sort_kv_dyn_async, d_n = d_info).  d_matrix: [world][world] counts (row = source); d_info: [0] pairs
// received, [1] != 0 when some rank would have received more than `capacity` (then nothing was sent).
extern "C" int b200_partition_exchange_kv_async(const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t N,
                                                const uint32_t *d_splitters, uint32_t sortDir,
                                                void *const *recvKeys, void *const *recvVals, size_t capacity,
                                                uint32_t *d_matrix, uint32_t *d_info, b200_stream_t stream_)
{
    cudaStream_t   stream = (cudaStream_t)stream_;
    const P2PArgs *a      = p2p_args();
    if (!a || a->world < 2 || N >= (1ull << 30) || capacity >= (1ull << 30))
        return (int)cudaErrorInvalidValue;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const unsigned world      = a->world, nspl = world - 1;
    const unsigned tiles      = (unsigned)((N + kSortTile - 1) / kSortTile);
    const size_t   desc_bytes = (size_t)(tiles ? tiles : 1) * kRadix * 4;
    const size_t   need       = kSortOffDesc + desc_bytes;
    B200_TRY((cudaError_t)ensure(c->sort, need, stream));
    char *base = reinterpret_cast<char *>(c->sort.ptr);
    B200_TRY(cudaMemsetAsync(base, 0, need, stream));
    unsigned *hticket = reinterpret_cast<unsigned *>(base);
    unsigned *ghist   = reinterpret_cast<unsigned *>(base + kSortOffHist);
    unsigned *counts  = ghist + kRadix; // second row of the digit table: raw bucket counts (the exchange payload)
    unsigned *tickets = reinterpret_cast<unsigned *>(base + kSortOffTickets);
    unsigned *desc    = reinterpret_cast<unsigned *>(base + kSortOffDesc);
    const unsigned flip = sortDir ? 0u : 0xffffffffu;

    HistPeer hp;
    hp.p2p      = *a;
    hp.matrix   = d_matrix;
    hp.info     = d_info;
    hp.capacity = (unsigned)capacity;
    size_t want = (N / 4 + kSortThreads - 1) / kSortThreads;
    size_t cap  = (size_t)c->num_sms * 4;
    int    grid = (int)(want < cap ? want : cap);
    if (grid < 1)
        grid = 1;
    sort_hist_kernel<true, true><<<grid, kSortThreads, 0, stream>>>(d_srcKey, N, flip, d_splitters, nspl, ghist,
                                                                  hticket, counts, hp);
    B200_TRY(cudaGetLastError());
    SortPeer sp;
    for (int r = 0; r < kP2PMaxRanks; r++) {
        sp.k[r] = r < (int)world ? (unsigned long long)(uintptr_t)recvKeys[r] : 0ull;
        sp.v[r] = (r < (int)world && recvVals) ? (unsigned long long)(uintptr_t)recvVals[r] : 0ull;
    }
    sp.info = d_info;
    sp.bulk = g_sort_peer_bulk ? 1u : 0u;
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPeer, sort_pass_kernel<true, 4, true>, sizeof(SortSmem)));
    // N == 0: one CTA that only takes part in the closing exchange
    sort_pass_kernel<true, 4, true><<<tiles ? tiles : 1, kSortThreads, sizeof(SortSmem), stream>>>(
        d_srcKey, d_srcVal, nullptr, nullptr, (unsigned)N, 0u, flip, d_splitters, nspl, ghist, tickets, desc, sp,
        nullptr);
    B200_TRY(cudaGetLastError());
    count_launch(2);
    // The pass kernel's peer stores are complete when it ends (no per-CTA system fence inside: that stalled every
    // tile for an NVLink round trip); a one-CTA kernel behind it then tells every peer and waits for every peer.
    return b200_p2p_barrier_async(stream_);
}

extern "C" int b200_tune_sort_peer_bulk(int on)
{
    g_sort_peer_bulk = on != 0;
    return 0;
}
extern "C" int b200_tune_sort_plan(int on)
{
    g_sort_plan = on != 0;
    return 0;
}
extern "C" int b200_tune_sort_pairs(int on)
{
    g_sort_pairs = on != 0;
    return 0;
}
// tuning knob for experiments (not part of the stable ABI)
extern "C" int b200_tune_sort_lookback(int per_round)
{
    g_sort_lb = per_round;
    return 0;
}

// Pre-size the sort scratch for up to N pairs (what initMergeSort() does for the sample's N <= 2^22, so the
// sample's single cold timed call does not pay for cudaMalloc: mergeSort.cu:455-461 allocates there too).
extern "C" int b200_sort_reserve(size_t N)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    const size_t tiles = (N + kSortTile - 1) / kSortTile;
    B200_TRY((cudaError_t)ensure(c->sort, kSortOffDesc + 4 * tiles * kRadix * 4, 0));
    if (g_sort_pairs) // (also what a pair-format input needs, whatever its size)
        B200_TRY((cudaError_t)ensure(c->sort_pairs, 2 * N * sizeof(uint2), 0));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc0, sort_pass_kernel<false, 4, false, 1>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc1, sort_pass_kernel<false, 4, false, 2>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc2, sort_pass_kernel<false, 4, false, 3>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPassDyn, sort_pass_plan_kernel, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortHist4, sort_hist4_kernel, 4 * kRadix * 32 * 4));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPass1, sort_pass_kernel<false, 4>, sizeof(SortSmem)));
    return 0;
}

// Splitters of the sharded sort on the device (b200_p2p_connect / b200_ctx_create first): d_splitters receives
// world-1 ascending values in ordered key space (key for ascending sorts, ~key for descending), identical on every
// rank.  Every rank must call it (it contains one exchange).
extern "C" int b200_sort_splitters_async(const uint32_t *d_srcKey, size_t N, uint32_t sortDir, uint32_t *d_splitters,
                                         b200_stream_t stream)
{
    const P2PArgs *a = p2p_args();
    if (!a || a->world < 1 || N >= (1ull << 32))
        return (int)cudaErrorInvalidValue;
    sort_splitters_kernel<<<1, kSplitThreads, 0, (cudaStream_t)stream>>>(*a, d_srcKey, (unsigned)N,
                                                                       sortDir ? 0u : 0xffffffffu, d_splitters);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}

namespace b200 {
// load every kernel of the sharded sort onto the current device and opt in to its shared memory, so that nothing
// but launches happens once kernels start waiting for each other (sharded.cu)
int sort_preload()
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    cudaFuncAttributes fa;
    B200_TRY(cudaFuncGetAttributes(&fa, sort_splitters_kernel));
    B200_TRY(cudaFuncGetAttributes(&fa, sort_hist_kernel<true, true>));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPeer, sort_pass_kernel<true, 4, true>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortHist4, sort_hist4_kernel, 4 * kRadix * 32 * 4));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPass1, sort_pass_kernel<false, 4>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPass0, sort_pass_kernel<false, 1>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPass2, sort_pass_kernel<false, 8>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc0, sort_pass_kernel<false, 4, false, 1>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc1, sort_pass_kernel<false, 4, false, 2>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrMisc2, sort_pass_kernel<false, 4, false, 3>, sizeof(SortSmem)));
    B200_TRY((cudaError_t)opt_in_smem(c, kAttrSortPassDyn, sort_pass_plan_kernel, sizeof(SortSmem)));
    return 0;
}
} // namespace b200

#ifdef B200_SCAN_TRACE
extern "C" int b200_debug_set_sort_trace(void *d_trace)
{
    unsigned long long *p = reinterpret_cast<unsigned long long *>(d_trace);
    B200_TRY(cudaMemcpyToSymbol(g_sort_trace, &p, sizeof(p)));
    return 0;
}
#endif
