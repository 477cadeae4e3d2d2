// p2p.cuh -- exchange of small per-rank payloads between the GPUs of one NVSwitch node, from INSIDE a kernel.
//
// The sharded primitives end in a tiny collective: 320 u32 bins, a few shard sums, G shard totals.
// Over NCCL that is a separate kernel launch plus a ring/tree protocol (~15-25 us); here the last CTA of the
// producing kernel stores its payload straight into every peer's window over NVLink (P2P stores), polls its own
// window for the peers' words and finishes the combine itself -- compute and collective in ONE kernel.
//
// Window (one per rank, cudaMalloc'ed, mapped into every other rank's process with CUDA IPC):
//   header  { u64 seq; u64 errors; }                       -- touched by the owning rank only
//   slot[parity 0..1][source rank 0..kP2PMaxRanks)  u64 word[kP2PSlotWords], word = { exchange number : data }
// Exchange number s (s = ++header.seq, every rank runs the same stream-ordered sequence of exchanges):
//   rank r stores payload word w as the single 64-bit value (s << 32 | data) into slot[s & 1][r].word[w] of EVERY
//   window; a reader polls the word in its OWN window until the upper half says s.  A 64-bit store is single-copy
//   atomic, so data and validity arrive together: no system-scope fence, no separate flag, one NVLink one-way
//   latency per exchange (the "low-latency" protocol idea of NCCL's LL, written here for our payloads).
//   Slot (s & 1) is rewritten by exchange s + 2, which a rank only starts after its kernel of exchange s + 1 has
//   finished, i.e. after every rank has published s + 1, i.e. after every rank's kernel of exchange s has ended.
// A bounded wait (kP2PTimeoutNs) turns a missing peer into an error count instead of a hung GPU.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int      kP2PMaxRanks  = 16;
constexpr unsigned kP2PSlotWords = 512;              // u32 payload words per rank per exchange
constexpr unsigned kP2PSlotBytes = 4 * kP2PSlotWords; // payload bytes
constexpr unsigned kP2PHeaderBytes = 64;
constexpr size_t   kP2PWindowBytes = kP2PHeaderBytes + 2ull * kP2PMaxRanks * kP2PSlotWords * 8;
constexpr unsigned long long kP2PTimeoutNs = 30ull * 1000 * 1000 * 1000; // default; b200_p2p_set_timeout_ms changes it

struct P2PArgs { // kernel parameter, by value
    unsigned long long win[kP2PMaxRanks]; // device addresses of all windows (own one included), 0 beyond world
    unsigned           rank, world;
    unsigned long long timeout_ns;        // bounded wait for a peer's word
};

__device__ __forceinline__ unsigned long long *p2p_slot(unsigned long long win, unsigned parity, unsigned src)
{
    return reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(win) + kP2PHeaderBytes) +
           (size_t)(parity * kP2PMaxRanks + src) * kP2PSlotWords;
}
__device__ __forceinline__ void st_sys_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long p2p_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Called by ALL threads of ONE CTA.  payload: nwords u32 (<= kP2PSlotWords) in global or shared memory, already
// visible to the calling CTA.  Publishes it to every rank (own window included) and returns the exchange number;
// any thread may then read any rank's words with p2p_peer_word(), which waits for the word to arrive.
__device__ __forceinline__ unsigned p2p_exchange(const P2PArgs &a, const unsigned *payload, unsigned nwords,
                                                 unsigned long long *s_seq /* one shared u64 */)
{
    unsigned long long *hdr = reinterpret_cast<unsigned long long *>(a.win[a.rank]);
    if (threadIdx.x == 0) {
        const unsigned long long s = hdr[0] + 1; // only this rank's (stream-ordered) kernels touch its header
        hdr[0]                     = s;
        *s_seq                     = s;
    }
    __syncthreads();
    const unsigned seq    = (unsigned)*s_seq; // 32 bits travel with every word
    const unsigned parity = seq & 1;
    for (unsigned i = threadIdx.x; i < a.world * nwords; i += blockDim.x) {
        const unsigned peer = i / nwords, w = i - peer * nwords;
        const unsigned v    = *reinterpret_cast<const volatile unsigned *>(payload + w); // never a stale L1 line
        st_sys_u64(p2p_slot(a.win[peer], parity, a.rank) + w, ((unsigned long long)seq << 32) | v);
    }
    return seq;
}

// word w of rank src's payload in exchange seq: polls this rank's own window until the word carries seq
__device__ __forceinline__ unsigned p2p_peer_word(const P2PArgs &a, unsigned seq, unsigned src, unsigned w)
{
    const unsigned long long *p  = p2p_slot(a.win[a.rank], seq & 1, src) + w;
    unsigned long long        v  = ld_sys_u64(p);
    if ((unsigned)(v >> 32) != seq) {
        const unsigned long long t0 = p2p_now();
        do {
            v = ld_sys_u64(p);
            if (p2p_now() - t0 > a.timeout_ns) {
                // A peer never arrived: this call's results are void.  The count is STICKY (never reset while the
                // windows live) and every host-side wrapper that synchronises checks it and fails loudly
                // (b200_p2p_errors; the b200_*_sharded entry points return cudaErrorLaunchTimeout).
                atomicAdd(reinterpret_cast<unsigned long long *>(a.win[a.rank]) + 1, 1ull);
                break;
            }
        } while ((unsigned)(v >> 32) != seq);
    }
    return (unsigned)v;
}

// host side (p2p.cu): the connected window set of the current device, or world == 0 when b200_p2p_connect() has
// not been called
const P2PArgs *p2p_args();
// ONE process driving all devices (sharded.cu): windows are plain cudaMalloc'ed blocks reached through peer access
int p2p_connect_local(int ndev, const int *devices);
int p2p_disconnect_local(int ndev, const int *devices);

} // namespace b200
