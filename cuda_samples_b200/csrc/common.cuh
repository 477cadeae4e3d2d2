// common.cuh -- shared device/host helpers for the sm_100a primitive library.
//
// Error convention (two layers, see include/b200prim.h):
//   * the extended C ABI (b200_*) returns an int status: 0 or a cudaError_t value;
//   * the drop-in entry points keep the samples' convention: print to stderr and
//     exit(EXIT_FAILURE) (reference: Common/helper_cuda.h:584-616).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define B200_TRY(expr)                                                                                       \
    do {                                                                                                     \
        cudaError_t _e = (expr);                                                                             \
        if (_e != cudaSuccess)                                                                               \
            return (int)_e;                                                                                  \
    } while (0)

#define B200_OR_DIE(expr, what)                                                                              \
    do {                                                                                                     \
        int _s = (int)(expr);                                                                                \
        if (_s != 0) {                                                                                       \
            fprintf(stderr, "%s(%d): %s failed: CUDA error %d (%s)\n", __FILE__, __LINE__, what, _s,         \
                    cudaGetErrorName((cudaError_t)_s));                                                      \
            exit(EXIT_FAILURE);                                                                              \
        }                                                                                                    \
    } while (0)

namespace b200 {

constexpr int kWarp = 32;

// ------------------------------------------------------------------------------------------------
// Per-device scratch.  The reference keeps file-static scratch allocated by init*() (e.g.
// histogram256.cu:133, scan.cu:182, mergeSort.cu:452); here every primitive asks the current
// device's context for grow-only scratch, so a missing init*() is harmless and sizes are not capped.
// Not thread-safe across host threads sharing a device -- same contract as the reference.
// ------------------------------------------------------------------------------------------------
struct Scratch {
    void  *ptr   = nullptr;
    size_t bytes = 0;
};

struct DeviceCtx {
    int     device   = -1;
    int     num_sms  = 0;
    int     smem_max = 0; // opt-in dynamic shared memory per block
    Scratch reduce;       // partials + ticket
    Scratch hist;         // accumulators + ticket
    Scratch scan;         // state + tile descriptors
    Scratch scan_tiles;   // sharded scan: [0] ticket, from byte 16 one u32 per 8192-element tile (sum, then prefix)
    Scratch sort;         // digit histograms + descriptors
    Scratch sort_pairs;   // two (key, value) pair buffers for the intermediate results of the radix passes
    Scratch image;        // integral image: band / block carries
    Scratch stage[2];     // device staging chunks of the host-buffer entry points
    Scratch host_work;    // their own reduce scratch + chunk rows (never shared with the *_async entry points)
    void   *host_slot = nullptr; // 4 KiB of mapped pinned host memory: results land here without a D2H copy
    cudaStream_t copy_stream[2] = {nullptr, nullptr}; // H2D / D2H engines of the host-buffer entry points
    cudaStream_t compute_stream = nullptr;
    cudaEvent_t  ev[6]          = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned     func_attr_mask = 0; // which kernels already have their >48 KiB dynamic-smem opt-in on THIS device
};

// cudaFuncSetAttribute is per device: every kernel that opts in to more than 48 KiB of dynamic shared memory does it
// once per DeviceCtx (bit = one of the kAttr* ids), not once per process.
enum FuncAttrBit : unsigned {
    kAttrScanTma0 = 1u << 0, kAttrScanTma1 = 1u << 1, kAttrScanTma2 = 1u << 2, kAttrScanTma3 = 1u << 3,
    kAttrSortHist4 = 1u << 4, kAttrSortPass0 = 1u << 5, kAttrSortPass1 = 1u << 6, kAttrSortPass2 = 1u << 7,
    kAttrSortPass3 = 1u << 8, kAttrSortPass4 = 1u << 9, kAttrSortPass5 = 1u << 10, kAttrSortPeer = 1u << 11,
    kAttrSegSort = 1u << 12, kAttrHist = 1u << 13, kAttrSortPassPk = 1u << 14, kAttrMisc0 = 1u << 15,
    kAttrMisc1 = 1u << 16, kAttrMisc2 = 1u << 17, kAttrMisc3 = 1u << 18, kAttrSortPassDyn = 1u << 19,
};
template <typename K> inline int opt_in_smem(DeviceCtx *c, unsigned bit, K kernel, size_t bytes)
{
    if (c->func_attr_mask & bit)
        return 0;
    B200_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    c->func_attr_mask |= bit;
    return 0;
}

// Returns the context of the current device (created on first use).  nullptr + *status on error.
DeviceCtx *ctx(int *status);
// Grow-only allocation; newly allocated memory is zero-filled (stream-ordered on `stream`).
int ensure(Scratch &s, size_t bytes, cudaStream_t stream, bool *fresh = nullptr);
void release(Scratch &s);

// ------------------------------------------------------------------------------------------------
// PTX helpers
// ------------------------------------------------------------------------------------------------
// 128-bit streaming load: read-only path, do not allocate in L1 (each byte is touched once).
__device__ __forceinline__ uint4 ld_stream_u4(const void *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
// L2 eviction-priority descriptors for ld/st/cp.async.bulk ... .L2::cache_hint (the fixed encodings createpolicy yields
// for fraction 1.0): evict_first = the line is dead once touched, evict_last = a later kernel reads it again
constexpr uint64_t kL2EvictFirst = 0x12F0000000000000ull, kL2EvictLast = 0x14F0000000000000ull;
__device__ __forceinline__ uint4 ld_stream_hint_u4(const void *p, uint64_t policy)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(policy));
    return r;
}
__device__ __forceinline__ uint4 ld_once_hint_u4(const void *p, uint64_t policy)
{
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p), "l"(policy)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_stream_hint_u4(void *p, uint4 v, uint64_t policy)
{
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w), "l"(policy)
                 : "memory");
}
// same without .nc: for kernels that may run in place (the loaded line is later written by this CTA)
__device__ __forceinline__ uint4 ld_once_u4(const void *p)
{
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_stream_u4(void *p, uint4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ unsigned lane_id()
{
    unsigned l;
    asm("mov.u32 %0, %%laneid;" : "=r"(l));
    return l;
}
__device__ __forceinline__ unsigned lanemask_lt()
{
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// relaxed / acquire / release accesses at gpu scope for the inter-CTA descriptors
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(unsigned *p, unsigned v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may
// start once every CTA of its predecessor on the stream has executed launch_dependents (or exited); it must not
// touch anything the predecessor produces before its own griddepcontrol.wait, which returns when the predecessor
// has completed and its memory operations are visible.  Both are no-ops in a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- mbarrier + 1-D bulk async copy (TMA engine; SASS: UBLKCP) -------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t *bar, unsigned parity);
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity)
{
#if B200_MBAR_GUARD // first runs of a new pipeline: a wait that cannot complete gives up after ~2 s (wrong results, no hang)
    const long long t0 = clock64();
    while (!mbar_test(bar, parity))
        if (clock64() - t0 > 4000000000ll)
            return;
    return;
#endif
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "WAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra DONE_%=;\n"
                 "bra WAIT_%=;\n"
                 "DONE_%=:\n"
                 "}" ::"r"(smem_u32(bar)),
                 "r"(parity)
                 : "memory");
}
// non-blocking: has the phase with this parity completed?
__device__ __forceinline__ bool mbar_test(uint64_t *bar, unsigned parity)
{
    unsigned ok;
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                 "selp.u32 %0, 1, 0, p;\n"
                 "}"
                 : "=r"(ok)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
    return ok != 0;
}
// global -> shared, completion signalled on an mbarrier (bytes: multiple of 16, both 16-B aligned)
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, unsigned bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s_hint(void *smem_dst, const void *gsrc, unsigned bytes, uint64_t *bar,
                                              uint64_t policy)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], "
                 "%4;" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
                 : "memory");
}
// global -> L2 only (no shared memory is reserved while the line is in flight)
__device__ __forceinline__ void bulk_prefetch_l2(const void *gsrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
// shared -> global, tracked by the bulk async-group
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_all()
{
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

} // namespace b200
