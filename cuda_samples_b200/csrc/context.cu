// context.cu -- per-device scratch, launch accounting and the small utility exports of b200prim.h.
#include "common.cuh"
#include "internal.h"

#include <atomic>
#include <mutex>

namespace b200 {

static constexpr int kMaxDevices = 64;
static DeviceCtx      g_ctx[kMaxDevices];
static std::mutex     g_mutex;
std::atomic<unsigned long long> g_launches{0};

DeviceCtx *ctx(int *status)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        if (status)
            *status = (int)(e != cudaSuccess ? e : cudaErrorInvalidDevice);
        return nullptr;
    }
    DeviceCtx &c = g_ctx[dev];
    if (c.device < 0) {
        std::lock_guard<std::mutex> lock(g_mutex);
        if (c.device < 0) {
            int major = 0;
            e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
            if (e == cudaSuccess)
                e = cudaDeviceGetAttribute(&c.num_sms, cudaDevAttrMultiProcessorCount, dev);
            if (e == cudaSuccess)
                e = cudaDeviceGetAttribute(&c.smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
            if (e == cudaSuccess && major != 10)
                e = cudaErrorNoKernelImageForDevice; // sm_100a only: fail loudly, there is no fallback
            if (e != cudaSuccess) {
                if (status)
                    *status = (int)e;
                return nullptr;
            }
            c.device = dev;
        }
    }
    if (status)
        *status = 0;
    return &c;
}

int ensure(Scratch &s, size_t bytes, cudaStream_t stream, bool *fresh)
{
    if (fresh)
        *fresh = false;
    if (s.bytes >= bytes)
        return 0;
    // Growing is rare (sizes settle after the first call; init*() / b200_*_reserve pre-size) and it is the one place
    // where an *_async entry point is not asynchronous: cudaFree synchronises the device (work enqueued on the old
    // block has finished before it goes away) and the zero-fill is waited for, so that launches on ANY stream --
    // also cudaStreamNonBlocking ones that do not order against the stream passed here -- see zeroed tickets and
    // epochs.  That cannot be done inside a stream capture: fail with a status instead of breaking the capture.
    if (stream != nullptr && stream != cudaStreamLegacy && stream != cudaStreamPerThread) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone)
            return (int)cudaErrorStreamCaptureUnsupported;
    }
    if (s.ptr)
        B200_TRY(cudaFree(s.ptr));
    s.ptr   = nullptr;
    s.bytes = 0;
    size_t want = bytes < 4096 ? 4096 : bytes;
    B200_TRY(cudaMalloc(&s.ptr, want));
    B200_TRY(cudaMemsetAsync(s.ptr, 0, want, stream));
    B200_TRY(cudaStreamSynchronize(stream));
    s.bytes = want;
    if (fresh)
        *fresh = true;
    return 0;
}

struct PdlNotes {
    uintptr_t lo[8] = {}, hi[8] = {};
    unsigned  next = 0;
};
static PdlNotes g_pdl_notes[kMaxDevices];

void pdl_note_output(int device, const void *p, size_t bytes)
{
    if (!p || !bytes || device < 0 || device >= kMaxDevices)
        return;
    PdlNotes &n      = g_pdl_notes[device];
    n.lo[n.next & 7] = (uintptr_t)p;
    n.hi[n.next & 7] = (uintptr_t)p + bytes;
    n.next++;
}

bool pdl_early_read_ok(int device, const void *p, size_t bytes)
{
    if (!p || !bytes)
        return true;
    if (device < 0 || device >= kMaxDevices)
        return false;
    const PdlNotes &n  = g_pdl_notes[device];
    const uintptr_t lo = (uintptr_t)p, hi = lo + bytes;
    for (int i = 0; i < 8; i++)
        if (lo < n.hi[i] && n.lo[i] < hi)
            return false;
    return true;
}

void release(Scratch &s)
{
    if (s.ptr)
        cudaFree(s.ptr);
    s.ptr   = nullptr;
    s.bytes = 0;
}

} // namespace b200

using namespace b200;

extern "C" int b200_version(void) { return 100; }

extern "C" const char *b200_error_string(int status)
{
    return status == 0 ? "success" : cudaGetErrorString((cudaError_t)status);
}

extern "C" int b200_device_info(int *num_sms, int *smem_optin, int *cc_major, int *cc_minor)
{
    int dev = 0, v = 0;
    B200_TRY(cudaGetDevice(&dev));
    if (num_sms) {
        B200_TRY(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
        *num_sms = v;
    }
    if (smem_optin) {
        B200_TRY(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        *smem_optin = v;
    }
    if (cc_major) {
        B200_TRY(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev));
        *cc_major = v;
    }
    if (cc_minor) {
        B200_TRY(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev));
        *cc_minor = v;
    }
    return 0;
}

extern "C" unsigned long long b200_launch_count(void) { return g_launches.load(); }

// what close*() of one primitive gives back: only that primitive's scratch (which: 0 histogram, 1 scan, 2 sort,
// 3 reduce); the other primitives' state stays
extern "C" int b200_release_scratch_of(int which)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY(cudaDeviceSynchronize());
    switch (which) {
    case 0: release(c->hist); break;
    case 1:
        release(c->scan);
        release(c->scan_tiles);
        break;
    case 2:
        release(c->sort);
        release(c->sort_pairs);
        break;
    case 3: release(c->reduce); break;
    default: return (int)cudaErrorInvalidValue;
    }
    return 0;
}

extern "C" int b200_release_scratch(void)
{
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    B200_TRY(cudaDeviceSynchronize());
    release(c->reduce);
    release(c->hist);
    release(c->scan);
    release(c->scan_tiles);
    release(c->sort);
    release(c->sort_pairs);
    release(c->image);
    release(c->stage[0]);
    release(c->stage[1]);
    release(c->host_work);
    return 0;
}
