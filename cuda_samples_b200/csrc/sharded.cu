// sharded.cu -- the sharded primitives behind a plain C ABI: ONE host process drives all GPUs of the node
// (the reference's own multi-GPU idiom: Samples/0_Introduction/simpleMultiGPU/simpleMultiGPU.cu:132-133,157-200 --
// a loop of cudaSetDevice + async launches per device, then a loop of synchronisations).
//
//   b200_ctx_create(ndev, devices)          peer access, one exchange window and one stream per device
//   b200_histogram_sharded / b200_reduce_sharded_{i32,f32,f64} / b200_scan_exclusive_sharded / b200_sort_kv_sharded
//                                           per-device pointer arrays in, int status out
// Nothing here is a new algorithm: every call is the per-device fused kernel of histogram.cu / reduce.cu / scan.cu /
// radix_sort.cu (compute + exchange over the NVLink peer windows of p2p.cuh) issued once per device; what this file
// owns is the sequencing that used to live in Python (multi_gpu.py): splitter selection on the device, receive-buffer
// management, the overflow check, and the rule that makes one-process operation safe --
//   kernels of different devices WAIT for each other, so between the first and the last launch of a step the host
//   must do nothing that can block on a device (no allocation, no lazy module load, no synchronisation).  Therefore
//   b200_ctx_create pre-loads every kernel on every device and each entry point reserves all scratch for all
//   devices BEFORE its first launch.
#include "common.cuh"
#include "internal.h"
#include "p2p.cuh"

#include <new>
#include <vector>

namespace b200 {
int p2p_preload();
int reduce_preload();
int hist_preload();
int sort_preload();
int scan_preload();
int scan_shard_preload();
} // namespace b200

using namespace b200;

struct b200_ctx {
    int          ndev = 0;
    int          dev[kP2PMaxRanks];
    cudaStream_t stream[kP2PMaxRanks];
    // small per-device work areas (device memory): doubles for the reduce / scan exchanges, u32 for the sort
    double   *work[kP2PMaxRanks];   // [0] slot, [1] sums, [2 .. 2+world) gathered
    uint32_t *meta[kP2PMaxRanks];   // [0, w*w) count matrix | [w*w] pairs received | [w*w+1] overflow | splitters | carry
    // receive buffers + ping-pong scratch of the sharded sort (grow-only)
    size_t    capacity = 0;
    // (several devices: one pair-format receive buffer each, 8 bytes per pair -- the partition pass stores (key, value)
    // with ONE peer store and the local sort reads pairs; one device: the ping-pong arrays b200_sort_kv_async wants)
    uint32_t *recvP[kP2PMaxRanks], *bufK[kP2PMaxRanks], *bufV[kP2PMaxRanks];
    // pinned host staging for results
    double   *h_work = nullptr;     // [ndev][4]
    uint32_t *h_meta = nullptr;     // [ndev][w*w + 2]
};

namespace {

struct DeviceGuard {
    int prev = 0;
    DeviceGuard() { cudaGetDevice(&prev); }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

constexpr size_t kMetaWords = kP2PMaxRanks * kP2PMaxRanks + 2 + kP2PMaxRanks + 2;
inline uint32_t *meta_matrix(b200_ctx *c, int i) { return c->meta[i]; }
inline uint32_t *meta_info(b200_ctx *c, int i) { return c->meta[i] + c->ndev * c->ndev; }
inline uint32_t *meta_splitters(b200_ctx *c, int i) { return c->meta[i] + kP2PMaxRanks * kP2PMaxRanks + 2; }
inline uint32_t *meta_carry(b200_ctx *c, int i) { return c->meta[i] + kP2PMaxRanks * kP2PMaxRanks + 2 + kP2PMaxRanks; }

int sync_all(b200_ctx *c)
{
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY(cudaStreamSynchronize(c->stream[i]));
    }
    return 0;
}

// a peer that never arrived makes every result of the step void: fail loudly (sticky count in the windows)
int check_exchange_errors(b200_ctx *c)
{
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        unsigned long long n = 0;
        B200_TRY((cudaError_t)b200_p2p_errors(&n));
        if (n)
            return (int)cudaErrorLaunchTimeout;
    }
    return 0;
}

int free_sort_buffers(b200_ctx *c)
{
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        for (uint32_t **p : {&c->recvP[i], &c->bufK[i], &c->bufV[i]}) {
            if (*p)
                B200_TRY(cudaFree(*p));
            *p = nullptr;
        }
    }
    c->capacity = 0;
    return 0;
}

int reserve_sort(b200_ctx *c, size_t capacity, size_t max_n)
{
    if (capacity > c->capacity) {
        B200_TRY((cudaError_t)free_sort_buffers(c));
        for (int i = 0; i < c->ndev; i++) {
            B200_TRY(cudaSetDevice(c->dev[i]));
            if (c->ndev == 1) { // one device: nothing is exchanged
                B200_TRY(cudaMalloc(&c->bufK[i], capacity * sizeof(uint32_t)));
                B200_TRY(cudaMalloc(&c->bufV[i], capacity * sizeof(uint32_t)));
            }
            else
                B200_TRY(cudaMalloc(&c->recvP[i], capacity * 2 * sizeof(uint32_t)));
        }
        c->capacity = capacity;
    }
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_sort_reserve(capacity > max_n ? capacity : max_n));
    }
    return 0;
}

} // namespace

extern "C" int b200_ctx_create(int ndev, const int *devices, b200_ctx **out)
{
    if (!out || !devices || ndev < 1 || ndev > kP2PMaxRanks)
        return (int)cudaErrorInvalidValue;
    for (int i = 0; i < ndev; i++)
        for (int j = 0; j < i; j++)
            if (devices[i] == devices[j])
                return (int)cudaErrorInvalidValue;
    DeviceGuard guard;
    b200_ctx   *c = new (std::nothrow) b200_ctx();
    if (!c)
        return (int)cudaErrorMemoryAllocation;
    c->ndev = ndev;
    for (int i = 0; i < kP2PMaxRanks; i++) {
        c->dev[i]    = i < ndev ? devices[i] : -1;
        c->stream[i] = nullptr;
        c->work[i]   = nullptr;
        c->meta[i]   = nullptr;
        c->recvP[i] = c->bufK[i] = c->bufV[i] = nullptr;
    }
    int st = p2p_connect_local(ndev, devices);
    if (st) {
        delete c;
        return st;
    }
    auto setup = [&]() -> int {
        for (int i = 0; i < ndev; i++) {
            B200_TRY(cudaSetDevice(c->dev[i]));
            B200_TRY(cudaStreamCreateWithFlags(&c->stream[i], cudaStreamNonBlocking));
            B200_TRY(cudaMalloc(&c->work[i], (4 + kP2PMaxRanks) * sizeof(double)));
            B200_TRY(cudaMalloc(&c->meta[i], kMetaWords * sizeof(uint32_t)));
            B200_TRY(cudaMemset(c->meta[i], 0, kMetaWords * sizeof(uint32_t)));
            // every kernel a sharded step may launch is loaded now, while no kernel is waiting for a peer
            B200_TRY((cudaError_t)p2p_preload());
            B200_TRY((cudaError_t)reduce_preload());
            B200_TRY((cudaError_t)hist_preload());
            B200_TRY((cudaError_t)sort_preload());
            B200_TRY((cudaError_t)scan_preload());
            B200_TRY((cudaError_t)scan_shard_preload());
            B200_TRY(cudaDeviceSynchronize());
        }
        B200_TRY(cudaHostAlloc(&c->h_work, ndev * 4 * sizeof(double), cudaHostAllocPortable));
        B200_TRY(cudaHostAlloc(&c->h_meta, ndev * (kP2PMaxRanks * kP2PMaxRanks + 2) * sizeof(uint32_t),
                               cudaHostAllocPortable));
        return 0;
    };
    st = setup();
    if (st) {
        b200_ctx_destroy(c);
        return st;
    }
    *out = c;
    return 0;
}

extern "C" int b200_ctx_destroy(b200_ctx *c)
{
    if (!c)
        return 0;
    DeviceGuard guard;
    for (int i = 0; i < c->ndev; i++) {
        cudaSetDevice(c->dev[i]);
        cudaDeviceSynchronize();
    }
    free_sort_buffers(c);
    for (int i = 0; i < c->ndev; i++) {
        cudaSetDevice(c->dev[i]);
        if (c->stream[i])
            cudaStreamDestroy(c->stream[i]);
        if (c->work[i])
            cudaFree(c->work[i]);
        if (c->meta[i])
            cudaFree(c->meta[i]);
    }
    if (c->h_work)
        cudaFreeHost(c->h_work);
    if (c->h_meta)
        cudaFreeHost(c->h_meta);
    int st = p2p_disconnect_local(c->ndev, c->dev);
    delete c;
    return st;
}

extern "C" int b200_ctx_num_devices(const b200_ctx *c) { return c ? c->ndev : 0; }

extern "C" int b200_ctx_stream(b200_ctx *c, int i, b200_stream_t *stream)
{
    if (!c || i < 0 || i >= c->ndev || !stream)
        return (int)cudaErrorInvalidValue;
    *stream = (b200_stream_t)c->stream[i];
    return 0;
}

// waits for everything enqueued on the context's streams; cudaErrorLaunchTimeout if an exchange lost a peer
extern "C" int b200_ctx_synchronize(b200_ctx *c)
{
    if (!c)
        return (int)cudaErrorInvalidValue;
    DeviceGuard guard;
    B200_TRY((cudaError_t)sync_all(c));
    return check_exchange_errors(c);
}

// histogram64 + histogram256 of the concatenation of all shards; EVERY device's d_hist256[i] / d_hist64[i] receives
// the global bins.  Asynchronous (context streams): b200_ctx_synchronize before reading.
extern "C" int b200_histogram_sharded(b200_ctx *c, uint32_t *const *d_hist256, uint32_t *const *d_hist64,
                                      const void *const *d_data, const size_t *byteCount)
{
    if (!c || !d_hist256 || !d_hist64 || !d_data || !byteCount)
        return (int)cudaErrorInvalidValue;
    DeviceGuard guard;
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_histogram_fused_allreduce_async(d_hist256[i], d_hist64[i], d_data[i], byteCount[i],
                                                                   (b200_stream_t)c->stream[i]));
    }
    return 0;
}

static int reduce_sharded(b200_ctx *c, int kind, const void *const *d_in, const size_t *n, double *total)
{
    if (!c || !d_in || !n || !total)
        return (int)cudaErrorInvalidValue;
    DeviceGuard guard;
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        double *w = c->work[i];
        B200_TRY((cudaError_t)b200_reduce_partial_exchange_async(kind, d_in[i], n[i], w, 0, 1, w + 2, w + 1, nullptr,
                                                                 (b200_stream_t)c->stream[i]));
        B200_TRY(cudaMemcpyAsync(c->h_work + 4 * i, w + 1, sizeof(double), cudaMemcpyDeviceToHost, c->stream[i]));
    }
    B200_TRY((cudaError_t)sync_all(c));
    B200_TRY((cudaError_t)check_exchange_errors(c));
    for (int i = 1; i < c->ndev; i++) // every device added the same partials in the same order
        if (c->h_work[4 * i] != c->h_work[0])
            return (int)cudaErrorUnknown;
    *total = c->h_work[0];
    return 0;
}

// Sum over all shards (blocking).  int: wrapping 32-bit sum like reduceCPU<int>; float / double: the shard sums are
// added in device order in fp64 (deterministic), then rounded once.
extern "C" int b200_reduce_sharded_i32(b200_ctx *c, const int32_t *const *d_in, const size_t *n, int32_t *h_result)
{
    double t  = 0.0;
    int    st = reduce_sharded(c, 0, (const void *const *)d_in, n, &t);
    if (!st && h_result)
        *h_result = (int32_t)(uint32_t)(long long)t; // exact integer in the double: wrap to 32 bits
    return st;
}
extern "C" int b200_reduce_sharded_f32(b200_ctx *c, const float *const *d_in, const size_t *n, float *h_result)
{
    double t  = 0.0;
    int    st = reduce_sharded(c, 1, (const void *const *)d_in, n, &t);
    if (!st && h_result)
        *h_result = (float)t;
    return st;
}
extern "C" int b200_reduce_sharded_f64(b200_ctx *c, const double *const *d_in, const size_t *n, double *h_result)
{
    double t  = 0.0;
    int    st = reduce_sharded(c, 2, (const void *const *)d_in, n, &t);
    if (!st && h_result)
        *h_result = t;
    return st;
}

// Exclusive scan of ONE array split into contiguous shards (device i holds the i-th slice, n[i] elements).
// Asynchronous.  Two kernels per device and no collective call: shard total + exchange + carry, then the chained scan
// seeded with the carry.
extern "C" int b200_scan_exclusive_sharded(b200_ctx *c, uint32_t *const *d_dst, const uint32_t *const *d_src,
                                           const size_t *n)
{
    if (!c || !d_dst || !d_src || !n)
        return (int)cudaErrorInvalidValue;
    DeviceGuard guard;
    for (int i = 0; i < c->ndev; i++) { // all scratch first (growing it synchronises the device)
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_scan_reserve(n[i]));
    }
    // per device: tile sums + exchange of the shard totals (its last CTA waits for the peers' totals), then the scan of
    // independent tiles -- both launches of a device are enqueued before the next device's: nothing here blocks the host
    for (int i = 0; i < c->ndev; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_scan_exclusive_shard_async(d_dst[i], d_src[i], n[i], c->work[i], meta_carry(c, i),
                                                              (b200_stream_t)c->stream[i]));
    }
    return 0;
}

// Stable key-value sort of the concatenation of all shards (sample sort).  Device i ends up with the i-th key range
// (ascending ranges for sortDir != 0, descending for 0) in d_dstKey[i] / d_dstVal[i], h_nOut[i] pairs of it.
// dstCapacity: pairs each destination array can hold.  Blocking.  Returns cudaErrorMemoryAllocation when some device
// would receive more than dstCapacity pairs (h_nOut then holds the sizes that are needed; nothing was moved).
extern "C" int b200_sort_kv_sharded(b200_ctx *c, uint32_t *const *d_dstKey, uint32_t *const *d_dstVal,
                                    size_t dstCapacity, const uint32_t *const *d_srcKey,
                                    const uint32_t *const *d_srcVal, const size_t *n, uint32_t sortDir, size_t *h_nOut)
{
    if (!c || !d_dstKey || !d_dstVal || !d_srcKey || !d_srcVal || !n || !h_nOut || dstCapacity >= (1ull << 30))
        return (int)cudaErrorInvalidValue;
    DeviceGuard  guard;
    const int    w     = c->ndev;
    size_t       max_n = 0;
    for (int i = 0; i < w; i++)
        max_n = n[i] > max_n ? n[i] : max_n;
    if (max_n >= (1ull << 30))
        return (int)cudaErrorInvalidValue;
    if (w == 1) {
        if (n[0] > dstCapacity) {
            h_nOut[0] = n[0];
            return (int)cudaErrorMemoryAllocation;
        }
        B200_TRY(cudaSetDevice(c->dev[0]));
        B200_TRY((cudaError_t)reserve_sort(c, n[0] ? n[0] : 1, n[0]));
        B200_TRY((cudaError_t)b200_sort_kv_async(d_dstKey[0], d_dstVal[0], c->bufK[0], c->bufV[0], d_srcKey[0],
                                                 d_srcVal[0], n[0], sortDir, (b200_stream_t)c->stream[0]));
        B200_TRY(cudaStreamSynchronize(c->stream[0]));
        h_nOut[0] = n[0];
        return 0;
    }
    // everything that may block happens before the first kernel that waits for a peer
    B200_TRY((cudaError_t)reserve_sort(c, dstCapacity ? dstCapacity : 1, max_n));
    const size_t cap = dstCapacity; // the receive buffers hold exactly what the destinations can take
    // 1. splitters, on the device (one exchange), same on every device
    for (int i = 0; i < w; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_sort_splitters_async(d_srcKey[i], n[i], sortDir, meta_splitters(c, i),
                                                        (b200_stream_t)c->stream[i]));
    }
    // 2. count + exchange of counts, stable partition whose bucket runs are stored straight into the destination
    //    device's receive buffer over NVLink, node-wide barrier
    void *rp[kP2PMaxRanks];
    for (int i = 0; i < w; i++)
        rp[i] = c->recvP[i];
    for (int i = 0; i < w; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_partition_exchange_kv_async(d_srcKey[i], d_srcVal[i], n[i], meta_splitters(c, i),
                                                               sortDir, rp, nullptr, cap, meta_matrix(c, i), meta_info(c, i),
                                                               (b200_stream_t)c->stream[i]));
    }
    // 3. local stable sort of what arrived, count read on the device
    for (int i = 0; i < w; i++) {
        B200_TRY(cudaSetDevice(c->dev[i]));
        B200_TRY((cudaError_t)b200_sort_kv_dyn_async(d_dstKey[i], d_dstVal[i], nullptr, nullptr, c->recvP[i], nullptr,
                                                     cap, meta_info(c, i), sortDir, (b200_stream_t)c->stream[i]));
        B200_TRY(cudaMemcpyAsync(c->h_meta + (size_t)i * (kP2PMaxRanks * kP2PMaxRanks + 2), c->meta[i],
                                 (w * w + 2) * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream[i]));
    }
    B200_TRY((cudaError_t)sync_all(c));
    B200_TRY((cudaError_t)check_exchange_errors(c));
    bool over = false;
    for (int i = 0; i < w; i++) {
        const uint32_t *m = c->h_meta + (size_t)i * (kP2PMaxRanks * kP2PMaxRanks + 2);
        size_t          col = 0;
        for (int r = 0; r < w; r++)
            col += m[r * w + i];
        h_nOut[i] = col; // what device i receives (from the matrix: valid whether or not anything was sent)
        over |= m[w * w + 1] != 0;
    }
    return over ? (int)cudaErrorMemoryAllocation : 0;
}
