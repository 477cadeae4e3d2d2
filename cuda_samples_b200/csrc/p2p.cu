// p2p.cu -- host side of the NVLink peer-window exchange (p2p.cuh) and the stand-alone all-gather kernel.
//
// One process per GPU: every rank creates its window (b200_p2p_window_create -> 64-byte CUDA IPC handle), the
// handles travel over whatever control plane the caller has (torch.distributed all_gather in multi_gpu.py), and
// b200_p2p_connect maps the peers' windows.  After that the fused entry points
//   b200_reduce_partial_exchange_async   (reduce.cu)      reduce -> all-gather of shard sums -> rank-order sum
//   b200_histogram_fused_allreduce_async (histogram.cu)   histogram -> all-reduce of the 320 bins
// and b200_p2p_allgather_async below need no NCCL call on the data path.
#include "p2p.cuh"

#include "internal.h"

namespace b200 {

constexpr int kMaxDevices = 64;
struct P2PState {
    P2PArgs args     = {};
    void   *own      = nullptr;
    void   *opened[kP2PMaxRanks] = {};
    bool    local    = false; // connected by p2p_connect_local: peers' windows are not IPC mappings
};
static P2PState           g_p2p[kMaxDevices];
static unsigned long long g_timeout_ns = kP2PTimeoutNs;

static P2PState *p2p_state(int *st)
{
    int dev = 0;
    *st     = (int)cudaGetDevice(&dev);
    if (*st != 0 || dev < 0 || dev >= kMaxDevices) {
        if (*st == 0)
            *st = (int)cudaErrorInvalidDevice;
        return nullptr;
    }
    return &g_p2p[dev];
}

const P2PArgs *p2p_args()
{
    int       st = 0;
    P2PState *s  = p2p_state(&st);
    return s ? &s->args : nullptr;
}

__global__ void __launch_bounds__(256, 1)
    p2p_allgather_kernel(P2PArgs a, const unsigned *__restrict__ payload, unsigned nwords, unsigned *__restrict__ gathered)
{
    __shared__ unsigned long long s_seq;
    const unsigned               seq = p2p_exchange(a, payload, nwords, &s_seq);
    for (unsigned i = threadIdx.x; i < a.world * nwords; i += blockDim.x) {
        const unsigned r = i / nwords;
        gathered[i]      = p2p_peer_word(a, seq, r, i - r * nwords);
    }
}

// Node-wide barrier on the stream: everything this rank enqueued before it (peer stores included) has completed
// when the kernel starts; it then tells every peer and returns once every peer has said the same.
__global__ void __launch_bounds__(32, 1) p2p_barrier_kernel(P2PArgs a)
{
    __shared__ unsigned long long s_seq;
    __shared__ unsigned           s_token;
    if (threadIdx.x == 0)
        s_token = 1;
    __syncthreads();
    __threadfence_system();
    const unsigned seq = p2p_exchange(a, &s_token, 1, &s_seq);
    if (threadIdx.x < a.world)
        (void)p2p_peer_word(a, seq, threadIdx.x, 0);
    __threadfence_system();
}

// One process, several devices: every device gets a window; with peer access enabled a peer's window is reachable
// through its ordinary device pointer (unified addressing), so "connecting" is just filling in the tables.
static int p2p_connect_local_try(int ndev, const int *devices, bool *fresh)
{
    for (int i = 0; i < ndev; i++) {
        if (devices[i] < 0 || devices[i] >= kMaxDevices)
            return (int)cudaErrorInvalidDevice;
        B200_TRY(cudaSetDevice(devices[i]));
        P2PState &s = g_p2p[devices[i]];
        if (s.args.world != 0)
            return (int)cudaErrorInvalidValue; // already part of a connected set
        for (int j = 0; j < ndev; j++) {
            if (j == i)
                continue;
            int can = 0;
            B200_TRY(cudaDeviceCanAccessPeer(&can, devices[i], devices[j]));
            if (!can)
                return (int)cudaErrorPeerAccessUnsupported;
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return (int)e;
            cudaGetLastError();
        }
        if (!s.own) {
            B200_TRY(cudaMalloc(&s.own, kP2PWindowBytes));
            fresh[i] = true;
            B200_TRY(cudaMemset(s.own, 0, kP2PWindowBytes));
            B200_TRY(cudaDeviceSynchronize());
        }
    }
    return 0;
}

int p2p_connect_local(int ndev, const int *devices)
{
    if (ndev < 1 || ndev > kP2PMaxRanks)
        return (int)cudaErrorInvalidValue;
    int prev = 0;
    B200_TRY(cudaGetDevice(&prev));
    bool      fresh[kP2PMaxRanks] = {};
    const int st                  = p2p_connect_local_try(ndev, devices, fresh);
    if (st != 0) { // give back the windows this call created on the devices it got through before it failed
        for (int i = 0; i < ndev; i++)
            if (fresh[i] && cudaSetDevice(devices[i]) == cudaSuccess) {
                cudaFree(g_p2p[devices[i]].own);
                g_p2p[devices[i]].own = nullptr;
            }
        cudaSetDevice(prev);
        return st;
    }
    for (int i = 0; i < ndev; i++) {
        P2PState &s = g_p2p[devices[i]];
        for (int j = 0; j < ndev; j++)
            s.args.win[j] = (unsigned long long)(uintptr_t)g_p2p[devices[j]].own;
        s.args.rank       = (unsigned)i;
        s.args.world      = (unsigned)ndev;
        s.args.timeout_ns = g_timeout_ns;
        s.local           = true;
    }
    B200_TRY(cudaSetDevice(prev));
    return 0;
}

int p2p_disconnect_local(int ndev, const int *devices)
{
    int prev = 0;
    B200_TRY(cudaGetDevice(&prev));
    for (int i = 0; i < ndev; i++) {
        B200_TRY(cudaSetDevice(devices[i]));
        B200_TRY(cudaDeviceSynchronize());
    }
    for (int i = 0; i < ndev; i++) {
        B200_TRY(cudaSetDevice(devices[i]));
        P2PState &s = g_p2p[devices[i]];
        if (s.own)
            cudaFree(s.own);
        s = P2PState();
    }
    B200_TRY(cudaSetDevice(prev));
    return 0;
}

// ---- NVLink probe: how fast can THIS device push bytes into a peer's memory, and with which kind of store? -----
// mode 0: 4-byte lane stores (what a scatter of u32 runs issues), 1: 16-byte lane stores,
// mode 2: 16 KiB tiles staged in shared memory and sent with cp.async.bulk shared -> (peer) global
__global__ void __launch_bounds__(512, 2)
    p2p_probe_store_kernel(unsigned *__restrict__ dst, const unsigned *__restrict__ src, size_t words, int mode)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
    if (mode == 0) {
        for (size_t i = tid; i < words; i += nthr)
            dst[i] = __ldg(src + i);
    }
    else if (mode == 1) {
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        uint4       *d4 = reinterpret_cast<uint4 *>(dst);
        for (size_t i = tid; i < words / 4; i += nthr)
            d4[i] = ld_stream_u4(s4 + i);
    }
    else {
        __shared__ __align__(128) unsigned tile[4096]; // 16 KiB
        const size_t ntiles = words / 4096;
        for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
            const uint4 *s4 = reinterpret_cast<const uint4 *>(src + t * 4096);
            for (unsigned i = threadIdx.x; i < 1024; i += blockDim.x)
                reinterpret_cast<uint4 *>(tile)[i] = ld_stream_u4(s4 + i);
            fence_proxy_async_smem();
            __syncthreads();
            if (threadIdx.x == 0) {
                bulk_s2g(dst + t * 4096, tile, 16384);
                bulk_commit();
                bulk_wait_read<0>(); // the tile may be overwritten once the engine has read it
            }
            __syncthreads();
        }
        if (threadIdx.x == 0)
            bulk_wait_all<0>();
    }
}

// force the (lazily loaded) exchange kernels onto the current device before any kernel may wait for a peer
int p2p_preload()
{
    cudaFuncAttributes a;
    B200_TRY(cudaFuncGetAttributes(&a, p2p_allgather_kernel));
    B200_TRY(cudaFuncGetAttributes(&a, p2p_barrier_kernel));
    return 0;
}

} // namespace b200

using namespace b200;

// How long a kernel waits for a peer's word before it gives up and counts an error (default 30 s).
// Current device connected (b200_p2p_connect / b200_ctx_create): applies to its connected set -- in the one-process
// mode that is EVERY device of the context, not only the current one -- and to nothing else.  Current device not
// connected: sets the value that sets connected later start with.
extern "C" int b200_p2p_set_timeout_ms(unsigned ms)
{
    const unsigned long long ns = (unsigned long long)(ms ? ms : 1) * 1000000ull;
    int                      st = 0;
    P2PState                *s  = p2p_state(&st);
    if (!s)
        return st;
    if (s->args.world == 0) {
        g_timeout_ns = ns;
        return 0;
    }
    s->args.timeout_ns = ns;
    if (s->local) // the other devices of the same one-process set share this device's window table
        for (int d = 0; d < kMaxDevices; d++)
            if (g_p2p[d].local && g_p2p[d].args.world == s->args.world && g_p2p[d].args.win[0] == s->args.win[0])
                g_p2p[d].args.timeout_ns = ns;
    return 0;
}

// GB/s at which the current device writes `bytes` from d_src (local) to d_dst (a peer-mapped buffer, or local memory
// for comparison); best of reps.  mode 0/1/2: the probe kernel above; mode 3: cudaMemcpyAsync (the copy engine; the
// method of p2pBandwidthLatencyTest.cu:113-146).  This is the measured NVLink roofline of the sort's exchange phase.
extern "C" int b200_p2p_probe_store(void *d_dst, const void *d_src, size_t bytes, int mode, int reps,
                                    double *best_gbs, b200_stream_t stream_)
{
    if (!d_dst || !d_src || bytes < 65536 || mode < 0 || mode > 3 || reps < 1 || !best_gbs)
        return (int)cudaErrorInvalidValue;
    cudaStream_t stream = (cudaStream_t)stream_;
    cudaEvent_t  a, b;
    B200_TRY(cudaEventCreate(&a));
    B200_TRY(cudaEventCreate(&b));
    int sms = 148;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double      best = 0.0;
    cudaError_t e    = cudaSuccess;
    for (int i = 0; i < reps + 1 && e == cudaSuccess; i++) {
        cudaEventRecord(a, stream);
        if (mode == 3)
            e = cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, stream);
        else {
            p2p_probe_store_kernel<<<2 * sms, 512, 0, stream>>>((unsigned *)d_dst, (const unsigned *)d_src, bytes / 4,
                                                                mode);
            e = cudaGetLastError();
        }
        cudaEventRecord(b, stream);
        if (e == cudaSuccess)
            e = cudaEventSynchronize(b);
        float ms = 0.f;
        if (e == cudaSuccess)
            e = cudaEventElapsedTime(&ms, a, b);
        if (i > 0 && ms > 0.f) {
            const double g = (double)bytes / (ms * 1e-3) / 1e9;
            best           = g > best ? g : best;
        }
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    *best_gbs = best;
    return (int)e;
}

extern "C" int b200_p2p_window_create(unsigned char *handle_out /* 64 bytes */)
{
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size is part of the ABI");
    int       st = 0;
    P2PState *s  = p2p_state(&st);
    if (!s)
        return st;
    if (!s->own) {
        B200_TRY(cudaMalloc(&s->own, kP2PWindowBytes));
        B200_TRY(cudaMemset(s->own, 0, kP2PWindowBytes));
        B200_TRY(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    B200_TRY(cudaIpcGetMemHandle(&h, s->own));
    memcpy(handle_out, &h, 64);
    return 0;
}

extern "C" int b200_p2p_connect(int rank, int world, const unsigned char *handles /* world x 64 bytes */)
{
    int       st = 0;
    P2PState *s  = p2p_state(&st);
    if (!s)
        return st;
    if (!s->own || world < 1 || world > kP2PMaxRanks || rank < 0 || rank >= world)
        return (int)cudaErrorInvalidValue;
    for (int r = 0; r < world; r++) {
        if (r == rank) {
            s->args.win[r] = (unsigned long long)(uintptr_t)s->own;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + 64 * r, 64);
        void *p = nullptr;
        B200_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        s->opened[r]   = p;
        s->args.win[r] = (unsigned long long)(uintptr_t)p;
    }
    s->args.rank       = (unsigned)rank;
    s->args.world      = (unsigned)world;
    s->args.timeout_ns = g_timeout_ns;
    return 0;
}

extern "C" int b200_p2p_disconnect(void)
{
    int       st = 0;
    P2PState *s  = p2p_state(&st);
    if (!s)
        return st;
    if (s->local)
        return (int)cudaErrorInvalidValue; // owned by a b200_ctx: b200_ctx_destroy releases it
    B200_TRY(cudaDeviceSynchronize());
    for (int r = 0; r < kP2PMaxRanks; r++)
        if (s->opened[r]) {
            cudaIpcCloseMemHandle(s->opened[r]);
            s->opened[r] = nullptr;
        }
    if (s->own)
        cudaFree(s->own);
    *s = P2PState();
    return 0;
}

extern "C" int b200_p2p_world(void)
{
    const P2PArgs *a = p2p_args();
    return a ? (int)a->world : 0;
}

// number of exchanges on this rank in which a peer did not arrive within the timeout
extern "C" int b200_p2p_errors(unsigned long long *count)
{
    int       st = 0;
    P2PState *s  = p2p_state(&st);
    if (!s || !s->own)
        return s ? (int)cudaErrorInvalidValue : st;
    B200_TRY(cudaMemcpy(count, reinterpret_cast<char *>(s->own) + 8, 8, cudaMemcpyDeviceToHost));
    return 0;
}

// d_gathered[r * nbytes ...) = rank r's d_payload, on every rank.  nbytes: multiple of 4, <= 2048 (kP2PSlotBytes).
extern "C" int b200_p2p_allgather_async(const void *d_payload, size_t nbytes, void *d_gathered, b200_stream_t stream)
{
    const P2PArgs *a = p2p_args();
    if (!a || a->world == 0 || nbytes == 0 || nbytes > kP2PSlotBytes || (nbytes & 3))
        return (int)cudaErrorInvalidValue;
    p2p_allgather_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*a, (const unsigned *)d_payload, (unsigned)(nbytes / 4),
                                                             (unsigned *)d_gathered);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}

// ---- large peer-mapped buffers (the receive buffers of the fused partition + exchange) --------------------------
extern "C" int b200_p2p_buffer_create(size_t bytes, void **d_ptr, unsigned char *handle_out /* 64 bytes */)
{
    B200_TRY(cudaMalloc(d_ptr, bytes ? bytes : 16));
    cudaIpcMemHandle_t h;
    B200_TRY(cudaIpcGetMemHandle(&h, *d_ptr));
    memcpy(handle_out, &h, 64);
    return 0;
}
extern "C" int b200_p2p_buffer_open(const unsigned char *handle, void **d_peer)
{
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    B200_TRY(cudaIpcOpenMemHandle(d_peer, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int b200_p2p_buffer_close(void *d_peer)
{
    B200_TRY(cudaIpcCloseMemHandle(d_peer));
    return 0;
}
extern "C" int b200_p2p_buffer_free(void *d_ptr)
{
    B200_TRY(cudaFree(d_ptr));
    return 0;
}

extern "C" int b200_p2p_barrier_async(b200_stream_t stream)
{
    const P2PArgs *a = p2p_args();
    if (!a || a->world == 0)
        return (int)cudaErrorInvalidValue;
    p2p_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*a);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}
