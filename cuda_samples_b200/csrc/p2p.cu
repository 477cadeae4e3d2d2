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
};
static P2PState g_p2p[kMaxDevices];

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

} // namespace b200

using namespace b200;

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
    s->args.rank  = (unsigned)rank;
    s->args.world = (unsigned)world;
    return 0;
}

extern "C" int b200_p2p_disconnect(void)
{
    int       st = 0;
    P2PState *s  = p2p_state(&st);
    if (!s)
        return st;
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
