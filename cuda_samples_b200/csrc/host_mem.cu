// host_mem.cu -- pinned host memory placed on the GPU's own NUMA node, and the PCIe probe the end-to-end numbers
// are reported against.
//
// The host-buffer entry points (host_path.cu) are bounded by the host<->device link, and on a multi-socket box that
// link runs at half speed when the pinned pages live on the other socket (round 1: 45 GB/s on a one-socket slice,
// 25 GB/s for the same code on an 8-GPU two-socket box).  cudaHostAlloc places pages by the calling thread's
// first-touch policy, i.e. wherever the scheduler happened to run it.  b200_host_alloc therefore
//   1. finds the NUMA node of the current device (sysfs: /sys/bus/pci/devices/<bdf>/numa_node; when a virtualised
//      PCI topology does not say, it MEASURES it: a 64 MiB pinned H2D copy from every node, fastest wins),
//   2. maps anonymous memory, asks for transparent huge pages, binds it to that node with mbind(MPOL_PREFERRED)
//      (raw syscall: no libnuma in this image); when mbind is not permitted it falls back to first-touching the
//      pages from a thread pinned to that node's CPUs (if the process is allowed to run there),
//   3. touches every page and registers the range with cudaHostRegister (mapped + portable).
// b200_host_alloc_info reports where the pages actually are (move_pages query), so bench.py can print it.
#include <chrono>
#include "common.cuh"
#include "internal.h"

#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include <errno.h>
#include <sched.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

namespace b200 {

struct HostBlock {
    size_t bytes;     // mapped length
    int    gpu_node;  // NUMA node of the device it was allocated for (-1 unknown)
    int    how;       // 0 no placement, 1 mbind, 2 first touch from the node's CPUs
};
static std::mutex                             g_host_mutex;
static std::unordered_map<void *, HostBlock>  g_host_blocks;

static int read_int_file(const char *path, int fallback)
{
    FILE *f = fopen(path, "r");
    if (!f)
        return fallback;
    int v = fallback;
    if (fscanf(f, "%d", &v) != 1)
        v = fallback;
    fclose(f);
    return v;
}

// NUMA node the current CUDA device hangs off, -1 when the platform does not say
static int device_numa_node()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess)
        return -1;
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof(bdf), dev) != cudaSuccess)
        return -1;
    for (char *p = bdf; *p; p++)
        if (*p >= 'A' && *p <= 'F')
            *p = (char)(*p - 'A' + 'a');
    char path[128];
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bdf);
    return read_int_file(path, -1);
}

// CPUs of a node that this process may run on
static int node_cpus(int node, cpu_set_t *out)
{
    CPU_ZERO(out);
    char path[128];
    snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
    FILE *f = fopen(path, "r");
    if (!f)
        return 0;
    cpu_set_t allowed;
    CPU_ZERO(&allowed);
    sched_getaffinity(0, sizeof(allowed), &allowed);
    int  n = 0, a = 0, b = 0;
    char sep = 0;
    while (fscanf(f, "%d", &a) == 1) {
        b = a;
        int c = fgetc(f);
        if (c == '-') {
            if (fscanf(f, "%d", &b) != 1)
                b = a;
            c = fgetc(f);
        }
        for (int i = a; i <= b && i < CPU_SETSIZE; i++)
            if (CPU_ISSET(i, &allowed)) {
                CPU_SET(i, out);
                n++;
            }
        sep = (char)c;
        if (sep != ',')
            break;
    }
    fclose(f);
    return n;
}

static void touch_range(char *p, size_t bytes, const cpu_set_t *cpus)
{
    if (cpus)
        sched_setaffinity(0, sizeof(*cpus), cpus); // this thread only
    const size_t page = 4096;
    for (size_t o = 0; o < bytes; o += page)
        p[o] = 0;
}

static int page_node(void *p)
{
    void *pages[1] = {(void *)((uintptr_t)p & ~(uintptr_t)4095)};
    int   status[1] = {-1};
    long  r = syscall(SYS_move_pages, 0, 1ul, pages, (const int *)nullptr, status, 0);
    return r == 0 ? status[0] : -1;
}

static long bind_to_node(void *p, size_t len, int node)
{
    unsigned long mask[16] = {0};
    mask[node / 64] |= 1ul << (node % 64);
    // MPOL_PREFERRED = 1: take the node when it has memory, do not fail when it has not
    return syscall(SYS_mbind, p, len, 1, mask, 1025ul, 0u);
}

static int count_nodes()
{
    int n = 0;
    for (int i = 0; i < 64; i++) {
        char path[96];
        snprintf(path, sizeof(path), "/sys/devices/system/node/node%d", i);
        if (access(path, F_OK) == 0)
            n = i + 1;
    }
    return n;
}

// When sysfs does not say which node the GPU hangs off (virtualised PCI topology) but the box has several, measure
// it: 64 MiB pinned on each node, one warm-up + one timed H2D copy each, fastest node wins.  Cached per device.
static int g_probed_node[64];
static bool g_probed[64];
static int probe_best_node(int dev)
{
    if (dev >= 0 && dev < 64 && g_probed[dev])
        return g_probed_node[dev];
    const int nodes = count_nodes();
    int       best = -1;
    if (nodes > 1) {
        const size_t len = 64u << 20;
        void        *d   = nullptr;
        cudaEvent_t  a = nullptr, b = nullptr;
        if (cudaMalloc(&d, len) == cudaSuccess && cudaEventCreate(&a) == cudaSuccess &&
            cudaEventCreate(&b) == cudaSuccess) {
            float best_ms = 1e30f;
            for (int node = 0; node < nodes; node++) {
                void *p = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
                if (p == MAP_FAILED)
                    continue;
                if (bind_to_node(p, len, node) == 0) {
                    memset(p, 1, len);
                    if (page_node(p) == node && cudaHostRegister(p, len, cudaHostRegisterDefault) == cudaSuccess) {
                        float ms = 1e30f;
                        for (int r = 0; r < 2; r++) {
                            cudaEventRecord(a, 0);
                            cudaMemcpyAsync(d, p, len, cudaMemcpyHostToDevice, 0);
                            cudaEventRecord(b, 0);
                            if (cudaEventSynchronize(b) == cudaSuccess)
                                cudaEventElapsedTime(&ms, a, b);
                        }
                        if (ms < best_ms) {
                            best_ms = ms;
                            best    = node;
                        }
                        cudaHostUnregister(p);
                    }
                }
                munmap(p, len);
            }
        }
        if (a)
            cudaEventDestroy(a);
        if (b)
            cudaEventDestroy(b);
        if (d)
            cudaFree(d);
        cudaGetLastError();
    }
    if (dev >= 0 && dev < 64) {
        g_probed[dev]      = true;
        g_probed_node[dev] = best;
    }
    return best;
}

} // namespace b200

using namespace b200;

// Pinned host memory for the *_host entry points, on the NUMA node of the CURRENT device when the platform allows.
extern "C" int b200_host_alloc(void **h_ptr, size_t bytes)
{
    if (!h_ptr)
        return (int)cudaErrorInvalidValue;
    const size_t huge = 2u << 20;
    const size_t len  = ((bytes ? bytes : 1) + huge - 1) / huge * huge;
    void        *p    = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED)
        return (int)cudaErrorMemoryAllocation;
    madvise(p, len, MADV_HUGEPAGE);
    int       node = device_numa_node();
    int       how  = 0;
    cpu_set_t cpus;
    int       ncpu = 0;
    if (node < 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        node = probe_best_node(dev); // -1 on a one-node box: nothing to choose
    }
    if (node >= 0 && node < 1024) {
        if (bind_to_node(p, len, node) == 0)
            how = 1;
        else {
            ncpu = node_cpus(node, &cpus);
            if (ncpu > 0)
                how = 2;
        }
    }
    {
        // first touch in parallel (the page faults of a GiB take ~0.3 s on one thread)
        const int                nthreads = 8;
        std::vector<std::thread> pool;
        const size_t             per = (len / nthreads + huge - 1) / huge * huge;
        for (int t = 0; t < nthreads; t++) {
            const size_t lo = (size_t)t * per;
            if (lo >= len)
                break;
            const size_t n = len - lo < per ? len - lo : per;
            pool.emplace_back(touch_range, (char *)p + lo, n, how == 2 ? &cpus : nullptr);
        }
        for (auto &t : pool)
            t.join();
    }
    cudaError_t e = cudaHostRegister(p, len, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e != cudaSuccess) {
        munmap(p, len);
        return (int)e;
    }
    {
        std::lock_guard<std::mutex> lock(g_host_mutex);
        g_host_blocks[p] = HostBlock{len, node, how};
    }
    *h_ptr = p;
    return 0;
}

extern "C" int b200_host_free(void *h_ptr)
{
    if (!h_ptr)
        return 0;
    HostBlock b{0, -1, 0};
    {
        std::lock_guard<std::mutex> lock(g_host_mutex);
        auto                        it = g_host_blocks.find(h_ptr);
        if (it == g_host_blocks.end())
            return (int)cudaErrorInvalidValue;
        b = it->second;
        g_host_blocks.erase(it);
    }
    cudaError_t e = cudaHostUnregister(h_ptr);
    munmap(h_ptr, b.bytes);
    return (int)e;
}

// Where a b200_host_alloc block is: NUMA node of the device it was made for, node its first and last pages are on
// (-1: unknown), placement method (0 none, 1 mbind, 2 first touch on the node's CPUs).
extern "C" int b200_host_alloc_info(const void *h_ptr, int *gpu_node, int *mem_node_first, int *mem_node_last,
                                    int *method)
{
    HostBlock b{0, -1, 0};
    {
        std::lock_guard<std::mutex> lock(g_host_mutex);
        auto                        it = g_host_blocks.find(const_cast<void *>(h_ptr));
        if (it == g_host_blocks.end())
            return (int)cudaErrorInvalidValue;
        b = it->second;
    }
    if (gpu_node)
        *gpu_node = b.gpu_node;
    if (mem_node_first)
        *mem_node_first = page_node(const_cast<void *>(h_ptr));
    if (mem_node_last)
        *mem_node_last = page_node((char *)const_cast<void *>(h_ptr) + b.bytes - 4096);
    if (method)
        *method = b.how;
    return 0;
}

// The PCIe roofline of the host-buffer entry points, measured in the same run: plain cudaMemcpyAsync of `bytes`
// from h_src (pinned) into a device buffer, best of `reps`, GB/s.  (The method of the reference's own
// p2pBandwidthLatencyTest / bandwidthTest: one large copy between events.)
extern "C" int b200_host_probe_h2d(const void *h_src, size_t bytes, int reps, double *best_gbs)
{
    if (!h_src || !bytes || reps < 1 || !best_gbs)
        return (int)cudaErrorInvalidValue;
    void *d = nullptr;
    B200_TRY(cudaMalloc(&d, bytes));
    cudaStream_t s;
    cudaEvent_t  a, b;
    B200_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    B200_TRY(cudaEventCreate(&a));
    B200_TRY(cudaEventCreate(&b));
    double      best = 0.0;
    cudaError_t e    = cudaSuccess;
    for (int i = 0; i < reps + 1 && e == cudaSuccess; i++) { // first copy is a warm-up
        cudaEventRecord(a, s);
        e = cudaMemcpyAsync(d, h_src, bytes, cudaMemcpyHostToDevice, s);
        cudaEventRecord(b, s);
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
    cudaStreamDestroy(s);
    cudaFree(d);
    *best_gbs = best;
    return (int)e;
}

// reps back-to-back pinned H2D copies into a caller-provided device buffer; *seconds = host time from the first
// cudaMemcpyAsync to the completion of the last (no allocation, stream or event creation inside the interval).  Ranks
// that call this between the same two barriers measure what the host feeds all their GPUs AT ONCE.
extern "C" int b200_host_probe_h2d_sustained(void *d_dst, const void *h_src, size_t bytes, int reps, double *seconds)
{
    if (!d_dst || !h_src || !bytes || reps < 1 || !seconds)
        return (int)cudaErrorInvalidValue;
    cudaStream_t s;
    B200_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaError_t e = cudaMemcpyAsync(d_dst, h_src, bytes < (1u << 20) ? bytes : (1u << 20), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(s); // warm the path
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < reps && e == cudaSuccess; i++)
        e = cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(s);
    *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    cudaStreamDestroy(s);
    return (int)e;
}
