/*
 * b200prim.h -- extended C ABI of libb200prim.so (sm_100a only).
 *
 * The drop-in entry points (same names and signatures as the cuda-samples `extern "C"` host
 * wrappers) are declared in histogram_common.h, scan_common.h, mergeSort_common.h and reduction.h
 * next to this file.  This header adds what the samples' API cannot express:
 *   - an explicit stream (the samples launch on the legacy default stream only),
 *   - size_t lengths (histogram byteCount is `uint` in histogram/histogram_common.h:80-82),
 *   - status codes instead of exit() (Common/helper_cuda.h:584-616 aborts),
 *   - host-buffer variants that own the host<->device copies (what a CPU caller of
 *     histogram256CPU / reduceCPU / scanExclusiveHost / mergeSortHost would switch to),
 *   - building blocks of the multi-GPU path (shard carry for scan, bucket partition for sort).
 *
 * Plain pointers and sizes only.  All `d_*` pointers are device pointers of the CURRENT device,
 * `h_*` pointers are host pointers (pinned memory gives full PCIe speed; pageable works).
 * Every function returns 0 on success or a cudaError_t value.  `*_async` functions only enqueue
 * work on `stream`; they never synchronise.  The library keeps grow-only per-device scratch, so
 * calls on one device must be stream-ordered with each other (same contract as the samples'
 * file-static scratch: histogram256.cu:133, scan.cu:182, mergeSort.cu:452).
 * Overlap between consecutive calls: the histogram kernels, and a scan given a d_carry, are launched with programmatic
 * stream serialization and start reading their INPUT while the library kernel in front of them on the stream is still
 * draining.  The library checks the input range against the outputs of its own recent launches on the device and
 * launches the ordinary, fully ordered way when they overlap, so chains such as reduce -> histogram of its output
 * buffer or histogram -> scan of the bins are safe; kernels that are not the library's never release early, so they
 * are always waited for in full.
 */
#ifndef B200PRIM_H
#define B200PRIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st *b200_stream_t; /* == cudaStream_t */

/* library / device ---------------------------------------------------------------------------- */
int         b200_version(void);
const char *b200_error_string(int status);
int         b200_device_info(int *num_sms, int *smem_per_block_optin, int *cc_major, int *cc_minor);
/* number of kernels this library has launched in this process (for bench.py's gpu_launches) */
unsigned long long b200_launch_count(void);
/* free all per-device scratch of the current device (the close*() calls do this too) */
int b200_release_scratch(void);

/* histogram: replaces histogram256()/histogram64() (histogram256.cu:145-155, histogram64.cu:186-200).
 * d_hist is OVERWRITTEN with u32 counts (wrap mod 2^32); any byteCount >= 0, any alignment.
 * b200_histogram_fused_async produces both histograms in ONE pass over the data
 * (hist64[v] = sum of hist256[4v..4v+3], exact: histogram_gold.cpp:41-44 vs :57-60). */
int b200_histogram256_async(uint32_t *d_hist256, const void *d_data, size_t byteCount, b200_stream_t stream);
int b200_histogram64_async(uint32_t *d_hist64, const void *d_data, size_t byteCount, b200_stream_t stream);
int b200_histogram_fused_async(uint32_t *d_hist256, uint32_t *d_hist64, const void *d_data, size_t byteCount,
                               b200_stream_t stream);

/* reduce: replaces reduce<T>() (reduction_kernel.cu:633-982) + the caller's multi-pass finish
 * (reduction.cpp:292-318).  Single pass: d_out[0] = sum, d_out[1..out_len) = 0, so callers that
 * go on to add up `blocks` partials still get the right answer.  int: wrapping 32-bit sum.
 * float/double: deterministic (fixed grid, fixed-order finish, no floating-point atomics);
 * |gpu - reduceCPU| <= 1e-5 (fp32) / 1e-12 (fp64) relative. n >= 0 (n == 0 gives 0). */
int b200_reduce_i32_async(const int32_t *d_in, size_t n, int32_t *d_out, size_t out_len, b200_stream_t stream);
int b200_reduce_f32_async(const float *d_in, size_t n, float *d_out, size_t out_len, b200_stream_t stream);
int b200_reduce_f64_async(const double *d_in, size_t n, double *d_out, size_t out_len, b200_stream_t stream);
/* multi-GPU building block: the shard's sum in its own type (kind 0 i32 wrapping, 1 f32, 2 f64), widened exactly
 * to a double in *d_slot, so partials of mixed types travel in ONE all_gather buffer. */
int b200_reduce_partial_async(int kind, const void *d_in, size_t n, double *d_slot, b200_stream_t stream);

/* batched dot products (scalarProd sample): d_C[v] = sum_i A[v][i]*B[v][i]; each product rounded to fp32, the sum
 * kept in fp64 and rounded once, like scalarProdCPU (scalarProd_cpu.cpp:32-46).  8 bytes of traffic per pair. */
int b200_scalar_prod_async(float *d_C, const float *d_A, const float *d_B, int vectorN, int elementN,
                           b200_stream_t stream);

/* scan: replaces scanExclusiveShort()/scanExclusiveLarge() (scan.cu:210-266): batchSize independent
 * exclusive +scans of arrayLength u32 each (wrap mod 2^32), single pass, 8 B of traffic per
 * element.  Any arrayLength >= 1 and any batchSize >= 0 (no power-of-two / size caps).
 * d_carry (nullable, device pointer to one u32) seeds element 0 of every array: the multi-GPU
 * path passes the allgathered prefix of the lower shards here.  In place (d_dst == d_src) is fine.
 * b200_scan_total_async writes sum(d_src[0..n)) mod 2^32 to d_total[0] (the shard total). */
int b200_scan_exclusive_async(uint32_t *d_dst, const uint32_t *d_src, size_t batchSize, size_t arrayLength,
                              const uint32_t *d_carry, b200_stream_t stream);
int b200_scan_total_async(uint32_t *d_total, const uint32_t *d_src, size_t n, b200_stream_t stream);
/* scan siblings.  Inclusive scan: dst[j] = sum src[0..j] -- what the shfl_scan sample computes
 * with shfl_scan_test x2 + uniform_add (shfl_scan.cu:55-147, host sequence :275-277, gold :157-161); same
 * single-pass kernels, same arguments as b200_scan_exclusive_async. */
int b200_scan_inclusive_async(uint32_t *d_dst, const uint32_t *d_src, size_t batchSize, size_t arrayLength,
                              const uint32_t *d_carry, b200_stream_t stream);
/* stream compaction: d_dst[0 .. *d_count) = the elements of d_src[0 .. n) that are > 0, IN INPUT ORDER (filter_arr of
 * warpAggregatedAtomicsCG.cu:59-68 keeps the same set in arbitrary order; its host gold :287-292 is ordered).
 * One single-pass chained-scan kernel; 4 B read per element + 4 B written per kept element.  n < 2^30. */
int b200_compact_positive_async(int32_t *d_dst, int32_t *d_count, const int32_t *d_src, size_t n,
                                b200_stream_t stream);
/* Integral image (summed-area table) of an 8-bit image, the shfl_scan sample's second test (shfl_intimage_rows +
 * shfl_vertical_shfl, shfl_integral_image.cuh:168-320, host sequence shfl_scan.cu:308-381):
 * d_out[r * width + c] = sum of d_img[r' * width + c'], r' <= r, c' <= c (u32, wraps).  Rows are `width` bytes
 * apart.  Any width, height >= 0 with width * height < 2^32; every output pixel is written once (5 B per pixel). */
int b200_integral_image_u8_async(uint32_t *d_out, const uint8_t *d_img, size_t width, size_t height,
                                 b200_stream_t stream);

/* sort: replaces mergeSort() (mergeSort.cu:471-537): STABLE key-value sort, sortDir != 0 ascending,
 * == 0 descending, ties keep source order.  Result in dst, buf is scratch, src is left untouched.
 * Any N >= 0 (the sample asserts N % 1024 == 0 and N <= 2^22).  src may not alias dst/buf.
 * Four 8-bit LSD passes; a pass whose digit is the same for every key is skipped (decided on the device from the
 * digit histograms, no host round trip): the sample's own keys, rand() % 65536, cost two passes.
 * d_srcVal == NULL: d_srcKey is an array of N interleaved (key, value) pairs (8 bytes each, 16-byte aligned) -- the
 * format of the sharded sort's receive buffers; buf is then unused and may be NULL. */
int b200_sort_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, uint32_t *d_bufKey, uint32_t *d_bufVal,
                       const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t N, uint32_t sortDir,
                       b200_stream_t stream);
/* stable partition of (key,val) into numBuckets <= 256 buckets by splitters: bucket(k) = number of
 * splitters s with s <= k (ascending) -- the local step of the multi-GPU sample sort.
 * d_splitters: numBuckets-1 ascending u32.  d_bucketCounts (numBuckets u32) receives the sizes. */
int b200_partition_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, const uint32_t *d_srcKey,
                            const uint32_t *d_srcVal, size_t N, const uint32_t *d_splitters, uint32_t numBuckets,
                            uint32_t sortDir, uint32_t *d_bucketCounts, b200_stream_t stream);

/* batched (segmented) stable key-value sort: batchSize independent arrays of arrayLength pairs (the sortingNetworks
 * sample's bitonicSort / oddEvenMergeSort contract, any length).  src untouched, result in dst. */
int b200_segmented_sort_kv_async(uint32_t *d_dstKey, uint32_t *d_dstVal, const uint32_t *d_srcKey,
                                 const uint32_t *d_srcVal, size_t batchSize, size_t arrayLength, uint32_t sortDir,
                                 b200_stream_t stream);

/* Pre-size the library's per-device scratch for scans of up to n elements / sorts of up to N pairs (what initScan /
 * initMergeSort do for the samples' own limits).  The *_async entry points grow their scratch on demand, and growing
 * synchronises the device and is refused inside a stream capture (cudaErrorStreamCaptureUnsupported): reserve
 * first when a call must stay asynchronous or is to be captured into a CUDA graph. */
int b200_scan_reserve(size_t n);
int b200_sort_reserve(size_t N);

/* Scan for callers that cannot allocate device memory -- NO COUNTERPART IN THE REFERENCE (the samples cudaMalloc their
 * scratch in initScan, scan.cu:182-186); added for users of this library, not to mirror a sample.
 * Who needs it: a plug-in whose host application gave the whole GPU to its own memory pool at start-up, a service
 * whose device-memory budget is fixed and audited once, a latency-critical loop that cannot afford the device-wide
 * synchronisation hidden in cudaMalloc / cudaFree.  b200_scan_exclusive_async grows its tile descriptors with
 * cudaMalloc on first use and whenever a longer array arrives; this entry point never allocates or frees anything.
 * Its descriptors live in a fixed array inside the library image: about 64 KiB of every GPU that loads
 * libb200prim.so, held until the process exits.
 * Same arguments and same result as b200_scan_exclusive_async.  Limit: the call may span at most
 * B200_SCAN_NOALLOC_MAX_TILES tiles of 8192 elements -- batchSize * ceil(arrayLength / 8192) tiles (power-of-two
 * lengths below 8192 are packed: ceil(batchSize * arrayLength / 8192)) -- so one array can be up to 2^26 elements;
 * larger requests return cudaErrorInvalidValue (1), print the limit on stderr and launch nothing.
 * The first call on each device also looks up the array's device address; make it before any stream capture.
 * Every call on a device shares the one array and inherits what the previous call left in it; the kernel itself
 * resets the only word that has to start at zero, so callers clear nothing -- but two calls that can run on the GPU
 * at the same time (different streams, no event between them) corrupt each other.  Keep them stream-ordered. */
#define B200_SCAN_NOALLOC_MAX_TILES 8192
int b200_scan_exclusive_noalloc_async(uint32_t *d_dst, const uint32_t *d_src, size_t batchSize, size_t arrayLength,
                                      const uint32_t *d_carry, b200_stream_t stream);

/* counter-based synthetic data, bit-identical to oracle_generate() (oracle/oracle.c): element i of
 * the stream depends only on (seed, first + i, kind, param), so shards are filled in place. */
int b200_generate_async(void *d_out, size_t first, size_t n, uint64_t seed, int kind, uint32_t param,
                        b200_stream_t stream);

/* ---- fused compute + collective over NVLink peer windows (one process per GPU, one NVSwitch node) -------------
 * The reference has no multi-GPU path for these primitives; the sharded versions end in a tiny
 * exchange (320 bins / a few shard sums / G shard totals).  Instead of a separate NCCL collective the last CTA of
 * the producing kernel stores its payload into every peer's window (P2P stores), flags it, waits for the peers'
 * flags and finishes the combine itself.  Setup: every rank calls b200_p2p_window_create (64-byte CUDA IPC
 * handle out), the handles are exchanged by the caller's control plane (multi_gpu.py: torch.distributed), then
 * b200_p2p_connect(rank, world, handles[world][64]).  Every rank must issue the same sequence of exchange calls,
 * stream-ordered.  A peer that never arrives is reported by b200_p2p_errors (bounded wait), not by a hang. */
int b200_p2p_window_create(unsigned char *handle_out /* 64 bytes */);
int b200_p2p_connect(int rank, int world, const unsigned char *handles);
int b200_p2p_disconnect(void);
int b200_p2p_world(void); /* 0 until connected */
int b200_p2p_errors(unsigned long long *count);
/* d_gathered[r * nbytes ..) = rank r's payload on every rank; nbytes % 4 == 0, <= 2048 */
int b200_p2p_allgather_async(const void *d_payload, size_t nbytes, void *d_gathered, b200_stream_t stream);
/* node-wide barrier on the stream: returns (stream order) once every rank's earlier work, peer stores included,
 * has completed */
int b200_p2p_barrier_async(b200_stream_t stream);
/* reduce -> exchange: d_slots[0..count) are this rank's shard partials (b200_reduce_partial_async fills the
 * others, this launch fills [index]); on every rank d_gathered[r*count + k] = rank r's slot k and
 * d_sums[k] = sum over ranks in rank order (fp64, deterministic).  d_carry (nullable) = wrapping u32 sum of slot
 * [index] over the LOWER ranks (the carry-in of this shard in a sharded scan). */
int b200_reduce_partial_exchange_async(int kind, const void *d_in, size_t n, double *d_slots, unsigned index,
                                       unsigned count, double *d_gathered, double *d_sums, uint32_t *d_carry,
                                       b200_stream_t stream);
/* sharded long scan, step 1: shard total -> peers -> *d_carry; d_work: >= 2 + world doubles of scratch */
int b200_scan_total_exchange_async(uint32_t *d_carry, const uint32_t *d_src, size_t n, double *d_work,
                                   b200_stream_t stream);
/* one shard of a sharded exclusive scan, both steps on the stream: d_dst = exclusive scan of d_src seeded with the
 * wrapping sum of the lower ranks' shards (also left in *d_carry).  16-byte aligned shards of n % 4 == 0 elements run
 * as reduce-then-scan over 8192-element tiles: pass 1 leaves one sum per tile, exchanges the shard total over the peer
 * windows and turns the sums into tile prefixes, pass 2 scans independent tiles (no look-back); other shapes take
 * b200_scan_total_exchange_async + the chained scan.  d_work: >= 2 + world doubles. */
int b200_scan_exclusive_shard_async(uint32_t *d_dst, const uint32_t *d_src, size_t n, double *d_work,
                                    uint32_t *d_carry, b200_stream_t stream);
/* large peer-mapped buffers (cudaMalloc + CUDA IPC): create locally, open the peers' handles */
int b200_p2p_buffer_create(size_t bytes, void **d_ptr, unsigned char *handle_out /* 64 bytes */);
int b200_p2p_buffer_open(const unsigned char *handle, void **d_peer);
int b200_p2p_buffer_close(void *d_peer);
int b200_p2p_buffer_free(void *d_ptr);
/* b200_sort_kv_async with the pair count read on the device: sorts the first min(*d_n, capacity) pairs; geometry
 * and scratch are sized for `capacity`.  (The multi-GPU sort runs it on a receive buffer whose fill level only the
 * GPU knows yet.) */
int b200_sort_kv_dyn_async(uint32_t *d_dstKey, uint32_t *d_dstVal, uint32_t *d_bufKey, uint32_t *d_bufVal,
                           const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t capacity, const uint32_t *d_n,
                           uint32_t sortDir, b200_stream_t stream);
/* partition + all-to-all of the multi-GPU sample sort without a collective call: bucket b == rank b, the stable
 * partition pass stores every bucket's run straight into rank b's receive buffer over NVLink.  recvKeys/recvVals:
 * HOST arrays of `world` device pointers (rank r's receive buffer as mapped here), `capacity` pairs each.
 * recvVals == NULL: recvKeys[r] are PAIR buffers (8 * capacity bytes): one 8-byte peer store per pair; sort them with
 * b200_sort_kv_dyn_async(..., d_srcKey = the own pair buffer, d_srcVal = NULL, ...).
 * d_matrix [world][world]
 * (row = source rank), d_info[0] = pairs received, d_info[1] != 0: a buffer would overflow, nothing was sent and
 * d_info[0] = 0 (a b200_sort_kv_dyn_async behind this call then sorts nothing); the sizes needed are the column sums
 * of d_matrix. */
int b200_partition_exchange_kv_async(const uint32_t *d_srcKey, const uint32_t *d_srcVal, size_t N,
                                     const uint32_t *d_splitters, uint32_t sortDir, void *const *recvKeys,
                                     void *const *recvVals, size_t capacity, uint32_t *d_matrix, uint32_t *d_info,
                                     b200_stream_t stream);
/* histogram64 + histogram256 of the concatenation of all ranks' shards, on every rank, one kernel per rank */
int b200_histogram_fused_allreduce_async(uint32_t *d_hist256, uint32_t *d_hist64, const void *d_data,
                                         size_t byteCount, b200_stream_t stream);

/* measured NVLink roofline of the exchange phases: GB/s at which the current device writes `bytes` from d_src into
 * d_dst (a peer-mapped buffer).  mode 0: 4-byte lane stores, 1: 16-byte lane stores, 2: cp.async.bulk from shared
 * memory, 3: cudaMemcpyAsync (copy engine; p2pBandwidthLatencyTest.cu:113-146).  Best of reps. */
int b200_p2p_probe_store(void *d_dst, const void *d_src, size_t bytes, int mode, int reps, double *best_gbs,
                         b200_stream_t stream);
/* how long a fused kernel waits for a peer's word before it gives up and counts an error (default 30 s).  With the
 * current device connected it applies to that device's whole connected set (every device of a b200_ctx); otherwise it
 * sets the value later connections start with. */
int b200_p2p_set_timeout_ms(unsigned ms);
/* splitters of the sharded sort chosen on the device (one exchange, no collective call, no host round trip):
 * d_splitters receives world-1 ascending values in ordered key space, identical on every rank */
int b200_sort_splitters_async(const uint32_t *d_srcKey, size_t N, uint32_t sortDir, uint32_t *d_splitters,
                              b200_stream_t stream);

/* ---- sharded primitives for ONE host process that drives all GPUs of the node (csrc/sharded.cu) ----------------
 * The reference's multi-GPU idiom (Samples/0_Introduction/simpleMultiGPU/simpleMultiGPU.cu:132-133,157-200): a
 * loop of cudaSetDevice + asynchronous launches, then a loop of synchronisations.  b200_ctx_create enables peer
 * access between the listed devices and gives each one an exchange window and a stream; the entry points take HOST
 * arrays with one device pointer / one length per device (index i = devices[i]) and return a status.  The exchange
 * each primitive needs is done by the kernels themselves over NVLink; there is no NCCL and no IPC in this mode.
 * A peer that never arrives is reported as cudaErrorLaunchTimeout by the blocking calls and by b200_ctx_synchronize.
 * Not thread-safe: one host thread per context; the devices of a context must not run other exchange calls. */
typedef struct b200_ctx b200_ctx;
int b200_ctx_create(int ndev, const int *devices, b200_ctx **ctx);
int b200_ctx_destroy(b200_ctx *ctx);
int b200_ctx_num_devices(const b200_ctx *ctx);
int b200_ctx_stream(b200_ctx *ctx, int i, b200_stream_t *stream); /* the stream device i's work is enqueued on */
int b200_ctx_synchronize(b200_ctx *ctx);
/* histogram64 + histogram256 of the concatenation of all shards; every device's d_hist256[i] (256 u32) and
 * d_hist64[i] (64 u32) receive the GLOBAL bins.  Asynchronous: b200_ctx_synchronize before reading. */
int b200_histogram_sharded(b200_ctx *ctx, uint32_t *const *d_hist256, uint32_t *const *d_hist64,
                           const void *const *d_data, const size_t *byteCount);
/* sum over all shards, blocking.  int wraps like reduceCPU<int>; float/double: every shard's sum is rounded to the
 * element type (it is that device's reduce<T> result), the shard sums are added in device order in fp64
 * (deterministic) and the total is rounded to the element type. */
int b200_reduce_sharded_i32(b200_ctx *ctx, const int32_t *const *d_in, const size_t *n, int32_t *h_result);
int b200_reduce_sharded_f32(b200_ctx *ctx, const float *const *d_in, const size_t *n, float *h_result);
int b200_reduce_sharded_f64(b200_ctx *ctx, const double *const *d_in, const size_t *n, double *h_result);
/* exclusive scan of ONE array split into contiguous shards (device i holds slice i, n[i] elements).  Asynchronous. */
int b200_scan_exclusive_sharded(b200_ctx *ctx, uint32_t *const *d_dst, const uint32_t *const *d_src, const size_t *n);
/* stable key-value sort of the concatenation of all shards; device i ends up with the i-th key range (h_nOut[i]
 * pairs) in d_dstKey[i] / d_dstVal[i] (dstCapacity pairs each; 1.25 x the largest shard is ample for random keys).
 * Blocking.  cudaErrorMemoryAllocation (2) when a device would receive more than dstCapacity pairs: nothing was
 * moved, the dst arrays are untouched and h_nOut holds the sizes that are needed.  The splitters are key values and
 * do not break ties: all copies of one key go to one device, so a key that occurs more often than dstCapacity can
 * only be sorted with a dstCapacity of at least that count. */
int b200_sort_kv_sharded(b200_ctx *ctx, uint32_t *const *d_dstKey, uint32_t *const *d_dstVal, size_t dstCapacity,
                         const uint32_t *const *d_srcKey, const uint32_t *const *d_srcVal, const size_t *n,
                         uint32_t sortDir, size_t *h_nOut);

/* host-buffer variants: H2D copies, kernels and D2H of the result are all inside the call
 * (chunked and double-buffered so the copy engine and the SMs overlap); they synchronise before
 * returning.  These are what bench.py's `e2e` number times. */
int b200_histogram256_host(uint32_t *h_hist256, const void *h_data, size_t byteCount);
int b200_histogram64_host(uint32_t *h_hist64, const void *h_data, size_t byteCount);
int b200_reduce_i32_host(const int32_t *h_in, size_t n, int32_t *h_result);
int b200_reduce_f32_host(const float *h_in, size_t n, float *h_result);
int b200_reduce_f64_host(const double *h_in, size_t n, double *h_result);
int b200_scan_exclusive_host(uint32_t *h_dst, const uint32_t *h_src, size_t batchSize, size_t arrayLength);
int b200_sort_kv_host(uint32_t *h_dstKey, uint32_t *h_dstVal, const uint32_t *h_srcKey, const uint32_t *h_srcVal,
                      size_t N, uint32_t sortDir);
/* Pinned (registered, device-mapped) host memory for the variants above, placed on the NUMA node of the CURRENT
 * device when the platform allows it (mbind, else first touch from that node's CPUs): a pinned buffer on the wrong
 * socket halves the PCIe rate.  (If sysfs does not name the GPU's node the allocator measures it.) */
int b200_host_alloc(void **h_ptr, size_t bytes);
int b200_host_free(void *h_ptr);
/* where a b200_host_alloc block lives: NUMA node of its device, node of its first / last page (-1 unknown),
 * placement method (0 none possible, 1 mbind, 2 first touch on the node's CPUs) */
int b200_host_alloc_info(const void *h_ptr, int *gpu_node, int *mem_node_first, int *mem_node_last, int *method);
/* the PCIe roofline of the *_host entry points, measured in place: one plain cudaMemcpyAsync of `bytes` from pinned
 * h_src to device memory between events, best of `reps`, in GB/s (method of the reference's
 * 5_Domain_Specific/p2pBandwidthLatencyTest/p2pBandwidthLatencyTest.cu:113-146) */
int b200_host_probe_h2d(const void *h_src, size_t bytes, int reps, double *best_gbs);
/* reps back-to-back copies of the pinned buffer into d_dst; *seconds = first copy issued -> last copy complete.  Called
 * by every rank between the same two barriers it measures what the host can feed all GPUs at once. */
int b200_host_probe_h2d_sustained(void *d_dst, const void *h_src, size_t bytes, int reps, double *seconds);

#ifdef __cplusplus
}
#endif
#endif /* B200PRIM_H */
