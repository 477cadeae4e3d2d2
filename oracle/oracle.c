/*
 * oracle.c -- CPU restatement of the cuda-samples gold paths for the data-parallel
 * primitive hot path (histogram64/256, reduce<T>, scanExclusive*, mergeSort key-value).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (cuda_samples_b200/, include/,
 * the C-ABI library) may include, link or call this file.  Only tests/, the smoke check in
 * __graft_entry__.py and the cpu_baseline / --impl reference legs of bench.py use it, and
 * there only as the checker / the reported CPU baseline.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle.py) against
 *   (1) the reference's own gold code compiled unmodified from /root/reference into
 *       oracle/_ref/libref_gold.so (oracle/build.py), on the samples' own seeded inputs, and
 *   (2) the known-answer hashes of SURVEY.md section 8c committed in tests/golden/kat.json.
 *
 * Citations are relative to /root/reference/Samples.
 * Build: gcc -O2 -fwrapv -fno-fast-math (never -ffast-math: Kahan summation must survive).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef uint32_t uint;

/* ------------------------------------------------------------------------------------------
 * histogram256CPU -- 2_Concepts_and_Techniques/histogram/histogram_gold.cpp:48-62
 * Bins are zeroed, then every byte b of the input increments bin b (u32 counters, wrap mod
 * 2^32).  The reference walks u32 words and peels 4 bytes; byte order is irrelevant for the
 * counts, so a byte loop is the same function.  64-bit length (the reference takes uint).
 * ---------------------------------------------------------------------------------------- */
void oracle_histogram256(uint32_t *hist, const void *data, uint64_t byteCount)
{
    const uint8_t *p = (const uint8_t *)data;
    for (int i = 0; i < 256; i++)
        hist[i] = 0;
    for (uint64_t i = 0; i < byteCount; i++)
        hist[p[i]]++;
}

/* histogram64CPU -- histogram_gold.cpp:32-46: bin = (byte >> 2) & 0x3F (the TOP six bits). */
void oracle_histogram64(uint32_t *hist, const void *data, uint64_t byteCount)
{
    const uint8_t *p = (const uint8_t *)data;
    for (int i = 0; i < 64; i++)
        hist[i] = 0;
    for (uint64_t i = 0; i < byteCount; i++)
        hist[(p[i] >> 2) & 0x3F]++;
}

/* ------------------------------------------------------------------------------------------
 * scanExclusiveHost -- 2_Concepts_and_Techniques/scan/scan_gold.cpp:30-38
 * batchSize independent arrays of arrayLength u32; dst[0]=0, dst[j]=src[j-1]+dst[j-1] (wraps).
 * `carry` seeds dst[0] of every array (0 reproduces the reference; the multi-GPU tests use it
 * to check a shard scanned with the allgathered carry).
 * ---------------------------------------------------------------------------------------- */
void oracle_scan_exclusive(uint32_t *dst, const uint32_t *src, uint64_t batchSize, uint64_t arrayLength,
                           uint32_t carry)
{
    for (uint64_t i = 0; i < batchSize; i++, src += arrayLength, dst += arrayLength) {
        if (arrayLength == 0)
            continue;
        dst[0] = carry;
        for (uint64_t j = 1; j < arrayLength; j++)
            dst[j] = src[j - 1] + dst[j - 1];
    }
}

/* ------------------------------------------------------------------------------------------
 * reduceCPU<T> -- 2_Concepts_and_Techniques/reduction/reduction.cpp:167-180 (Kahan summation).
 * T=int: the compensation term is identically 0 and the result is the wrapping 32-bit sum
 * (SURVEY.md 8c KAT-RED: wraps negative from 2^26 elements upward).  Done in unsigned so the
 * wrap is defined behaviour.
 * ---------------------------------------------------------------------------------------- */
int32_t oracle_reduce_i32(const int32_t *data, uint64_t size)
{
    uint32_t sum = (uint32_t)data[0];
    uint32_t c   = 0;
    for (uint64_t i = 1; i < size; i++) {
        uint32_t y = (uint32_t)data[i] - c;
        uint32_t t = sum + y;
        c          = (t - sum) - y;
        sum        = t;
    }
    return (int32_t)sum;
}

float oracle_reduce_f32(const float *data, uint64_t size)
{
    volatile float sum = data[0]; /* volatile: forbid x87/FMA style re-association of the Kahan terms */
    volatile float c   = 0.0f;
    for (uint64_t i = 1; i < size; i++) {
        float y = data[i] - c;
        float t = sum + y;
        c       = (t - sum) - y;
        sum     = t;
    }
    return sum;
}

double oracle_reduce_f64(const double *data, uint64_t size)
{
    volatile double sum = data[0];
    volatile double c   = 0.0;
    for (uint64_t i = 1; i < size; i++) {
        double y = data[i] - c;
        double t = sum + y;
        c        = (t - sum) - y;
        sum      = t;
    }
    return sum;
}

/* ------------------------------------------------------------------------------------------
 * scalarProdCPU -- 2_Concepts_and_Techniques/scalarProd/scalarProd_cpu.cpp:32-46
 * vectorN dot products of elementN floats: the product is a float (float * float), the running sum a
 * double, the result rounded to float once.
 * ---------------------------------------------------------------------------------------- */
void oracle_scalar_prod(float *C, const float *A, const float *B, int vectorN, int elementN)
{
    for (int vec = 0; vec < vectorN; vec++) {
        const float *a = A + (size_t)elementN * vec, *b = B + (size_t)elementN * vec;
        double       sum = 0;
        for (int pos = 0; pos < elementN; pos++) {
            volatile float prod = a[pos] * b[pos]; /* volatile: keep the product a rounded float (no fused multiply-add) */
            sum += prod;
        }
        C[vec] = (float)sum;
    }
}

/* ------------------------------------------------------------------------------------------
 * Stable key-value sort: the SPECIFICATION of mergeSort (0_Introduction/mergeSort/mergeSort.cu:
 * 471-537; stable because the tile sort ranks A exclusive / B inclusive, :135-136, and the merges
 * do the same, :304,310).  sortDir != 0 ascending, == 0 descending; ties keep source order in
 * both directions.  Bottom-up merge sort, left run wins ties.  Deliberately NOT a radix sort so
 * that the checker and the checked do not share an algorithm.
 * ---------------------------------------------------------------------------------------- */
static inline int before(uint32_t a, uint32_t b, uint32_t dir) { return dir ? (a < b) : (a > b); }

void oracle_stable_sort_kv(uint32_t *dstKey, uint32_t *dstVal, const uint32_t *srcKey, const uint32_t *srcVal,
                           uint64_t N, uint32_t sortDir)
{
    if (N == 0)
        return;
    uint32_t *tk = (uint32_t *)malloc(N * sizeof(uint32_t));
    uint32_t *tv = (uint32_t *)malloc(N * sizeof(uint32_t));
    uint32_t *ik = dstKey, *iv = dstVal, *ok = tk, *ov = tv;
    memcpy(ik, srcKey, N * sizeof(uint32_t));
    memcpy(iv, srcVal, N * sizeof(uint32_t));
    for (uint64_t w = 1; w < N; w <<= 1) {
        for (uint64_t lo = 0; lo < N; lo += 2 * w) {
            uint64_t mid = lo + w < N ? lo + w : N;
            uint64_t hi  = lo + 2 * w < N ? lo + 2 * w : N;
            uint64_t a = lo, b = mid, o = lo;
            while (a < mid && b < hi) {
                if (before(ik[b], ik[a], sortDir)) { /* strictly before: ties take the left run */
                    ok[o] = ik[b];
                    ov[o++] = iv[b++];
                }
                else {
                    ok[o] = ik[a];
                    ov[o++] = iv[a++];
                }
            }
            while (a < mid) {
                ok[o] = ik[a];
                ov[o++] = iv[a++];
            }
            while (b < hi) {
                ok[o] = ik[b];
                ov[o++] = iv[b++];
            }
        }
        uint32_t *t;
        t = ik, ik = ok, ok = t;
        t = iv, iv = ov, ov = t;
    }
    if (ik != dstKey) {
        memcpy(dstKey, ik, N * sizeof(uint32_t));
        memcpy(dstVal, iv, N * sizeof(uint32_t));
    }
    free(tk);
    free(tv);
}

/* ------------------------------------------------------------------------------------------
 * mergeSortHost -- 0_Introduction/mergeSort/mergeSort_host.cpp:288-363, restated by what it
 * computes rather than how:
 *   (1) :305-315  copy src, then sort every 1024-aligned tile with "bubbleSort" (:257-283), which
 *       is a SELECTION sort that swaps the minimum into place -- it reorders equal keys, so the
 *       whole function is NOT stable when a tile holds duplicates (SURVEY.md 8c caveat);
 *   (2) :327-358  log2(N/1024) rounds of pairwise merges of adjacent runs of length `stride`.
 *       Sample ranks / limits / elementary intervals (:113-252) only cut each A|B pair into
 *       independent pieces; merge() (:179-205) places A[i] after the B elements strictly before
 *       it and B[i] after the A elements before-or-equal, i.e. an ordinary stable two-way merge
 *       in which A wins ties.  A trailing run without a partner is passed through (:342-350).
 * Result lands in dst (the reference ping-pongs so that the last round writes dst, :292-303);
 * buf is scratch.
 * ---------------------------------------------------------------------------------------- */
static void selection_sort_tile(uint32_t *key, uint32_t *val, uint64_t n, uint32_t dir)
{
    if (n <= 1)
        return;
    for (uint64_t bottom = 0; bottom < n - 1; bottom++) {
        uint64_t savePos = bottom;
        uint32_t saveKey = key[bottom];
        for (uint64_t i = bottom + 1; i < n; i++)
            if (before(key[i], saveKey, dir)) {
                savePos = i;
                saveKey = key[i];
            }
        if (savePos != bottom) {
            uint32_t t;
            t = key[savePos], key[savePos] = key[bottom], key[bottom] = t;
            t = val[savePos], val[savePos] = val[bottom], val[bottom] = t;
        }
    }
}

void oracle_merge_sort_host(uint32_t *dstKey, uint32_t *dstVal, uint32_t *bufKey, uint32_t *bufVal,
                            const uint32_t *srcKey, const uint32_t *srcVal, uint64_t N, uint32_t sortDir)
{
    const uint64_t TILE = 1024; /* SHARED_SIZE_LIMIT, mergeSort_common.h:33 */
    uint32_t *ik, *iv, *ok, *ov;
    uint32_t stageCount = 0;
    for (uint64_t stride = TILE; stride < N; stride <<= 1)
        stageCount++;
    if (stageCount & 1) {
        ik = bufKey, iv = bufVal, ok = dstKey, ov = dstVal;
    }
    else {
        ik = dstKey, iv = dstVal, ok = bufKey, ov = bufVal;
    }
    memcpy(ik, srcKey, N * sizeof(uint32_t));
    memcpy(iv, srcVal, N * sizeof(uint32_t));
    for (uint64_t pos = 0; pos < N; pos += TILE)
        selection_sort_tile(ik + pos, iv + pos, (N - pos < TILE) ? N - pos : TILE, sortDir);

    for (uint64_t stride = TILE; stride < N; stride <<= 1) {
        for (uint64_t lo = 0; lo < N; lo += 2 * stride) {
            uint64_t mid = lo + stride < N ? lo + stride : N;
            uint64_t hi  = lo + 2 * stride < N ? lo + 2 * stride : N;
            uint64_t a = lo, b = mid, o = lo;
            while (a < mid && b < hi) {
                if (before(ik[b], ik[a], sortDir)) {
                    ok[o] = ik[b];
                    ov[o++] = iv[b++];
                }
                else {
                    ok[o] = ik[a];
                    ov[o++] = iv[a++];
                }
            }
            while (a < mid) {
                ok[o] = ik[a];
                ov[o++] = iv[a++];
            }
            while (b < hi) {
                ok[o] = ik[b];
                ov[o++] = iv[b++];
            }
        }
        uint32_t *t;
        t = ik, ik = ok, ok = t;
        t = iv, iv = ov, ov = t;
    }
}

/* ------------------------------------------------------------------------------------------
 * Validators -- 0_Introduction/mergeSort/mergeSort_validate.cpp
 * validateSortedKeys :38-97 (multiset equality through a numValues-bin histogram + order);
 * validateSortedValues :108-128 (resKey[j]==srcKey[resVal[j]]; stability only reported there --
 * returned here through *stable).  Silent (the reference prints).
 * ---------------------------------------------------------------------------------------- */
uint32_t oracle_validate_sorted_keys(const uint32_t *resKey, const uint32_t *srcKey, uint64_t batchSize,
                                     uint64_t arrayLength, uint32_t numValues, uint32_t sortDir)
{
    if (arrayLength < 2)
        return 1;
    uint32_t *srcHist = (uint32_t *)malloc((size_t)numValues * sizeof(uint32_t));
    uint32_t *resHist = (uint32_t *)malloc((size_t)numValues * sizeof(uint32_t));
    uint32_t  flag    = 1;
    for (uint64_t j = 0; j < batchSize && flag; j++, srcKey += arrayLength, resKey += arrayLength) {
        memset(srcHist, 0, (size_t)numValues * sizeof(uint32_t));
        memset(resHist, 0, (size_t)numValues * sizeof(uint32_t));
        for (uint64_t i = 0; i < arrayLength; i++) {
            if (srcKey[i] < numValues && resKey[i] < numValues) {
                srcHist[srcKey[i]]++;
                resHist[resKey[i]]++;
            }
            else {
                flag = 0;
                break;
            }
        }
        for (uint32_t i = 0; i < numValues && flag; i++)
            if (srcHist[i] != resHist[i])
                flag = 0;
        for (uint64_t i = 0; i + 1 < arrayLength && flag; i++)
            if ((sortDir && resKey[i] > resKey[i + 1]) || (!sortDir && resKey[i] < resKey[i + 1]))
                flag = 0;
    }
    free(resHist);
    free(srcHist);
    return flag;
}

int oracle_validate_sorted_values(const uint32_t *resKey, const uint32_t *resVal, const uint32_t *srcKey,
                                  uint64_t batchSize, uint64_t arrayLength, int *stable)
{
    int correctFlag = 1, stableFlag = 1;
    for (uint64_t i = 0; i < batchSize; i++, resKey += arrayLength, resVal += arrayLength)
        for (uint64_t j = 0; j < arrayLength; j++) {
            if (resKey[j] != srcKey[resVal[j]])
                correctFlag = 0;
            if (j + 1 < arrayLength && resKey[j] == resKey[j + 1] && resVal[j] > resVal[j + 1])
                stableFlag = 0;
        }
    if (stable)
        *stable = stableFlag;
    return correctFlag;
}

/* fillValues -- mergeSort_validate.cpp:102-106 */
void oracle_fill_values(uint32_t *val, uint64_t N)
{
    for (uint64_t i = 0; i < N; i++)
        val[i] = (uint32_t)i;
}

/* ------------------------------------------------------------------------------------------
 * The samples' own input generators (glibc rand(); same libc on the GPU box).
 *   mode 0: bytes  rand() % 256                 histogram/main.cpp:92-96  (seed 2009)
 *   mode 1: u32    rand()                       scan/main.cpp:52-56       (seed 2009)
 *   mode 2: i32    rand() & 0xFF                reduction/reduction.cpp:483-487 (never seeded = seed 1)
 *   mode 3: f32    (rand() & 0xFF)/(float)RAND_MAX   reduction.cpp:488-490
 *   mode 4: u32    rand() % 65536               mergeSort/main.cpp:65-69  (seed 2009)
 *   mode 5: f64    (rand() & 0xFF)/(double)RAND_MAX  reduction.cpp:488-490 with T=double
 * ---------------------------------------------------------------------------------------- */
void oracle_fill_sample_input(void *out, uint64_t n, uint32_t seed, int mode)
{
    srand(seed);
    switch (mode) {
    case 0:
        for (uint64_t i = 0; i < n; i++)
            ((uint8_t *)out)[i] = (uint8_t)(rand() % 256);
        break;
    case 1:
        for (uint64_t i = 0; i < n; i++)
            ((uint32_t *)out)[i] = (uint32_t)rand();
        break;
    case 2:
        for (uint64_t i = 0; i < n; i++)
            ((int32_t *)out)[i] = (int32_t)(rand() & 0xFF);
        break;
    case 3:
        for (uint64_t i = 0; i < n; i++)
            ((float *)out)[i] = (rand() & 0xFF) / (float)RAND_MAX;
        break;
    case 4:
        for (uint64_t i = 0; i < n; i++)
            ((uint32_t *)out)[i] = (uint32_t)(rand() % 65536);
        break;
    case 5:
        for (uint64_t i = 0; i < n; i++)
            ((double *)out)[i] = (rand() & 0xFF) / (double)RAND_MAX;
        break;
    }
}

/* ------------------------------------------------------------------------------------------
 * Counter-based generator: splitmix64(seed + global index of the 64-bit word).
 * The CUDA twin lives in cuda_samples_b200/csrc/generate.cu; the two must stay bit-identical
 * (tests/test_generate.py).  Lets every GPU fill its shard in place and the CPU regenerate any
 * chunk for gold without a 32 GiB host buffer.
 *   kind 0: raw bytes (uniform)                      -- histogram inputs
 *   kind 1: u32 = low/high half >> 1 (31-bit)        -- scan inputs (like rand())
 *   kind 2: i32 = half & 0xFF                        -- reduce<int> inputs
 *   kind 3: f32 = (half & 0xFF) / 2147483647.0f      -- reduce<float> inputs
 *   kind 4: u32 = bijection P(index ^ seed)          -- sort dataset D1 (all keys distinct)
 *   kind 5: u32 = P(index mod M), M = `param`        -- sort dataset D2 (tile-unique, duplicated)
 *   kind 6: u32 = half % param                       -- sort dataset D3-like (keys < param)
 *   kind 7: f64 = (half & 0xFF) / 2147483647.0       -- reduce<double> inputs (one per u32 half)
 * `first` is the global element index of out[0] (elements of the kind's own width).
 * ---------------------------------------------------------------------------------------- */
static inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

static inline uint32_t bijection32(uint32_t x)
{
    x *= 2654435761u;
    x ^= x >> 15;
    x *= 2246822519u;
    x ^= x >> 13;
    x *= 3266489917u;
    x ^= x >> 16;
    return x;
}

static inline uint32_t gen_half(uint64_t seed, uint64_t idx32)
{
    uint64_t w = splitmix64(seed + (idx32 >> 1));
    return (uint32_t)((idx32 & 1) ? (w >> 32) : w);
}

void oracle_generate(void *out, uint64_t first, uint64_t n, uint64_t seed, int kind, uint32_t param)
{
    switch (kind) {
    case 0:
        for (uint64_t i = 0; i < n; i++) {
            uint64_t g = first + i;
            uint64_t w = splitmix64(seed + (g >> 3));
            ((uint8_t *)out)[i] = (uint8_t)(w >> (8 * (g & 7)));
        }
        break;
    case 1:
        for (uint64_t i = 0; i < n; i++)
            ((uint32_t *)out)[i] = gen_half(seed, first + i) >> 1;
        break;
    case 2:
        for (uint64_t i = 0; i < n; i++)
            ((int32_t *)out)[i] = (int32_t)(gen_half(seed, first + i) & 0xFF);
        break;
    case 3:
        for (uint64_t i = 0; i < n; i++)
            ((float *)out)[i] = (float)(gen_half(seed, first + i) & 0xFF) / 2147483647.0f;
        break;
    case 4:
        for (uint64_t i = 0; i < n; i++)
            ((uint32_t *)out)[i] = bijection32((uint32_t)(first + i) ^ (uint32_t)seed);
        break;
    case 5:
        for (uint64_t i = 0; i < n; i++)
            ((uint32_t *)out)[i] = bijection32((uint32_t)((first + i) % param));
        break;
    case 6:
        for (uint64_t i = 0; i < n; i++)
            ((uint32_t *)out)[i] = gen_half(seed, first + i) % param;
        break;
    case 7:
        for (uint64_t i = 0; i < n; i++)
            ((double *)out)[i] = (double)(gen_half(seed, first + i) & 0xFF) / 2147483647.0;
        break;
    }
}

/* FNV-1a 64 over bytes (the hash the SURVEY.md 8c known answers are quoted in). */
/* shfl_scan sample, CPUverify (Samples/2_Concepts_and_Techniques/shfl_scan/shfl_scan.cu:157-161): the gold is the
 * running sum h_data[i + 1] = h_data[i] + h_data[i + 1], i.e. an INCLUSIVE scan (int arithmetic, wraps). */
void oracle_scan_inclusive(uint32_t *dst, const uint32_t *src, uint64_t n)
{
    uint32_t run = 0;
    for (uint64_t i = 0; i < n; i++) {
        run += src[i];
        dst[i] = run;
    }
}

/* warpAggregatedAtomicsCG sample, host loop of main() (Samples/3_CUDA_Features/warpAggregatedAtomicsCG/
 * warpAggregatedAtomicsCG.cu:287-292): if (data_to_filter[i] > 0) host_filtered_data[host_flt_count++] = ... */
uint64_t oracle_filter_positive(int32_t *dst, const int32_t *src, uint64_t n)
{
    uint64_t count = 0;
    for (uint64_t i = 0; i < n; i++)
        if (src[i] > 0)
            dst[count++] = src[i];
    return count;
}

/* shfl_scan sample, integral image test (Samples/2_Concepts_and_Techniques/shfl_scan/shfl_scan.cu:308-381): the
 * first kernel leaves the prefix sums of every scanline (what verifyDataRowSums checks, :193-205: gold = i + 1 on
 * the all-ones image), the second turns every column into its running sum in place (shfl_integral_image.cuh:268-320),
 * and the corner must equal w * h (:372-381).  Restated as those two steps, u32 arithmetic. */
void oracle_integral_image_u8(uint32_t *dst, const uint8_t *img, uint32_t w, uint32_t h)
{
    for (uint64_t r = 0; r < h; r++) {
        uint32_t run = 0;
        for (uint64_t c = 0; c < w; c++) {
            run += img[r * w + c];
            dst[r * w + c] = run;
        }
    }
    for (uint64_t r = 1; r < h; r++)
        for (uint64_t c = 0; c < w; c++)
            dst[r * w + c] += dst[(r - 1) * w + c];
}

uint64_t oracle_fnv1a64(const void *data, uint64_t nbytes)
{
    const uint8_t *p = (const uint8_t *)data;
    uint64_t       h = 0xcbf29ce484222325ull;
    for (uint64_t i = 0; i < nbytes; i++) {
        h ^= p[i];
        h *= 0x100000001b3ull;
    }
    return h;
}
