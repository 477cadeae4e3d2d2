// integral_image.cu -- summed-area table of an 8-bit image: out[r][c] = sum of img[0..r][0..c] (u32, wraps).
//
// What the shfl_scan sample builds with shfl_intimage_rows (row prefix sums, one CTA per 1920-pixel row:
// Samples/2_Concepts_and_Techniques/shfl_scan/shfl_integral_image.cuh:168-257) followed by shfl_vertical_shfl
// (column running sums IN PLACE over the u32 image, 60 CTAs walking 1080 rows in 135 dependent steps, :268-320;
// host sequence shfl_scan.cu:345-362).  That is 1 + 4 bytes per pixel for the first kernel and 4 + 4 for the
// second: 13 B/pixel, and the second kernel is a serial chain.
//
// Here: reduce-then-scan over warp tiles, every output pixel written exactly once (5 B/pixel + the carries).
//   tile          R rows (a "band", R = 8/16/32) x 128 columns (a "block"): one warp, lane l owns columns 4l..4l+3
//                 of the block for all R rows, so the vertical running sum lives in four registers per lane and
//                 the horizontal prefix of a row is one warp scan -- no shared memory, no CTA barrier.
//   ii_sums       reads the image once: per tile the column sums over the band (packed 16-bit SIMD adds) and the
//                 row sums over the block (dp4a + redux.sync).
//   ii_prefix     one launch, two roles: exclusive prefix of the column sums down the bands (matrix = 1/R of the
//                 output) and exclusive prefix of the row sums along the blocks (1/128 of the output).
//   scan          the library's own batched inclusive scan along each band row of the column carries: after it
//                 carry[b][c] = out[first row of band b - 1][c], the full integral above the band.
//   ii_write      per tile: acc(c) = carry[b][c]; for each row: 4-pixel prefix, warp scan (two rows per scan, 16
//                 bits each), + the row's carry from the blocks to its left, acc += it, one 16-byte store per lane.
// Rows whose width or base address is not a multiple of 4 bytes / 16 bytes take the guarded scalar variant.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr unsigned kIICols  = 128; // columns per tile
#ifndef B200_II_BATCH
#define B200_II_BATCH 4
#endif
constexpr unsigned kIIBatch = B200_II_BATCH; // rows loaded ahead per iteration of ii_write (even)
constexpr unsigned kIIWarps = 8;   // warps per CTA (tiles are independent; the CTA is only a scheduling unit)

struct IIParams {
    const uint8_t *img;
    uint32_t      *out;
    uint32_t      *colcarry; // [nbands][wp]
    uint32_t      *rowcarry; // [h][nblk]
    unsigned       w, h, R, nbands, nblk, wp, ntiles;
};

// four pixels of row r starting at column c (c % 4 == 0); pixels past the row end read as 0
template <bool kAligned>
__device__ __forceinline__ unsigned load_px4(const IIParams &p, unsigned r, unsigned c)
{
    const uint8_t *row = p.img + (size_t)r * p.w;
    if (kAligned)
        return c < p.w ? __ldg(reinterpret_cast<const unsigned *>(row + c)) : 0u;
    unsigned x = 0;
#pragma unroll
    for (unsigned k = 0; k < 4; k++)
        if (c + k < p.w)
            x |= (unsigned)__ldg(row + c + k) << (8 * k);
    return x;
}

template <bool kAligned>
__global__ void __launch_bounds__(kIIWarps * 32) ii_sums_kernel(const IIParams p)
{
    const unsigned lane = lane_id();
    const unsigned tile = blockIdx.x * kIIWarps + (threadIdx.x >> 5);
    if (tile >= p.ntiles)
        return;
    const unsigned b = tile / p.nblk, j = tile - b * p.nblk;
    const unsigned r0 = b * p.R, rows = min(p.R, p.h - r0), c = j * kIICols + 4 * lane;
    unsigned       lo = 0, hi = 0, mine = 0; // lo: columns 0,2 as 16-bit fields; hi: columns 1,3 (R*255 < 2^16)
#pragma unroll 8
    for (unsigned i = 0; i < rows; i++) {
        const unsigned x = load_px4<kAligned>(p, r0 + i, c);
        lo += x & 0x00ff00ffu;
        hi += (x >> 8) & 0x00ff00ffu;
        const unsigned t = __reduce_add_sync(0xffffffffu, __dp4a(x, 0x01010101u, 0u));
        if (lane == i)
            mine = t;
    }
    if (lane < rows)
        p.rowcarry[(size_t)(r0 + lane) * p.nblk + j] = mine;
    *reinterpret_cast<uint4 *>(p.colcarry + (size_t)b * p.wp + c) =
        make_uint4(lo & 0xffffu, hi & 0xffffu, lo >> 16, hi >> 16);
}

// The same sums from 16-byte loads (rows 16-byte aligned): a warp covers R rows x 512 columns = four blocks, lane l
// owns columns 16l..16l+15; the row sums of a block are a reduction over its 8 lanes.
__global__ void __launch_bounds__(kIIWarps * 32) ii_sums16_kernel(const IIParams p)
{
    const unsigned lane = lane_id();
    const unsigned nsb  = (p.nblk + 3) / 4; // 512-column super blocks per band
    const unsigned tile = blockIdx.x * kIIWarps + (threadIdx.x >> 5);
    if (tile >= p.nbands * nsb)
        return;
    const unsigned b = tile / nsb, sb = tile - b * nsb;
    const unsigned r0 = b * p.R, rows = min(p.R, p.h - r0), c = sb * 512 + 16 * lane;
    const bool     in = c < p.w;
    const uint8_t *src = p.img + (size_t)r0 * p.w + c;
    unsigned       lo[4] = {0, 0, 0, 0}, hi[4] = {0, 0, 0, 0};
    for (unsigned i0 = 0; i0 < rows; i0 += 8) {
        uint4 x[8];
#pragma unroll
        for (unsigned k = 0; k < 8; k++)
            x[k] = (in && i0 + k < rows) ? ld_stream_u4(src + (size_t)(i0 + k) * p.w) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (unsigned k = 0; k < 8; k++) {
            const unsigned v[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
            unsigned       t = 0;
#pragma unroll
            for (unsigned q = 0; q < 4; q++) {
                lo[q] += v[q] & 0x00ff00ffu;
                hi[q] += (v[q] >> 8) & 0x00ff00ffu;
                t = __dp4a(v[q], 0x01010101u, t);
            }
            t += __shfl_xor_sync(0xffffffffu, t, 1);
            t += __shfl_xor_sync(0xffffffffu, t, 2);
            t += __shfl_xor_sync(0xffffffffu, t, 4);
            // every lane of an 8-lane group now holds its block's sum of this row: lane k of the group stores it
            // (one 4-byte store per 128 pixels)
            if ((lane & 7) == k && i0 + k < rows && sb * 4 + (lane >> 3) < p.nblk)
                p.rowcarry[(size_t)(r0 + i0 + k) * p.nblk + sb * 4 + (lane >> 3)] = t;
        }
    }
    if (c < p.wp) {
        uint4 *dst = reinterpret_cast<uint4 *>(p.colcarry + (size_t)b * p.wp + c);
#pragma unroll
        for (unsigned q = 0; q < 4; q++)
            dst[q] = make_uint4(lo[q] & 0xffffu, hi[q] & 0xffffu, lo[q] >> 16, hi[q] >> 16);
    }
}

// Column carries: exclusive prefix down the bands.  One CTA per 32 columns, warp k takes a contiguous run of bands
// (128-byte row segments), the runs are chained through shared memory; the second read of a run hits L1/L2.
__device__ __forceinline__ void ii_rowprefix(const IIParams &p, unsigned r, unsigned lane);

__global__ void __launch_bounds__(1024) ii_prefix_kernel(const IIParams p, const unsigned col_ctas)
{
    __shared__ unsigned s_tot[32][33];
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (blockIdx.x >= col_ctas) { // the other role of this launch: 32 image rows per CTA, one warp each
        ii_rowprefix(p, (blockIdx.x - col_ctas) * 32 + wid, lane);
        return;
    }
    const unsigned per = (p.nbands + 31) / 32, b0 = wid * per, b1 = min(p.nbands, b0 + per);
    uint32_t      *m = p.colcarry + (size_t)blockIdx.x * 32 + lane;
    // the first 16 bands of the run stay in registers between the two phases (up to 512 bands = 16384 rows at
    // R = 32 that is the whole run); longer runs re-read the rest, which the first phase left in L2
    unsigned t[16], sum = 0;
#pragma unroll
    for (unsigned k = 0; k < 16; k++) {
        t[k] = b0 + k < b1 ? m[(size_t)(b0 + k) * p.wp] : 0u;
        sum += t[k];
    }
    for (unsigned b = b0 + 16; b < b1; b++)
        sum += m[(size_t)b * p.wp];
    s_tot[wid][lane] = sum;
    __syncthreads();
    unsigned run = 0;
    for (unsigned k = 0; k < wid; k++)
        run += s_tot[k][lane];
#pragma unroll
    for (unsigned k = 0; k < 16; k++) {
        if (b0 + k < b1)
            m[(size_t)(b0 + k) * p.wp] = run;
        run += t[k];
    }
    for (unsigned b = b0 + 16; b < b1; b++) {
        const unsigned x    = m[(size_t)b * p.wp];
        m[(size_t)b * p.wp] = run;
        run += x;
    }
}

// Row carries: exclusive prefix along the blocks of every image row, one warp per row, four blocks per lane (one
// round covers 128 blocks = 16384 columns).
__device__ __forceinline__ void ii_rowprefix(const IIParams &p, unsigned r, unsigned lane)
{
    if (r >= p.h)
        return;
    uint32_t *m   = p.rowcarry + (size_t)r * p.nblk;
    unsigned  run = 0;
    for (unsigned j0 = 0; j0 < p.nblk; j0 += 128) {
        const unsigned j = j0 + 4 * lane;
        unsigned       t[4];
#pragma unroll
        for (unsigned k = 0; k < 4; k++)
            t[k] = j + k < p.nblk ? m[j + k] : 0u;
        const unsigned tot = (t[0] + t[1]) + (t[2] + t[3]);
        unsigned       inc = tot;
#pragma unroll
        for (unsigned d = 1; d < 32; d <<= 1) {
            const unsigned n = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d)
                inc += n;
        }
        unsigned e = run + inc - tot;
#pragma unroll
        for (unsigned k = 0; k < 4; k++) {
            if (j + k < p.nblk)
                m[j + k] = e;
            e += t[k];
        }
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
}

template <bool kAligned>
__global__ void __launch_bounds__(kIIWarps * 32) ii_write_kernel(const IIParams p)
{
    const unsigned lane = lane_id();
    const unsigned tile = blockIdx.x * kIIWarps + (threadIdx.x >> 5);
    if (tile >= p.ntiles)
        return;
    const unsigned b = tile / p.nblk, j = tile - b * p.nblk;
    const unsigned r0 = b * p.R, rows = min(p.R, p.h - r0), c = j * kIICols + 4 * lane;
    const unsigned left = lane < rows ? p.rowcarry[(size_t)(r0 + lane) * p.nblk + j] : 0u;
    uint4          acc  = *reinterpret_cast<const uint4 *>(p.colcarry + (size_t)b * p.wp + c);

    auto emit = [&](unsigned row, unsigned base, unsigned q0, unsigned q1, unsigned q2, unsigned q3) {
        acc.x += base + q0;
        acc.y += base + q1;
        acc.z += base + q2;
        acc.w += base + q3;
        if (row >= rows)
            return;
        uint32_t *o = p.out + (size_t)(r0 + row) * p.w + c;
        if (kAligned) {
            if (c < p.w)
                st_stream_u4(o, acc);
        } else {
            if (c < p.w) o[0] = acc.x;
            if (c + 1 < p.w) o[1] = acc.y;
            if (c + 2 < p.w) o[2] = acc.z;
            if (c + 3 < p.w) o[3] = acc.w;
        }
    };
    // kIIBatch rows per iteration (all loads first); two rows share one 32-bit warp scan: their lane totals
    // (<= 1020 each, <= 32640 per warp) sit in the two 16-bit halves
    for (unsigned i0 = 0; i0 < rows; i0 += kIIBatch) {
        unsigned x[kIIBatch];
#pragma unroll
        for (unsigned k = 0; k < kIIBatch; k++)
            x[k] = i0 + k < rows ? load_px4<kAligned>(p, r0 + i0 + k, c) : 0u;
#pragma unroll
        for (unsigned k = 0; k < kIIBatch; k += 2) {
            const unsigned i = i0 + k, xa = x[k], xb = x[k + 1];
            const unsigned a0 = xa & 0xff, a1 = a0 + ((xa >> 8) & 0xff), a2 = a1 + ((xa >> 16) & 0xff), a3 = a2 + (xa >> 24);
            const unsigned b0 = xb & 0xff, b1 = b0 + ((xb >> 8) & 0xff), b2 = b1 + ((xb >> 16) & 0xff), b3 = b2 + (xb >> 24);
            const unsigned tot = a3 | (b3 << 16);
            unsigned       inc = tot;
#pragma unroll
            for (unsigned d = 1; d < 32; d <<= 1) {
                const unsigned n = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d)
                    inc += n;
            }
            const unsigned exc = inc - tot;
            emit(i, (exc & 0xffffu) + __shfl_sync(0xffffffffu, left, i & 31), a0, a1, a2, a3);
            emit(i + 1, (exc >> 16) + __shfl_sync(0xffffffffu, left, (i + 1) & 31), b0, b1, b2, b3);
        }
    }
}

} // namespace b200

using namespace b200;

int g_ii_rows = 0; // tuning knob (B200_II_ROWS = 8 / 16 / 32): rows per band; 0 = by image size

// d_out[r * width + c] = sum of d_img[r' * width + c'] over r' <= r, c' <= c (u32, wraps).  Any width, height >= 0;
// rows are `width` bytes apart (no padding), as in the sample.  d_out may not alias d_img.
extern "C" int b200_integral_image_u8_async(uint32_t *d_out, const uint8_t *d_img, size_t width, size_t height,
                                            b200_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    if (width == 0 || height == 0)
        return 0;
    if (width > (1u << 24) || height > (1u << 24) || width * height >= (1ull << 32))
        return (int)cudaErrorInvalidValue;
    int        st = 0;
    DeviceCtx *c  = ctx(&st);
    if (!c)
        return st;
    static const int env_rows = [] {
        const char *e = getenv("B200_II_ROWS");
        return e ? atoi(e) : 0;
    }();
    IIParams p;
    p.img  = d_img;
    p.out  = d_out;
    p.w    = (unsigned)width;
    p.h    = (unsigned)height;
    p.nblk = (p.w + kIICols - 1) / kIICols;
    p.wp   = p.nblk * kIICols;
    // tall bands amortise the carries (the carry matrix is 1/R of the output); small images take shorter bands
    // so that every SM still gets a few warps
    unsigned R = 32;
    const unsigned want_tiles = (unsigned)c->num_sms * 16;
    while (R > 8 && ((p.h + R - 1) / R) * p.nblk < want_tiles)
        R >>= 1;
    const int forced = g_ii_rows ? g_ii_rows : env_rows;
    if (forced == 8 || forced == 16 || forced == 32)
        R = (unsigned)forced;
    p.R      = R;
    p.nbands = (p.h + R - 1) / R;
    p.ntiles = p.nbands * p.nblk;
    const size_t col_bytes = (size_t)p.nbands * p.wp * 4, row_bytes = (size_t)p.h * p.nblk * 4;
    B200_TRY((cudaError_t)ensure(c->image, col_bytes + row_bytes + 256, stream));
    p.colcarry = reinterpret_cast<uint32_t *>(c->image.ptr);
    p.rowcarry = reinterpret_cast<uint32_t *>(reinterpret_cast<char *>(c->image.ptr) + ((col_bytes + 255) & ~(size_t)255));
    const bool aligned = (p.w % 4 == 0) && (((uintptr_t)d_img) % 4 == 0) && (((uintptr_t)d_out) % 16 == 0);
    const unsigned grid = (p.ntiles + kIIWarps - 1) / kIIWarps;
    const bool aligned16 = (p.w % 16 == 0) && (((uintptr_t)d_img) % 16 == 0);
    if (aligned16) {
        const unsigned t16 = p.nbands * ((p.nblk + 3) / 4);
        ii_sums16_kernel<<<(t16 + kIIWarps - 1) / kIIWarps, kIIWarps * 32, 0, stream>>>(p);
    } else if (aligned)
        ii_sums_kernel<true><<<grid, kIIWarps * 32, 0, stream>>>(p);
    else
        ii_sums_kernel<false><<<grid, kIIWarps * 32, 0, stream>>>(p);
    B200_TRY(cudaGetLastError());
    ii_prefix_kernel<<<p.wp / 32 + (p.h + 31) / 32, 1024, 0, stream>>>(p, p.wp / 32);
    B200_TRY(cudaGetLastError());
    count_launch(2);
    {
        const int r = b200_scan_inclusive_async(p.colcarry, p.colcarry, p.nbands, p.wp, nullptr, stream_);
        if (r)
            return r;
    }
    if (aligned)
        ii_write_kernel<true><<<grid, kIIWarps * 32, 0, stream>>>(p);
    else
        ii_write_kernel<false><<<grid, kIIWarps * 32, 0, stream>>>(p);
    count_launch();
    B200_TRY(cudaGetLastError());
    return 0;
}
