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
//   ii_prefix     exclusive prefix of the column sums down the bands, exclusive prefix of the row sums along the
//                 blocks (both matrices are 1/R resp. 1/128 of the image).
//   scan          the library's own batched inclusive scan along each band row of the column carries: after it
//                 carry[b][c] = out[first row of band b - 1][c], the full integral above the band.
//   ii_write      per tile: acc(c) = carry[b][c]; for each row: 4-pixel prefix, warp scan (two rows per scan, 16
//                 bits each), + the row's carry from the blocks to its left, acc += it, one 16-byte store per lane.
// Rows whose width or base address is not a multiple of 4 bytes / 16 bytes take the guarded scalar variant.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr unsigned kIICols  = 128; // columns per tile
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

// blocks [0, wp/128): one thread per column, exclusive prefix down the bands; the rest: one thread per image row,
// exclusive prefix along the blocks.  Both loops only chain adds; the loads are independent.
__global__ void __launch_bounds__(128) ii_prefix_kernel(const IIParams p)
{
    if (blockIdx.x < p.nblk) {
        uint32_t *m   = p.colcarry + blockIdx.x * 128 + threadIdx.x;
        unsigned  run = 0;
        for (unsigned b0 = 0; b0 < p.nbands; b0 += 8) {
            unsigned t[8];
#pragma unroll
            for (unsigned k = 0; k < 8; k++)
                t[k] = b0 + k < p.nbands ? m[(size_t)(b0 + k) * p.wp] : 0u;
#pragma unroll
            for (unsigned k = 0; k < 8; k++) {
                if (b0 + k < p.nbands)
                    m[(size_t)(b0 + k) * p.wp] = run;
                run += t[k];
            }
        }
    } else {
        const unsigned r = (blockIdx.x - p.nblk) * 128 + threadIdx.x;
        if (r >= p.h)
            return;
        uint32_t *m   = p.rowcarry + (size_t)r * p.nblk;
        unsigned  run = 0;
        for (unsigned j0 = 0; j0 < p.nblk; j0 += 8) {
            unsigned t[8];
#pragma unroll
            for (unsigned k = 0; k < 8; k++)
                t[k] = j0 + k < p.nblk ? m[j0 + k] : 0u;
#pragma unroll
            for (unsigned k = 0; k < 8; k++) {
                if (j0 + k < p.nblk)
                    m[j0 + k] = run;
                run += t[k];
            }
        }
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

    // two rows per iteration: their lane totals (<= 1020 each, <= 32640 per warp) share one 32-bit warp scan
    for (unsigned i = 0; i < rows; i += 2) {
        const bool     two = i + 1 < rows;
        const unsigned xa = load_px4<kAligned>(p, r0 + i, c);
        const unsigned xb = two ? load_px4<kAligned>(p, r0 + i + 1, c) : 0u;
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
        const unsigned exc   = inc - tot;
        const unsigned baseA = (exc & 0xffffu) + __shfl_sync(0xffffffffu, left, i);
        const unsigned baseB = (exc >> 16) + __shfl_sync(0xffffffffu, left, (i + 1) & 31);
        acc.x += baseA + a0;
        acc.y += baseA + a1;
        acc.z += baseA + a2;
        acc.w += baseA + a3;
        uint32_t *o = p.out + (size_t)(r0 + i) * p.w + c;
        if (kAligned) {
            if (c < p.w)
                st_stream_u4(o, acc);
        } else {
            if (c < p.w) o[0] = acc.x;
            if (c + 1 < p.w) o[1] = acc.y;
            if (c + 2 < p.w) o[2] = acc.z;
            if (c + 3 < p.w) o[3] = acc.w;
        }
        if (two) {
            acc.x += baseB + b0;
            acc.y += baseB + b1;
            acc.z += baseB + b2;
            acc.w += baseB + b3;
            o += p.w;
            if (kAligned) {
                if (c < p.w)
                    st_stream_u4(o, acc);
            } else {
                if (c < p.w) o[0] = acc.x;
                if (c + 1 < p.w) o[1] = acc.y;
                if (c + 2 < p.w) o[2] = acc.z;
                if (c + 3 < p.w) o[3] = acc.w;
            }
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
    if (aligned)
        ii_sums_kernel<true><<<grid, kIIWarps * 32, 0, stream>>>(p);
    else
        ii_sums_kernel<false><<<grid, kIIWarps * 32, 0, stream>>>(p);
    B200_TRY(cudaGetLastError());
    ii_prefix_kernel<<<p.nblk + (p.h + 127) / 128, 128, 0, stream>>>(p);
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
