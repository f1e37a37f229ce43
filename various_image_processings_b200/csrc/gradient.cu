// Gradient magnitude for sm_100a.  Replaces compute_gradient_kernel (u8 / f32) and cuda_gradient_impl
// (src/gradient_impl.cu:7-112): dst = sqrtf(sum_c (I(x+1,y)-I(x-1,y))^2 + sum_c (I(x,y+1)-I(x,y-1))^2) with
// clamped coordinates and no 1/2 factor.  The stage is HBM-bound (3 B in + 4 B out per pixel for BGR):
//   * one CTA owns a 1024-column x 16-row band; three source rows are kept in a rolling shared-memory window
//     so every source byte is read from global memory once per CTA row (plus 2 rows of halo per band);
//   * one thread produces 4 adjacent outputs and stores them as one float4 (coalesced 512 B per warp).
// u8: integer differences and squares (<= 6*255^2, exact) then the correctly rounded sqrtf -> bit-exact.
// f32: del_x and del_y are accumulated with fmaf per channel in channel order, then added, like the
// reference's nvcc build (diff += d*d contracts to FFMA).
#include <cuda_runtime.h>

#include <cstdint>

#include "vip_internal.h"

namespace vip {
namespace {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

constexpr int kGradThreads = 256;
constexpr int kGradCols = 4 * kGradThreads;  // outputs per CTA row segment
constexpr int kGradRowsPerCta = 16;

// Shared rows hold pixels [X0-1, X0+kGradCols+1) x C bytes, clamp-replicated at the image border.
template <int C>
__global__ void __launch_bounds__(kGradThreads) gradient_u8_kernel(const std::uint8_t* __restrict__ src,
                                                                  float* __restrict__ dst, const int W, const int H,
                                                                  const int src_y0, const int dst_y0, const int out_y0,
                                                                  const int out_rows) {
    constexpr int kRowBytes = (kGradCols + 2) * C;
    constexpr int kRowPitch = (kRowBytes + 15) & ~15;
    __shared__ __align__(16) std::uint8_t rows[3][kRowPitch];

    const int X0 = blockIdx.x * kGradCols;
    const int Y0 = out_y0 + blockIdx.y * kGradRowsPerCta;
    const int Y1 = min(Y0 + kGradRowsPerCta, out_y0 + out_rows);
    const int tid = threadIdx.x;
    const int ncols = min(kGradCols, W - X0);

    auto load_row = [&](int slot, int y) {
        const int yy = clampi(y, 0, H - 1) - src_y0;
        const std::uint8_t* g = src + static_cast<size_t>(yy) * W * C;
        // pixel X0-1 .. X0+ncols (inclusive), clamped
        const int nbytes = (ncols + 2) * C;
        for (int b = tid; b < nbytes; b += kGradThreads) {
            const int px = b / C, ch = b - px * C;
            const int gx = clampi(X0 - 1 + px, 0, W - 1);
            rows[slot][b] = __ldg(g + static_cast<size_t>(gx) * C + ch);
        }
    };

    load_row(0, Y0 - 1);
    load_row(1, Y0);
    for (int y = Y0; y < Y1; ++y) {
        const int up = (y - Y0) % 3, mid = (y - Y0 + 1) % 3, dn = (y - Y0 + 2) % 3;
        load_row(dn, y + 1);
        __syncthreads();
        const int x = 4 * tid;
        if (x < ncols) {
            float out[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                // smem pixel index of output x+j is x+j+1
                const int p = (x + j + 1) * C;
                int acc = 0;
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const int dx = static_cast<int>(rows[mid][p + C + c]) - static_cast<int>(rows[mid][p - C + c]);
                    const int dy = static_cast<int>(rows[dn][p + c]) - static_cast<int>(rows[up][p + c]);
                    acc += dx * dx + dy * dy;
                }
                out[j] = sqrtf(static_cast<float>(acc));
            }
            float* d = dst + static_cast<size_t>(y - dst_y0) * W + X0 + x;
            if (x + 4 <= ncols && (reinterpret_cast<std::uintptr_t>(d) & 15u) == 0) {
                *reinterpret_cast<float4*>(d) = make_float4(out[0], out[1], out[2], out[3]);
            } else {
                for (int j = 0; j < 4 && x + j < ncols; ++j) d[j] = out[j];
            }
        }
        __syncthreads();
    }
}

// Fast path for packed BGR rows that are 4-byte aligned (width % 4 == 0, aligned base): one thread produces 4 adjacent
// outputs of one row from 32-bit words -- 5 words of its own row (pixels x-1 .. x+4) and 3 words of the rows above and
// below -- and stores one float4.  Pixels are cut out of the words with funnel shifts; the squared differences are
// VABSDIFF4 + IDP4A (exact integers).  A CTA covers 128 x 8 outputs, so the rows above/below hit L1.  HBM-bound:
// 3 B in + 4 B out per pixel.
__device__ __forceinline__ int sq3(const std::uint32_t a, const std::uint32_t b) {
    const std::uint32_t d = __vabsdiffu4(a, b) & 0x00ffffffu;
    return __dp4a(d, d, 0u);
}

// A thread produces 4 columns x kFastRows rows: the kFastRows + 2 source rows it needs are read once as three aligned
// words each and cut into pixels once; every row then serves as the upper, the centre and the lower neighbour.
#ifndef VIP_GRAD_ROWS
#define VIP_GRAD_ROWS 4
#endif
constexpr int kFastRows = VIP_GRAD_ROWS;   // -DVIP_GRAD_ROWS=8: A/B build
__global__ void __launch_bounds__(256) gradient_u8c3_fast_kernel(const std::uint8_t* __restrict__ src,
                                                                 float* __restrict__ dst, const int W, const int H,
                                                                 const int src_y0, const int src_rows, const int dst_y0,
                                                                 const int out_y0, const int out_rows) {
    const int x = 4 * (blockIdx.x * 32 + threadIdx.x);
    const int y0 = out_y0 + (blockIdx.y * 8 + threadIdx.y) * kFastRows;
    // launched through launch_pdl: the kernel is 10 - 50 us long, so the launch gap between it and its predecessor is a
    // visible share of it; the previous grid has completed and flushed from pdl_wait() on
    pdl_wait();
    pdl_release();
    if (x >= W || y0 >= out_y0 + out_rows) return;
    const int wpr = (W * 3) >> 2;  // words per row
    auto row_ptr = [&](int yy) {
        const int r = clampi(clampi(yy, 0, H - 1), src_y0, src_y0 + src_rows - 1) - src_y0;
        return reinterpret_cast<const std::uint32_t*>(src) + static_cast<size_t>(r) * wpr + (x >> 2) * 3;
    };
    // pixels x .. x+3 of rows y0-1 .. y0+kFastRows, plus pixel x-1 / x+4 of the centre rows (clamped)
    std::uint32_t p[kFastRows + 2][4], pm[kFastRows], pp[kFastRows];
#pragma unroll
    for (int j = 0; j < kFastRows + 2; ++j) {
        const std::uint32_t* m = row_ptr(y0 - 1 + j);
        const std::uint32_t w0 = __ldg(m), w1 = __ldg(m + 1), w2 = __ldg(m + 2);
        p[j][0] = w0;
        p[j][1] = __funnelshift_r(w0, w1, 24);
        p[j][2] = __funnelshift_r(w1, w2, 16);
        p[j][3] = w2 >> 8;
        if (j >= 1 && j <= kFastRows) {
            pm[j - 1] = x > 0 ? (__ldg(m - 1) >> 8) : w0;
            pp[j - 1] = x + 4 < W ? __ldg(m + 3) : p[j][3];
        }
    }
#pragma unroll
    for (int i = 0; i < kFastRows; ++i) {
        const int y = y0 + i;
        if (y >= out_y0 + out_rows) break;
        const std::uint32_t* c = p[i + 1];
        const std::uint32_t* u = p[i];
        const std::uint32_t* d = p[i + 2];
        float4 o;
        o.x = sqrtf(static_cast<float>(sq3(c[1], pm[i]) + sq3(d[0], u[0])));
        o.y = sqrtf(static_cast<float>(sq3(c[2], c[0]) + sq3(d[1], u[1])));
        o.z = sqrtf(static_cast<float>(sq3(c[3], c[1]) + sq3(d[2], u[2])));
        o.w = sqrtf(static_cast<float>(sq3(pp[i], c[2]) + sq3(d[3], u[3])));
        *reinterpret_cast<float4*>(dst + static_cast<size_t>(y - dst_y0) * W + x) = o;
    }
}

// Generic channel count (u8) and float input: one thread per output pixel, clamped global reads.
template <typename T>
__global__ void gradient_generic_kernel(const T* __restrict__ src, float* __restrict__ dst, const int W, const int H,
                                        const int C, const int src_y0, const int dst_y0, const int out_y0,
                                        const int out_rows) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = out_y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= out_y0 + out_rows) return;
    const int xm = clampi(x - 1, 0, W - 1), xp = clampi(x + 1, 0, W - 1);
    const int ym = clampi(y - 1, 0, H - 1) - src_y0, yp = clampi(y + 1, 0, H - 1) - src_y0, yc = y - src_y0;
    const T* rc = src + static_cast<size_t>(yc) * W * C;
    const T* ru = src + static_cast<size_t>(ym) * W * C;
    const T* rd = src + static_cast<size_t>(yp) * W * C;
    float r;
    if constexpr (sizeof(T) == 1) {
        int acc = 0;
        for (int c = 0; c < C; ++c) {
            const int dx = static_cast<int>(rc[xp * C + c]) - static_cast<int>(rc[xm * C + c]);
            const int dy = static_cast<int>(rd[x * C + c]) - static_cast<int>(ru[x * C + c]);
            acc += dx * dx + dy * dy;
        }
        r = sqrtf(static_cast<float>(acc));
    } else {
        float gx = 0.f, gy = 0.f;
        for (int c = 0; c < C; ++c) {
            const float dx = __fsub_rn(rc[xp * C + c], rc[xm * C + c]);
            gx = __fmaf_rn(dx, dx, gx);
        }
        for (int c = 0; c < C; ++c) {
            const float dy = __fsub_rn(rd[x * C + c], ru[x * C + c]);
            gy = __fmaf_rn(dy, dy, gy);
        }
        r = sqrtf(__fadd_rn(gx, gy));
    }
    dst[static_cast<size_t>(y - dst_y0) * W + x] = r;
}

}  // namespace

cudaError_t launch_gradient_u8(const std::uint8_t* src, float* dst, int width, int height, int channels, int src_y0,
                               int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream) {
    if (out_rows <= 0 || width <= 0) return cudaSuccess;
    if (channels == 3 && (width & 3) == 0 && (reinterpret_cast<std::uintptr_t>(src) & 3u) == 0 &&
        (reinterpret_cast<std::uintptr_t>(dst) & 15u) == 0) {
        const dim3 block(32, 8);
        const dim3 grid((width / 4 + 31) / 32, (out_rows + 8 * kFastRows - 1) / (8 * kFastRows));
        const cudaError_t e = launch_pdl(gradient_u8c3_fast_kernel, grid, block, 0, stream, src, dst, width, height, src_y0,
                                         src_rows, dst_y0, out_y0, out_rows);
        count_launch();
        return e;
    } else if (channels == 3 || channels == 1) {
        const dim3 grid((width + kGradCols - 1) / kGradCols, (out_rows + kGradRowsPerCta - 1) / kGradRowsPerCta);
        if (channels == 3)
            gradient_u8_kernel<3><<<grid, kGradThreads, 0, stream>>>(src, dst, width, height, src_y0, dst_y0, out_y0,
                                                                    out_rows);
        else
            gradient_u8_kernel<1><<<grid, kGradThreads, 0, stream>>>(src, dst, width, height, src_y0, dst_y0, out_y0,
                                                                    out_rows);
    } else {
        const dim3 block(32, 8);
        const dim3 grid((width + 31) / 32, (out_rows + 7) / 8);
        gradient_generic_kernel<std::uint8_t><<<grid, block, 0, stream>>>(src, dst, width, height, channels, src_y0,
                                                                          dst_y0, out_y0, out_rows);
    }
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_gradient_f32(const float* src, float* dst, int width, int height, int channels,
                                cudaStream_t stream) {
    if (height <= 0 || width <= 0) return cudaSuccess;
    const dim3 block(32, 8);
    const dim3 grid((width + 31) / 32, (height + 7) / 8);
    gradient_generic_kernel<float><<<grid, block, 0, stream>>>(src, dst, width, height, channels, 0, 0, 0, height);
    count_launch();
    return cudaGetLastError();
}

}  // namespace vip
