// Border-replicated integral image on the GPU (sm_100a).  GPU twin of the reference's CPU-only
// BorderReplicatedIntegralImage<uint8_t,3> (include/cpp/border_replicated_integral_image.hpp:6-85): a
// (H+2r+1) x (W+2r+1) x 3 table over the clamp-padded image, first row / column zero; get(x0,y0,x1,y1) is the
// inclusive box sum from four corners (:34-47).  All sums are uint32 modulo 2^32: differences of corners are
// exact for any image size (the reference's int32 table overflows -- UB -- beyond ~4K).
//
//  K1 row pass    one CTA per padded row; 256-pixel chunks are scanned with warp shuffles (__shfl_up_sync
//                 inclusive scan inside each warp, one shared-memory hop across the 8 warps) and a running carry,
//                 reading the clamped source row once and writing coalesced 12-byte records;
//  K2 column pass 64-row bands, one thread per (column, channel): band-local running sums, band totals to a
//                 small aux array;
//  K3 band fix-up adds the exclusive prefix of the band totals.
// Traffic: 3 B read + 12 B written (K1) + 24 B (K2) + 24 B (K3) per padded cell.
#include <cuda_runtime.h>

#include <cstdint>

#include "vip_internal.h"

namespace vip {
namespace {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

constexpr int kScanThreads = 256;
constexpr int kBandRows = 64;

__global__ void __launch_bounds__(kScanThreads) integral_rows_kernel(const std::uint8_t* __restrict__ src,
                                                                    std::uint32_t* __restrict__ table, const int W,
                                                                    const int H, const int r) {
    const int tw = W + 2 * r + 1;
    const int Y = blockIdx.x;  // padded row, 0 .. H+2r
    std::uint32_t* trow = table + static_cast<size_t>(Y) * tw * 3;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (Y == 0) {
        for (int i = tid; i < tw * 3; i += kScanThreads) trow[i] = 0u;
        return;
    }
    const std::uint8_t* srow = src + static_cast<size_t>(clampi(Y - 1 - r, 0, H - 1)) * W * 3;
    __shared__ std::uint32_t warp_tot[3][kScanThreads / 32];
    __shared__ std::uint32_t carry_s[3];
    if (tid < 3) {
        carry_s[tid] = 0u;
        trow[tid] = 0u;  // column 0
    }
    __syncthreads();
    for (int X0 = 1; X0 < tw; X0 += kScanThreads) {
        const int X = X0 + tid;
        std::uint32_t v0 = 0, v1 = 0, v2 = 0;
        if (X < tw) {
            const std::uint8_t* p = srow + clampi(X - 1 - r, 0, W - 1) * 3;
            v0 = __ldg(p);
            v1 = __ldg(p + 1);
            v2 = __ldg(p + 2);
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const std::uint32_t a0 = __shfl_up_sync(0xffffffffu, v0, d);
            const std::uint32_t a1 = __shfl_up_sync(0xffffffffu, v1, d);
            const std::uint32_t a2 = __shfl_up_sync(0xffffffffu, v2, d);
            if (lane >= d) {
                v0 += a0;
                v1 += a1;
                v2 += a2;
            }
        }
        if (lane == 31) {
            warp_tot[0][warp] = v0;
            warp_tot[1][warp] = v1;
            warp_tot[2][warp] = v2;
        }
        __syncthreads();
        std::uint32_t o0 = carry_s[0], o1 = carry_s[1], o2 = carry_s[2];
        for (int w = 0; w < warp; ++w) {
            o0 += warp_tot[0][w];
            o1 += warp_tot[1][w];
            o2 += warp_tot[2][w];
        }
        v0 += o0;
        v1 += o1;
        v2 += o2;
        if (X < tw) {
            trow[X * 3 + 0] = v0;
            trow[X * 3 + 1] = v1;
            trow[X * 3 + 2] = v2;
        }
        __syncthreads();
        if (tid == kScanThreads - 1) {
            carry_s[0] = v0;
            carry_s[1] = v1;
            carry_s[2] = v2;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kScanThreads) integral_cols_kernel(std::uint32_t* __restrict__ table,
                                                                    std::uint32_t* __restrict__ aux, const int pitch,
                                                                    const int th) {
    const int col = blockIdx.x * kScanThreads + threadIdx.x;  // column*3 + channel
    if (col >= pitch) return;
    const int band = blockIdx.y;
    const int y0 = band * kBandRows, y1 = min(y0 + kBandRows, th);
    std::uint32_t acc = 0;
    std::uint32_t* p = table + static_cast<size_t>(y0) * pitch + col;
    for (int y = y0; y < y1; ++y, p += pitch) {
        acc += *p;
        *p = acc;
    }
    aux[static_cast<size_t>(band) * pitch + col] = acc;
}

__global__ void __launch_bounds__(kScanThreads) integral_fix_kernel(std::uint32_t* __restrict__ table,
                                                                   const std::uint32_t* __restrict__ aux,
                                                                   const int pitch, const int th) {
    const int col = blockIdx.x * kScanThreads + threadIdx.x;
    if (col >= pitch) return;
    const int band = blockIdx.y + 1;
    std::uint32_t off = 0;
    for (int b = 0; b < band; ++b) off += __ldg(aux + static_cast<size_t>(b) * pitch + col);
    const int y0 = band * kBandRows, y1 = min(y0 + kBandRows, th);
    std::uint32_t* p = table + static_cast<size_t>(y0) * pitch + col;
    for (int y = y0; y < y1; ++y, p += pitch) *p += off;
}

__device__ __forceinline__ void box3(const std::uint32_t* __restrict__ table, const int tw, const int r, const int x0,
                                     const int y0, const int x1, const int y1, std::uint32_t out[3]) {
    const std::uint32_t* a = table + (static_cast<size_t>(y1 + r + 1) * tw + (x1 + r + 1)) * 3;
    const std::uint32_t* b = table + (static_cast<size_t>(y1 + r + 1) * tw + (x0 + r)) * 3;
    const std::uint32_t* c = table + (static_cast<size_t>(y0 + r) * tw + (x1 + r + 1)) * 3;
    const std::uint32_t* d = table + (static_cast<size_t>(y0 + r) * tw + (x0 + r)) * 3;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) out[ch] = __ldg(a + ch) - __ldg(b + ch) - __ldg(c + ch) + __ldg(d + ch);
}

__global__ void integral_get_kernel(const std::uint32_t* __restrict__ table, const int W, const int r,
                                    const std::int32_t* __restrict__ rects, const int n,
                                    std::uint32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    std::uint32_t v[3];
    box3(table, W + 2 * r + 1, r, rects[4 * i], rects[4 * i + 1], rects[4 * i + 2], rects[4 * i + 3], v);
    out[3 * i] = v[0];
    out[3 * i + 1] = v[1];
    out[3 * i + 2] = v[2];
}

__global__ void integral_box_mean_kernel(const std::uint32_t* __restrict__ table, float* __restrict__ mean,
                                         const int W, const int H, const int r) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    std::uint32_t v[3];
    box3(table, W + 2 * r + 1, r, x - r, y - r, x + r, y + r, v);
    const float kk = static_cast<float>((2 * r + 1) * (2 * r + 1));
    float* m = mean + (static_cast<size_t>(y) * W + x) * 3;
    m[0] = __fdiv_rn(static_cast<float>(v[0]), kk);
    m[1] = __fdiv_rn(static_cast<float>(v[1]), kk);
    m[2] = __fdiv_rn(static_cast<float>(v[2]), kk);
}

}  // namespace

cudaError_t launch_integral_u8c3(const std::uint8_t* src, std::uint32_t* table, int width, int height, int radius,
                                 cudaStream_t stream) {
    const int tw = width + 2 * radius + 1, th = height + 2 * radius + 1;
    const int pitch = tw * 3;
    const int nbands = (th + kBandRows - 1) / kBandRows;
    std::uint32_t* aux = nullptr;
    cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(&aux), static_cast<size_t>(nbands) * pitch * 4, stream);
    if (e != cudaSuccess) return e;
    integral_rows_kernel<<<th, kScanThreads, 0, stream>>>(src, table, width, height, radius);
    const dim3 grid((pitch + kScanThreads - 1) / kScanThreads, nbands);
    integral_cols_kernel<<<grid, kScanThreads, 0, stream>>>(table, aux, pitch, th);
    count_launch(2);
    if (nbands > 1) {
        const dim3 grid2(grid.x, nbands - 1);
        integral_fix_kernel<<<grid2, kScanThreads, 0, stream>>>(table, aux, pitch, th);
        count_launch();
    }
    e = cudaGetLastError();
    cudaFreeAsync(aux, stream);
    return e;
}

cudaError_t launch_integral_get(const std::uint32_t* table, int width, int height, int radius,
                                const std::int32_t* rects, int n, std::uint32_t* out, cudaStream_t stream) {
    (void)height;
    if (n <= 0) return cudaSuccess;
    integral_get_kernel<<<(n + 255) / 256, 256, 0, stream>>>(table, width, radius, rects, n, out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_integral_box_mean(const std::uint32_t* table, float* mean, int width, int height, int radius,
                                     cudaStream_t stream) {
    const dim3 block(32, 8);
    const dim3 grid((width + 31) / 32, (height + 7) / 8);
    integral_box_mean_kernel<<<grid, block, 0, stream>>>(table, mean, width, height, radius);
    count_launch();
    return cudaGetLastError();
}

}  // namespace vip
