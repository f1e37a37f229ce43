// Host-I/O edge of the filter path: bring whatever a decoder produced into the packed 8-bit BGR
// layout the filters take.  Device twin of convert_to_3ch (sample/benchmark/main.cpp:91-103: GRAY2BGR replicates the
// channel, BGRA2BGR drops alpha, 3 channels are copied) plus the R<->B swap a PNG/PPM decoder (RGB order) needs to reach
// OpenCV's BGR, and of convert_to_u8 (sample/gradient/main.cpp:9-18) for writing a gradient magnitude as an image.
// Pure streaming kernels: 4 pixels per thread, 32-bit loads and stores when the buffers allow it.
#include <cuda_runtime.h>

#include <cstdint>

#include "vip_internal.h"

namespace vip {
namespace {

constexpr int kConvThreads = 256;

template <int SRC_C, bool SWAP>
__device__ __forceinline__ void convert_px(const std::uint8_t* __restrict__ s, std::uint8_t* __restrict__ d) {
    if constexpr (SRC_C == 1) {
        d[0] = d[1] = d[2] = s[0];
    } else {
        d[0] = SWAP ? s[2] : s[0];
        d[1] = s[1];
        d[2] = SWAP ? s[0] : s[2];
    }
}

// one thread = 4 consecutive pixels = 4*SRC_C bytes in (1, 3 or 4 words), 12 bytes out (3 words)
template <int SRC_C, bool SWAP>
__global__ void __launch_bounds__(kConvThreads) convert_to_bgr_kernel(const std::uint8_t* __restrict__ src,
                                                                     std::uint8_t* __restrict__ dst, const size_t n_px,
                                                                     const bool aligned) {
    const size_t p0 = (static_cast<size_t>(blockIdx.x) * kConvThreads + threadIdx.x) * 4;
    if (p0 >= n_px) return;
    if (aligned && p0 + 4 <= n_px) {
        std::uint8_t in[4 * SRC_C], out[12];
        const std::uint32_t* s32 = reinterpret_cast<const std::uint32_t*>(src + p0 * SRC_C);
#pragma unroll
        for (int i = 0; i < SRC_C; ++i) {
            const std::uint32_t v = __ldg(s32 + i);
            in[4 * i + 0] = v & 0xff;
            in[4 * i + 1] = (v >> 8) & 0xff;
            in[4 * i + 2] = (v >> 16) & 0xff;
            in[4 * i + 3] = v >> 24;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) convert_px<SRC_C, SWAP>(in + q * SRC_C, out + q * 3);
        std::uint32_t* d32 = reinterpret_cast<std::uint32_t*>(dst + p0 * 3);
#pragma unroll
        for (int i = 0; i < 3; ++i)
            d32[i] = out[4 * i] | (out[4 * i + 1] << 8) | (out[4 * i + 2] << 16) |
                     (static_cast<std::uint32_t>(out[4 * i + 3]) << 24);
    } else {
        for (size_t p = p0; p < n_px && p < p0 + 4; ++p) convert_px<SRC_C, SWAP>(src + p * SRC_C, dst + p * 3);
    }
}

// dst = saturate_u8(round_half_even(src * 255 / max)) -- cv::Mat::convertTo(CV_8UC1) of img * 255 / max
__global__ void __launch_bounds__(kConvThreads) scale_to_u8_kernel(const float* __restrict__ src,
                                                                  std::uint8_t* __restrict__ dst, const size_t n,
                                                                  const float scale_num, const float max_value) {
    const size_t i = static_cast<size_t>(blockIdx.x) * kConvThreads + threadIdx.x;
    if (i >= n) return;
    const float v = __fdiv_rn(__fmul_rn(__ldg(src + i), scale_num), max_value);
    const int q = __float2int_rn(v);
    dst[i] = static_cast<std::uint8_t>(q < 0 ? 0 : (q > 255 ? 255 : q));
}

template <int SRC_C>
cudaError_t launch_convert_c(const std::uint8_t* src, std::uint8_t* dst, size_t n_px, bool swap, cudaStream_t stream) {
    const bool aligned = ((reinterpret_cast<std::uintptr_t>(src) | reinterpret_cast<std::uintptr_t>(dst)) & 3u) == 0;
    const unsigned grid = static_cast<unsigned>((n_px + 4 * kConvThreads - 1) / (4 * kConvThreads));
    if (swap)
        convert_to_bgr_kernel<SRC_C, true><<<grid, kConvThreads, 0, stream>>>(src, dst, n_px, aligned);
    else
        convert_to_bgr_kernel<SRC_C, false><<<grid, kConvThreads, 0, stream>>>(src, dst, n_px, aligned);
    count_launch();
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_convert_to_bgr(const std::uint8_t* src, int src_channels, std::uint8_t* dst, int width, int height,
                                  int swap_rb, cudaStream_t stream) {
    const size_t n_px = static_cast<size_t>(width) * height;
    if (n_px == 0) return cudaSuccess;
    switch (src_channels) {
        case 1: return launch_convert_c<1>(src, dst, n_px, false, stream);
        case 3: return launch_convert_c<3>(src, dst, n_px, swap_rb != 0, stream);
        case 4: return launch_convert_c<4>(src, dst, n_px, swap_rb != 0, stream);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_scale_to_u8(const float* src, std::uint8_t* dst, size_t n, float max_value, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    scale_to_u8_kernel<<<static_cast<unsigned>((n + kConvThreads - 1) / kConvThreads), kConvThreads, 0, stream>>>(
        src, dst, n, 255.f, max_value);
    count_launch();
    return cudaGetLastError();
}

}  // namespace vip
