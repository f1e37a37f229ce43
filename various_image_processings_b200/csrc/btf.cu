// Bilateral-texture-filter stage kernels for sm_100a.  Replace compute_blur_and_rtv_kernel
// (src/bilateral_texture_filter_impl.cu:10-104) and compute_guide_kernel (:106-177).  New design:
//
//  blur / mRTV   Everything that is order-independent is made separable and exact: the k x k box sums of the
//                three channels and the min / max of (b+g+r) are integers (the reference's float running sums
//                of u8 values are exact integers < 2^24, and (b+g+r)/3.f is monotone, so min/max commute with
//                it); the max of the gradient magnitude is a float max.  A horizontal pass over the shared tile
//                writes one 16-byte record per (row, column), a vertical pass folds k records per output.
//                The SUM of the magnitudes is the one quantity whose fp32 rounding depends on the order, and the
//                argmin in the next stage amplifies a 1-ulp difference into a different guide pixel
//                (hard part 3), so it is accumulated in the reference's raster order (ky outer, kx
//                inner), with each lane owning an image row and 4 adjacent outputs so that one shared load
//                feeds up to 4 accumulators.  The final divide is done in fp64 like the reference
//                (its epsilon is a double literal, :8, :103).
//  guide         The k x k argmin with "first in raster order wins" is separable as a lexicographic
//                (value, ky, kx) minimum: a horizontal pass keeps the smallest kx among equal values, the
//                vertical pass the smallest ky.  alpha, the blend (fma(alpha, Bq, (1-alpha)*Bp) + 0.5f, the
//                contraction nvcc applies to the reference), truncation and clamp follow :167-176.
//
// In both kernels a lane owns one image row (odd shared pitches make that bank-conflict free) and a warp
// owns 4 columns x 32 rows; a 256-thread CTA produces 32 x 32 outputs.
#include <cuda_runtime.h>

#include <cstdint>

#include "vip_internal.h"

namespace vip {
namespace {

constexpr int kThreads = 256;
constexpr int kTile = 32;
constexpr int kP = 4;

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

__device__ __forceinline__ std::uint32_t load_packed(const std::uint8_t* __restrict__ p) {
    return static_cast<std::uint32_t>(__ldg(p)) | (static_cast<std::uint32_t>(__ldg(p + 1)) << 8) |
           (static_cast<std::uint32_t>(__ldg(p + 2)) << 16);
}

struct BlurLayout {
    int tw, th, tp, hp, pix_off, mag_off, hs_off, total;
};
__host__ __device__ inline BlurLayout blur_layout(int k) {
    BlurLayout L;
    L.tw = kTile + k - 1;
    L.th = kTile + k - 1;
    L.tp = L.tw | 1;
    L.hp = kTile + 1;
    L.pix_off = 0;
    L.mag_off = L.tp * L.th * 4;
    L.hs_off = (2 * L.tp * L.th * 4 + 15) & ~15;
    L.total = L.hs_off + L.hp * L.th * 16;
    return L;
}

// KT > 0: compile-time window size (unrolled, no predication); KT == 0: runtime k.
template <int KT>
__global__ void __launch_bounds__(kThreads) blur_rtv_kernel(const std::uint8_t* __restrict__ image,
                                                           const float* __restrict__ magnitude,
                                                           float* __restrict__ blurred, float* __restrict__ rtv,
                                                           const int W, const int H, const int k_rt, const int src_y0,
                                                           const int dst_y0, const int out_y0, const int out_rows) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int k = KT > 0 ? KT : k_rt;
    const int r = k / 2;
    const BlurLayout L = blur_layout(k);
    std::uint32_t* const pix = reinterpret_cast<std::uint32_t*>(smem + L.pix_off);
    float* const mag = reinterpret_cast<float*>(smem + L.mag_off);
    uint4* const hs = reinterpret_cast<uint4*>(smem + L.hs_off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int X0 = blockIdx.x * kTile;
    const int Y0 = out_y0 + blockIdx.y * kTile;

    for (int e = tid; e < L.tw * L.th; e += kThreads) {
        const int ey = e / L.tw, ex = e - ey * L.tw;
        const int gx = clampi(X0 - r + ex, 0, W - 1);
        const int gy = clampi(Y0 - r + ey, 0, H - 1) - src_y0;
        const size_t o = static_cast<size_t>(gy) * W + gx;
        pix[ey * L.tp + ex] = load_packed(image + o * 3);
        mag[ey * L.tp + ex] = __ldg(magnitude + o);
    }
    __syncthreads();

    // horizontal pass: record = {B | R<<16, G, min(b+g+r) | max(b+g+r)<<16, max magnitude} over k columns
    for (int e = tid; e < L.th * kTile; e += kThreads) {
        const int row = e / kTile, col = e - row * kTile;
        const std::uint32_t* pr = pix + row * L.tp + col;
        const float* mr = mag + row * L.tp + col;
        std::uint32_t br = 0, g = 0, mn = 0xffffu, mx = 0;
        float mm = 0.f;
#pragma unroll
        for (int j = 0; j < (KT > 0 ? KT : 1); ++j) {
            if (KT == 0) break;
            const std::uint32_t p = pr[j];
            const std::uint32_t s = __vsadu4(p, 0u);
            br += p & 0x00ff00ffu;
            g += (p >> 8) & 0xffu;
            mn = min(mn, s);
            mx = max(mx, s);
            mm = fmaxf(mm, mr[j]);
        }
        if (KT == 0) {
            for (int j = 0; j < k; ++j) {
                const std::uint32_t p = pr[j];
                const std::uint32_t s = __vsadu4(p, 0u);
                br += p & 0x00ff00ffu;
                g += (p >> 8) & 0xffu;
                mn = min(mn, s);
                mx = max(mx, s);
                mm = fmaxf(mm, mr[j]);
            }
        }
        hs[row * L.hp + col] = make_uint4(br, g, mn | (mx << 16), __float_as_uint(mm));
    }
    __syncthreads();

    // vertical pass + raster-order magnitude sum; lane <-> row, 4 adjacent outputs per thread
    const int xl = warp * kP;
    std::uint32_t sb[kP], sg[kP], sr[kP], mn[kP], mx[kP];
    float mm[kP], msum[kP];
#pragma unroll
    for (int i = 0; i < kP; ++i) {
        sb[i] = sg[i] = sr[i] = 0;
        mn[i] = 0xffffu;
        mx[i] = 0;
        mm[i] = 0.f;
        msum[i] = 0.f;
    }
    if (KT > 0) {
#pragma unroll
        for (int ky = 0; ky < (KT > 0 ? KT : 1); ++ky) {
            const uint4* hrow = hs + (lane + ky) * L.hp + xl;
#pragma unroll
            for (int i = 0; i < kP; ++i) {
                const uint4 h = hrow[i];
                sb[i] += h.x & 0xffffu;
                sr[i] += h.x >> 16;
                sg[i] += h.y;
                mn[i] = min(mn[i], h.z & 0xffffu);
                mx[i] = max(mx[i], h.z >> 16);
                mm[i] = fmaxf(mm[i], __uint_as_float(h.w));
            }
            const float* mrow = mag + (lane + ky) * L.tp + xl;
#pragma unroll
            for (int j = 0; j < (KT > 0 ? KT : 1) + kP - 1; ++j) {
                const float v = mrow[j];
#pragma unroll
                for (int i = 0; i < kP; ++i) {
                    if (j - i >= 0 && j - i < KT) msum[i] = __fadd_rn(msum[i], v);
                }
            }
        }
    } else {
        for (int ky = 0; ky < k; ++ky) {
            const uint4* hrow = hs + (lane + ky) * L.hp + xl;
#pragma unroll
            for (int i = 0; i < kP; ++i) {
                const uint4 h = hrow[i];
                sb[i] += h.x & 0xffffu;
                sr[i] += h.x >> 16;
                sg[i] += h.y;
                mn[i] = min(mn[i], h.z & 0xffffu);
                mx[i] = max(mx[i], h.z >> 16);
                mm[i] = fmaxf(mm[i], __uint_as_float(h.w));
            }
            const float* mrow = mag + (lane + ky) * L.tp + xl;
            for (int j = 0; j < k + kP - 1; ++j) {
                const float v = mrow[j];
#pragma unroll
                for (int i = 0; i < kP; ++i) {
                    if (j - i >= 0 && j - i < k) msum[i] = __fadd_rn(msum[i], v);
                }
            }
        }
    }

    const int y = Y0 + lane;
    const int x0 = X0 + xl;
    if (y >= out_y0 + out_rows || x0 >= W) return;
    const float kk = static_cast<float>(k * k);
    float bl[3 * kP], rt[kP];
#pragma unroll
    for (int i = 0; i < kP; ++i) {
        bl[3 * i + 0] = __fdiv_rn(static_cast<float>(sb[i]), kk);
        bl[3 * i + 1] = __fdiv_rn(static_cast<float>(sg[i]), kk);
        bl[3 * i + 2] = __fdiv_rn(static_cast<float>(sr[i]), kk);
        const float imax = __fdiv_rn(static_cast<float>(mx[i]), 3.f);
        const float imin = __fdiv_rn(static_cast<float>(mn[i]), 3.f);
        const float num = __fmul_rn(__fsub_rn(imax, imin), mm[i]);
        rt[i] = static_cast<float>(static_cast<double>(num) / (static_cast<double>(msum[i]) + 1e-9));
    }
    const size_t o = static_cast<size_t>(y - dst_y0) * W + x0;
    float* bp = blurred + o * 3;
    float* rp = rtv + o;
    if (x0 + kP <= W && ((reinterpret_cast<std::uintptr_t>(bp) | reinterpret_cast<std::uintptr_t>(rp)) & 15u) == 0) {
        reinterpret_cast<float4*>(bp)[0] = make_float4(bl[0], bl[1], bl[2], bl[3]);
        reinterpret_cast<float4*>(bp)[1] = make_float4(bl[4], bl[5], bl[6], bl[7]);
        reinterpret_cast<float4*>(bp)[2] = make_float4(bl[8], bl[9], bl[10], bl[11]);
        *reinterpret_cast<float4*>(rp) = make_float4(rt[0], rt[1], rt[2], rt[3]);
    } else {
#pragma unroll
        for (int i = 0; i < kP; ++i) {
            if (x0 + i < W) {
                bp[3 * i + 0] = bl[3 * i + 0];
                bp[3 * i + 1] = bl[3 * i + 1];
                bp[3 * i + 2] = bl[3 * i + 2];
                rp[i] = rt[i];
            }
        }
    }
}

struct GuideLayout {
    int tw, th, tp, hp, hs_off, total;
};
__host__ __device__ inline GuideLayout guide_layout(int k) {
    GuideLayout L;
    L.tw = kTile + k - 1;
    L.th = kTile + k - 1;
    L.tp = L.tw | 1;
    L.hp = kTile + 1;
    L.hs_off = (L.tp * L.th * 4 + 15) & ~15;
    L.total = L.hs_off + L.hp * L.th * 8;
    return L;
}

__global__ void __launch_bounds__(kThreads) guide_kernel(const float* __restrict__ blurred,
                                                        const float* __restrict__ rtv, std::uint8_t* __restrict__ guide,
                                                        const int W, const int H, const int k, const int src_y0,
                                                        const int dst_y0, const int out_y0, const int out_rows) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int r = k / 2;
    const GuideLayout L = guide_layout(k);
    float* const tile = reinterpret_cast<float*>(smem);
    float2* const hs = reinterpret_cast<float2*>(smem + L.hs_off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int X0 = blockIdx.x * kTile;
    const int Y0 = out_y0 + blockIdx.y * kTile;

    for (int e = tid; e < L.tw * L.th; e += kThreads) {
        const int ey = e / L.tw, ex = e - ey * L.tw;
        const int gx = clampi(X0 - r + ex, 0, W - 1);
        const int gy = clampi(Y0 - r + ey, 0, H - 1) - src_y0;
        tile[ey * L.tp + ex] = __ldg(rtv + static_cast<size_t>(gy) * W + gx);
    }
    __syncthreads();

    // horizontal pass: (min value, smallest kx attaining it); strict '<' keeps the first
    for (int e = tid; e < L.th * kTile; e += kThreads) {
        const int row = e / kTile, col = e - row * kTile;
        const float* tr = tile + row * L.tp + col;
        float best = 1e10f;  // src/bilateral_texture_filter_impl.cu:152
        int bj = -1;
        for (int j = 0; j < k; ++j) {
            const float v = tr[j];
            if (best > v) {
                best = v;
                bj = j;
            }
        }
        hs[row * L.hp + col] = make_float2(best, __int_as_float(bj));
    }
    __syncthreads();

    const int xl = warp * kP;
    const int y = Y0 + lane;
    const int x0 = X0 + xl;
    if (y >= out_y0 + out_rows || x0 >= W) return;
    const float sigma_alpha = __fdiv_rn(1.f, static_cast<float>(5 * k));  // :121
    std::uint8_t out[3 * kP];
#pragma unroll
    for (int i = 0; i < kP; ++i) {
        float best = 1e10f;
        int bj = -1, bky = 0;
        const float2* hcol = hs + lane * L.hp + xl + i;
        for (int ky = 0; ky < k; ++ky) {
            const float2 h = hcol[ky * L.hp];
            if (best > h.x) {
                best = h.x;
                bj = __float_as_int(h.y);
                bky = ky;
            }
        }
        const int x = min(x0 + i, W - 1);
        int mx = 0, my = 0;  // the reference's initial coordinates when nothing is below 1e10f
        if (bj >= 0) {
            mx = clampi(x + bj - r, 0, W - 1);
            my = clampi(y + bky - r, 0, H - 1);
        } else {
            best = __ldg(rtv + static_cast<size_t>(clampi(0, 0, H - 1) - src_y0) * W);
        }
        const float rp = tile[(lane + r) * L.tp + xl + i + r];
        const float e = expf(__fmul_rn(sigma_alpha, __fsub_rn(rp, best)));
        const float alpha = __fsub_rn(__fdiv_rn(2.f, __fadd_rn(1.f, e)), 1.f);  // :167
        const float one_m = __fsub_rn(1.f, alpha);
        const float* bq = blurred + (static_cast<size_t>(my - src_y0) * W + mx) * 3;
        const float* bp = blurred + (static_cast<size_t>(y - src_y0) * W + x) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float t = __fadd_rn(__fmaf_rn(alpha, __ldg(bq + c), __fmul_rn(one_m, __ldg(bp + c))), 0.5f);
            out[3 * i + c] = static_cast<std::uint8_t>(clampi(__float2int_rz(t), 0, 255));
        }
    }
    std::uint8_t* gp = guide + (static_cast<size_t>(y - dst_y0) * W + x0) * 3;
    if (x0 + kP <= W && (reinterpret_cast<std::uintptr_t>(gp) & 3u) == 0) {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            reinterpret_cast<std::uint32_t*>(gp)[j] = out[4 * j] | (out[4 * j + 1] << 8) | (out[4 * j + 2] << 16) |
                                                       (static_cast<std::uint32_t>(out[4 * j + 3]) << 24);
        }
    } else {
#pragma unroll
        for (int i = 0; i < kP; ++i) {
            if (x0 + i < W) {
                gp[3 * i + 0] = out[3 * i + 0];
                gp[3 * i + 1] = out[3 * i + 1];
                gp[3 * i + 2] = out[3 * i + 2];
            }
        }
    }
}

}  // namespace

cudaError_t launch_blur_rtv(const std::uint8_t* image, const float* magnitude, float* blurred, float* rtv, int width,
                            int height, int ksize, int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows,
                            cudaStream_t stream) {
    (void)src_rows;
    if (out_rows <= 0 || width <= 0) return cudaSuccess;
    const BlurLayout L = blur_layout(ksize);
    const dim3 grid((width + kTile - 1) / kTile, (out_rows + kTile - 1) / kTile);
    cudaError_t e;
    if (ksize == 9) {
        e = cudaFuncSetAttribute(blur_rtv_kernel<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total);
        if (e != cudaSuccess) return e;
        blur_rtv_kernel<9><<<grid, kThreads, L.total, stream>>>(image, magnitude, blurred, rtv, width, height, ksize,
                                                               src_y0, dst_y0, out_y0, out_rows);
    } else {
        e = cudaFuncSetAttribute(blur_rtv_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total);
        if (e != cudaSuccess) return e;
        blur_rtv_kernel<0><<<grid, kThreads, L.total, stream>>>(image, magnitude, blurred, rtv, width, height, ksize,
                                                               src_y0, dst_y0, out_y0, out_rows);
    }
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_guide(const float* blurred, const float* rtv, std::uint8_t* guide, int width, int height, int ksize,
                         int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream) {
    (void)src_rows;
    if (out_rows <= 0 || width <= 0) return cudaSuccess;
    const GuideLayout L = guide_layout(ksize);
    const dim3 grid((width + kTile - 1) / kTile, (out_rows + kTile - 1) / kTile);
    cudaError_t e = cudaFuncSetAttribute(guide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total);
    if (e != cudaSuccess) return e;
    guide_kernel<<<grid, kThreads, L.total, stream>>>(blurred, rtv, guide, width, height, ksize, src_y0, dst_y0, out_y0,
                                                     out_rows);
    count_launch();
    return cudaGetLastError();
}

}  // namespace vip
