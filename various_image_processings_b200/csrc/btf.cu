// Bilateral-texture-filter stage kernels for sm_100a.  Replace compute_blur_and_rtv_kernel
// (src/bilateral_texture_filter_impl.cu:10-104), compute_guide_kernel (:106-177) and -- fused into the first --
// compute_gradient_kernel (src/gradient_impl.cu:7-77).  New design:
//
//  blur / mRTV   One CTA (256 threads) produces a 64 x 32 tile; a lane owns a COLUMN (global loads/stores are
//                coalesced, shared-memory rows are read conflict-free) and 8 consecutive rows.
//                * the image tile (+ radius + 1 halo) is staged as packed 32-bit pixels; the gradient magnitude of
//                  the tile (+ radius halo) is computed from it in shared memory (VABSDIFF4 + IDP4A: exact integers,
//                  then the correctly rounded sqrtf), so the magnitude image never goes through HBM;
//                * everything order-independent is made separable and exact: k x k box sums of the three channels
//                  and min / max of (b+g+r) are integers (the reference's float running sums of u8 values are exact
//                  integers < 2^24, and (b+g+r)/3.f is monotone, so min/max commute with it), the max of the
//                  magnitude is a float max.  A horizontal pass writes one 16-byte record per (row, column), the
//                  vertical pass folds k records per output while the records slide through registers;
//                * the SUM of the magnitudes is the one quantity whose fp32 rounding depends on the order, and the
//                  argmin of the next stage amplifies a 1-ulp difference into a different guide pixel
//                  (hard part 3): it is accumulated in the reference's raster order (ky outer, kx
//                  inner), one shared-memory row load feeding the chains of up to 8 outputs.  The final divide is
//                  fp64 like the reference (its epsilon is a double literal, :8, :103).
//  guide         The k x k argmin with "first in raster order wins" is separable as a lexicographic
//                (value, ky, kx) minimum: the horizontal pass keeps the smallest kx among equal values, the
//                vertical pass the smallest ky.  alpha, the blend (fma(alpha, Bq, (1-alpha)*Bp) + 0.5f, the
//                contraction nvcc applies to the reference), truncation and clamp follow :167-176.  The tile's
//                output bytes are staged in shared memory and stored as coalesced 32-bit words.
//
// Window sizes 3..15 run the tiled kernels (instantiated per k: the loops are fully unrolled); anything else runs
// a plain one-thread-per-pixel kernel with the same arithmetic.
#include <cuda_runtime.h>

#include <cstdint>

#include "vip_internal.h"

namespace vip {
namespace {

constexpr int kThreads = 256;
constexpr int kTW = 64;    // tile columns: 2 warps side by side (the blur/mRTV horizontal pass assumes 8 groups of 8)
constexpr int kTH = 32;    // tile rows:    4 warp rows x 8 rows per thread
constexpr int kRows = 8;   // consecutive output rows per thread

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

__device__ __forceinline__ std::uint32_t load_packed(const std::uint8_t* __restrict__ p) {
    return static_cast<std::uint32_t>(__ldg(p)) | (static_cast<std::uint32_t>(__ldg(p + 1)) << 8) |
           (static_cast<std::uint32_t>(__ldg(p + 2)) << 16);
}

// Four consecutive packed BGR pixels (12 bytes at any byte alignment) from aligned 32-bit loads (see stencil.cu)
__device__ __forceinline__ void load_packed4(const std::uint8_t* __restrict__ p, std::uint32_t out[4]) {
    const unsigned m = static_cast<unsigned>(reinterpret_cast<std::uintptr_t>(p) & 3u);
    const std::uint32_t* w = reinterpret_cast<const std::uint32_t*>(p - m);
    const std::uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    const std::uint32_t w3 = m ? __ldg(w + 3) : 0u;
    const unsigned sh = 8u * m;
    const std::uint32_t v0 = __funnelshift_r(w0, w1, sh), v1 = __funnelshift_r(w1, w2, sh), v2 = __funnelshift_r(w2, w3, sh);
    out[0] = v0 & 0x00ffffffu;
    out[1] = __funnelshift_r(v0, v1, 24) & 0x00ffffffu;
    out[2] = __funnelshift_r(v1, v2, 16) & 0x00ffffffu;
    out[3] = v2 >> 8;
}

// sum over the three colour bytes of (a_c - b_c)^2, exact
__device__ __forceinline__ int sq_diff(const std::uint32_t a, const std::uint32_t b) {
    const std::uint32_t d = __vabsdiffu4(a, b);
    return __dp4a(d, d, 0u);
}

// x / c for a compile-time constant divisor and an integer-valued x: q0 = x * RN(1/c), one FMA residual, one FMA
// correction -- the correctly rounded IEEE quotient (Markstein), i.e. the bits of __fdiv_rn(x, c), in 3 instructions
// instead of the ~10 + slow-path call of the generic divide.  tests/div_const_check.c verifies it exhaustively for every
// dividend this file can produce (0 .. 255*K*K for K*K, 0 .. 765 for 3).
__device__ __forceinline__ float div_const(const float x, const float c, const float rc) {
    const float q0 = __fmul_rn(x, rc);
    const float r = __fmaf_rn(-c, q0, x);
    return __fmaf_rn(r, rc, q0);
}

template <int K>
struct BlurLayout {
    static constexpr int r = K / 2;
    static constexpr int PW = kTW + 2 * r + 2, PH = kTH + 2 * r + 2;  // packed pixels, radius + 1 halo
    static constexpr int MW = kTW + 2 * r, MH = kTH + 2 * r;          // magnitudes / records, radius halo
    static constexpr int pix_off = 0;
    static constexpr int mag_off = (PW * PH * 4 + 15) & ~15;
    static constexpr int MP = MW + 2;   // magnitude row pitch: rows 10 banks apart (as the pixels), not 8
    static constexpr int hs_off = (mag_off + MP * MH * 4 + 15) & ~15;
    static constexpr int total = hs_off + MH * kTW * 16;
};

struct RowMap {
    int H, src_y0, src_rows;
    // buffer row of full-image row y (clamped to the image, then -- for rows no valid output depends on -- to the strip)
    __device__ __forceinline__ int operator()(int y) const {
        return clampi(clampi(y, 0, H - 1), src_y0, src_y0 + src_rows - 1) - src_y0;
    }
};

// FUSED: magnitude == nullptr, the gradient magnitude is computed from the image tile.
template <int K, bool FUSED>
__global__ void __launch_bounds__(kThreads, 3) blur_rtv_kernel(const std::uint8_t* __restrict__ image,
                                                              const float* __restrict__ magnitude,
                                                              float* __restrict__ blurred, float* __restrict__ rtv,
                                                              const int W, const int H, const int src_y0,
                                                              const int src_rows, const int dst_y0, const int out_y0,
                                                              const int out_rows) {
    using L = BlurLayout<K>;
    constexpr int r = L::r;
    extern __shared__ __align__(16) unsigned char smem[];
    std::uint32_t* const pix = reinterpret_cast<std::uint32_t*>(smem + L::pix_off);
    float* const mag = reinterpret_cast<float*>(smem + L::mag_off);
    uint4* const hs = reinterpret_cast<uint4*>(smem + L::hs_off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int X0 = blockIdx.x * kTW;
    const int Y0 = out_y0 + blockIdx.y * kTH;
    const RowMap rowmap{H, src_y0, src_rows};
    pdl_wait();     // launched through launch_pdl: the previous kernel of the stream has completed from here on
    pdl_release();  // the next one may be scheduled as SMs free up

    // A. packed pixels of [X0-r-1, X0+kTW+r+1) x [Y0-r-1, Y0+kTH+r+1), coordinates clamped to the image
    //    interior tiles: 4 pixels per thread from aligned 32-bit loads; border tiles / leftover columns pixel by pixel
    const bool interior = X0 - r - 1 >= 0 && X0 - r - 1 + L::PW < W;
    constexpr int kGroups = L::PW / 4;
    const int vec_cols = interior ? kGroups * 4 : 0;
    if (interior) {
        for (int item = tid; item < L::PH * kGroups; item += kThreads) {
            const int ey = item / kGroups, g = item - ey * kGroups;
            const int gy = rowmap(Y0 - r - 1 + ey);
            std::uint32_t v[4];
            load_packed4(image + (static_cast<size_t>(gy) * W + (X0 - r - 1 + 4 * g)) * 3, v);
#pragma unroll
            for (int j = 0; j < 4; ++j) pix[ey * L::PW + 4 * g + j] = v[j];
        }
    }
    {
        const int rem = L::PW - vec_cols;
        for (int item = tid; item < L::PH * rem; item += kThreads) {
            const int ey = item / rem, ex = vec_cols + (item - ey * rem);
            const int gx = clampi(X0 - r - 1 + ex, 0, W - 1);
            const int gy = rowmap(Y0 - r - 1 + ey);
            pix[ey * L::PW + ex] = load_packed(image + (static_cast<size_t>(gy) * W + gx) * 3);
        }
    }
    if constexpr (!FUSED) {
        for (int e = tid; e < L::MW * L::MH; e += kThreads) {
            const int my = e / L::MW, mx = e - my * L::MW;
            const int gx = clampi(X0 - r + mx, 0, W - 1);
            const int gy = rowmap(Y0 - r + my);
            mag[my * L::MP + mx] = __ldg(magnitude + static_cast<size_t>(gy) * W + gx);
        }
    }
    __syncthreads();

    // B. gradient magnitude at the CLAMPED coordinate of every tile position (src/gradient_impl.cu:52-65):
    //    central differences with clamped neighbours, integer squares, sqrtf
    if constexpr (FUSED) {
        for (int e = tid; e < L::MW * L::MH; e += kThreads) {
            const int my = e / L::MW, mx = e - my * L::MW;
            const int cx = clampi(X0 - r + mx, 0, W - 1), cy = clampi(Y0 - r + my, 0, H - 1);
            const int tx = cx - (X0 - r - 1), ty = cy - (Y0 - r - 1);
            const std::uint32_t* c = pix + ty * L::PW + tx;
            const int xm = cx > 0 ? -1 : 0, xp = cx < W - 1 ? 1 : 0;
            const int ym = cy > 0 ? -L::PW : 0, yp = cy < H - 1 ? L::PW : 0;
            const int acc = sq_diff(c[xp], c[xm]) + sq_diff(c[yp], c[ym]);
            mag[my * L::MP + mx] = sqrtf(static_cast<float>(acc));
        }
        __syncthreads();
    }

    // C. horizontal pass: record = {B | R<<16, G, min(b+g+r) | max(b+g+r)<<16, max magnitude} over K columns.
    //    A thread produces the records of 8 adjacent columns of one row from K+7 pixels / magnitudes held in
    //    registers: the two sums slide (add the entering, subtract the leaving pixel), min / max / fmax of a K-window
    //    are two overlapping power-of-two windows built by doubling (all exact, order-independent operations).
    {
        constexpr int N = K + 7;                       // inputs of 8 outputs
        constexpr int Q = K >= 8 ? 8 : (K >= 4 ? 4 : (K >= 2 ? 2 : 1));   // largest power of two <= K
        for (int item = tid; item < L::MH * (kTW / 8); item += kThreads) {
            const int cg = item & 7, row = item >> 3, col0 = cg * 8;   // kTW / 8 == 8 column groups
            const std::uint32_t* pr = pix + (row + 1) * L::PW + col0 + 1;
            const float* mr = mag + row * L::MP + col0;
            std::uint32_t br[8], g[8], lh[8];
            float fm[8];
            {
                std::uint32_t p[N];
#pragma unroll
                for (int j = 0; j < N; ++j) p[j] = pr[j];
                std::uint32_t sb = 0, sg = 0;
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    sb += __byte_perm(p[j], 0u, 0x4240);
                    sg += __byte_perm(p[j], 0u, 0x4441);
                }
                br[0] = sb;
                g[0] = sg;
#pragma unroll
                for (int i = 1; i < 8; ++i) {
                    sb += __byte_perm(p[i + K - 1], 0u, 0x4240) - __byte_perm(p[i - 1], 0u, 0x4240);
                    sg += __byte_perm(p[i + K - 1], 0u, 0x4441) - __byte_perm(p[i - 1], 0u, 0x4441);
                    br[i] = sb;
                    g[i] = sg;
                }
                {
                    std::uint32_t a[N];
#pragma unroll
                    for (int j = 0; j < N; ++j) a[j] = __vsadu4(p[j], 0u);
#pragma unroll
                    for (int d = 1; d < Q; d <<= 1) {
#pragma unroll
                        for (int j = 0; j + d < N; ++j) a[j] = min(a[j], a[j + d]);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) lh[i] = min(a[i], a[i + K - Q]);
                }
                {
                    std::uint32_t c[N];
#pragma unroll
                    for (int j = 0; j < N; ++j) c[j] = __vsadu4(p[j], 0u);
#pragma unroll
                    for (int d = 1; d < Q; d <<= 1) {
#pragma unroll
                        for (int j = 0; j + d < N; ++j) c[j] = max(c[j], c[j + d]);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) lh[i] |= max(c[i], c[i + K - Q]) << 16;
                }
            }
            {
                float f[N];
#pragma unroll
                for (int j = 0; j < N; ++j) f[j] = mr[j];
#pragma unroll
                for (int d = 1; d < Q; d <<= 1) {
#pragma unroll
                    for (int j = 0; j + d < N; ++j) f[j] = fmaxf(f[j], f[j + d]);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) fm[i] = fmaxf(fmaxf(f[i], f[i + K - Q]), 0.f);   // the reference's max starts at 0
            }
            uint4* hrow = hs + row * kTW + col0;
#pragma unroll
            for (int i = 0; i < 8; ++i)   // rotated inside the group of 8: the 8 lanes of a quarter warp hit 8 different 16-byte slots
                hrow[(i + cg) & 7] = make_uint4(br[i], g[i], lh[i], __float_as_uint(fm[i]));
        }
    }
    __syncthreads();

    // D. vertical pass + raster-order magnitude sum; lane <-> column, kRows consecutive rows per thread
    const int col = (warp & 1) * 32 + lane;
    const int r0 = (warp >> 1) * kRows;
    std::uint32_t sbr[kRows], sg[kRows], mn[kRows], mx[kRows];
    float mm[kRows];
    // raster-order magnitude sums of the outputs, two outputs (rows 2p, 2p+1) per packed fp32x2 register: while a tile row
    // serves both, their chains advance with ONE FFMA2 per magnitude, acc = 1 * {m, m} + acc -- each half rounds exactly
    // like the scalar add (fma(1, m, s) == s + m), half the issue slots of the K*K adds per output
    unsigned long long msum2[kRows / 2];
#pragma unroll
    for (int i = 0; i < kRows; ++i) {
        sbr[i] = sg[i] = 0;
        mn[i] = 0xffffu;
        mx[i] = 0;
        mm[i] = 0.f;
    }
#pragma unroll
    for (int p = 0; p < kRows / 2; ++p) msum2[p] = 0ull;
    unsigned long long one2;
    asm("mov.b64 %0, {%1, %1};" : "=l"(one2) : "f"(1.f));
#pragma unroll
    for (int R = 0; R < K + kRows - 1; ++R) {
        const uint4 h = hs[(r0 + R) * kTW + (col & ~7) + ((col + (col >> 3)) & 7)];
        float v[K];
#pragma unroll
        for (int j = 0; j < K; ++j) v[j] = mag[(r0 + R) * L::MP + col + j];
#pragma unroll
        for (int i = 0; i < kRows; ++i) {
            if (R - i >= 0 && R - i < K) {
                sbr[i] += h.x;
                sg[i] += h.y;
                mn[i] = min(mn[i], h.z & 0xffffu);
                mx[i] = max(mx[i], h.z >> 16);
                mm[i] = fmaxf(mm[i], __uint_as_float(h.w));
            }
        }
#pragma unroll
        for (int p = 0; p < kRows / 2; ++p) {
            const bool lo = R - 2 * p >= 0 && R - 2 * p < K, hi = R - 2 * p - 1 >= 0 && R - 2 * p - 1 < K;
            if (lo && hi) {
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    unsigned long long vv;
                    asm("mov.b64 %0, {%1, %1};" : "=l"(vv) : "f"(v[j]));
                    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(msum2[p]) : "l"(one2), "l"(vv));
                }
            } else if (lo || hi) {
                float a, b;
                asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(msum2[p]));
#pragma unroll
                for (int j = 0; j < K; ++j) {
                    if (lo) a = __fadd_rn(a, v[j]);
                    else b = __fadd_rn(b, v[j]);
                }
                asm("mov.b64 %0, {%1, %2};" : "=l"(msum2[p]) : "f"(a), "f"(b));
            }
        }
    }
    float msum[kRows];
#pragma unroll
    for (int p = 0; p < kRows / 2; ++p) asm("mov.b64 {%0, %1}, %2;" : "=f"(msum[2 * p]), "=f"(msum[2 * p + 1]) : "l"(msum2[p]));

    // E. epilogue (:88-103) and coalesced stores
    const int x = X0 + col;
    if (x >= W) return;
    constexpr float kk = static_cast<float>(K * K), rkk = 1.f / kk, r3 = 1.f / 3.f;
#pragma unroll
    for (int i = 0; i < kRows; ++i) {
        const int y = Y0 + r0 + i;
        if (y >= out_y0 + out_rows) break;
        const size_t o = static_cast<size_t>(y - dst_y0) * W + x;
        blurred[o * 3 + 0] = div_const(static_cast<float>(sbr[i] & 0xffffu), kk, rkk);
        blurred[o * 3 + 1] = div_const(static_cast<float>(sg[i]), kk, rkk);
        blurred[o * 3 + 2] = div_const(static_cast<float>(sbr[i] >> 16), kk, rkk);
        const float imax = div_const(static_cast<float>(mx[i]), 3.f, r3);
        const float imin = div_const(static_cast<float>(mn[i]), 3.f, r3);
        const float num = __fmul_rn(__fsub_rn(imax, imin), mm[i]);
        rtv[o] = static_cast<float>(static_cast<double>(num) / (static_cast<double>(msum[i]) + 1e-9));
    }
}

// Any window size: one thread per pixel straight from global memory, same arithmetic.
__global__ void blur_rtv_generic_kernel(const std::uint8_t* __restrict__ image, const float* __restrict__ magnitude,
                                        float* __restrict__ blurred, float* __restrict__ rtv, const int W, const int H,
                                        const int k, const int src_y0, const int src_rows, const int dst_y0,
                                        const int out_y0, const int out_rows) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = out_y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= out_y0 + out_rows) return;
    const RowMap rowmap{H, src_y0, src_rows};
    const int r = k / 2;
    int s0 = 0, s1 = 0, s2 = 0, mn = 0xffff, mx = 0;
    float mm = 0.f, msum = 0.f;
    for (int ky = -r; ky <= r; ++ky) {
        const int cy = clampi(y + ky, 0, H - 1);
        const int yy = rowmap(cy);
        for (int kx = -r; kx <= r; ++kx) {
            const int cx = clampi(x + kx, 0, W - 1);
            const std::uint8_t* p = image + (static_cast<size_t>(yy) * W + cx) * 3;
            const int b = p[0], g = p[1], rr = p[2];
            s0 += b;
            s1 += g;
            s2 += rr;
            mn = min(mn, b + g + rr);
            mx = max(mx, b + g + rr);
            float m;
            if (magnitude) {
                m = __ldg(magnitude + static_cast<size_t>(yy) * W + cx);
            } else {
                const int xm = max(cx - 1, 0), xp = min(cx + 1, W - 1);
                const std::uint8_t* pu = image + (static_cast<size_t>(rowmap(cy - 1)) * W + cx) * 3;
                const std::uint8_t* pd = image + (static_cast<size_t>(rowmap(cy + 1)) * W + cx) * 3;
                const std::uint8_t* pl = image + (static_cast<size_t>(yy) * W + xm) * 3;
                const std::uint8_t* pr = image + (static_cast<size_t>(yy) * W + xp) * 3;
                int acc = 0;
                for (int c = 0; c < 3; ++c) {
                    const int dx = static_cast<int>(pr[c]) - static_cast<int>(pl[c]);
                    const int dy = static_cast<int>(pd[c]) - static_cast<int>(pu[c]);
                    acc += dx * dx + dy * dy;
                }
                m = sqrtf(static_cast<float>(acc));
            }
            mm = fmaxf(mm, m);
            msum = __fadd_rn(msum, m);
        }
    }
    const float kk = static_cast<float>(k * k);
    const size_t o = static_cast<size_t>(y - dst_y0) * W + x;
    blurred[o * 3 + 0] = __fdiv_rn(static_cast<float>(s0), kk);
    blurred[o * 3 + 1] = __fdiv_rn(static_cast<float>(s1), kk);
    blurred[o * 3 + 2] = __fdiv_rn(static_cast<float>(s2), kk);
    const float imax = __fdiv_rn(static_cast<float>(mx), 3.f);
    const float imin = __fdiv_rn(static_cast<float>(mn), 3.f);
    const float num = __fmul_rn(__fsub_rn(imax, imin), mm);
    rtv[o] = static_cast<float>(static_cast<double>(num) / (static_cast<double>(msum) + 1e-9));
}

// 4-byte asynchronous global -> shared copy (LDGSTS): no register staging, any number in flight
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<std::uint32_t>(__cvta_generic_to_shared(smem_dst))),
                 "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

template <int K>
struct GuideLayout {
    static constexpr int r = K / 2;
    static constexpr int MW = kTW + 2 * r, MH = kTH + 2 * r;
    static constexpr int TP = MW + 2;   // rtv tile pitch: consecutive rows 10 banks apart
    static constexpr int hs_off = (TP * MH * 4 + 15) & ~15;
    static constexpr int out_off = hs_off + MH * kTW * 8;
    static constexpr int bl_off = (out_off + kTH * kTW * 3 + 15) & ~15;   // blurred tile + halo: MH x MW x 3 floats
    static constexpr int total = bl_off + MW * MH * 12;
};

// alpha and the blend of one guide pixel (:167-176); bq / bp are the 3 blurred floats of the argmin / centre pixel
__device__ __forceinline__ void guide_pixel(const float rtv_p, const float rtv_min, const float sigma_alpha,
                                            const float bq[3], const float bp[3], std::uint8_t* out) {
    const float e = expf(__fmul_rn(sigma_alpha, __fsub_rn(rtv_p, rtv_min)));
    const float alpha = __fsub_rn(__fdiv_rn(2.f, __fadd_rn(1.f, e)), 1.f);
    const float one_m = __fsub_rn(1.f, alpha);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float t = __fadd_rn(__fmaf_rn(alpha, bq[c], __fmul_rn(one_m, bp[c])), 0.5f);
        out[c] = static_cast<std::uint8_t>(clampi(__float2int_rz(t), 0, 255));
    }
}

template <int K>
__global__ void __launch_bounds__(kThreads, 3) guide_kernel(const float* __restrict__ blurred,
                                                           const float* __restrict__ rtv, std::uint8_t* __restrict__ guide,
                                                           const int W, const int H, const int src_y0, const int src_rows,
                                                           const int dst_y0, const int out_y0, const int out_rows) {
    using L = GuideLayout<K>;
    constexpr int r = L::r;
    extern __shared__ __align__(16) unsigned char smem[];
    float* const tile = reinterpret_cast<float*>(smem);
    float2* const hs = reinterpret_cast<float2*>(smem + L::hs_off);
    std::uint8_t* const obuf = smem + L::out_off;
    float* const bl = reinterpret_cast<float*>(smem + L::bl_off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int X0 = blockIdx.x * kTW;
    const int Y0 = out_y0 + blockIdx.y * kTH;
    const RowMap rowmap{H, src_y0, src_rows};
    pdl_wait();     // launched through launch_pdl: the previous kernel of the stream has completed from here on
    pdl_release();  // the next one may be scheduled as SMs free up

    // The rtv tile and the blurred tile (+ radius halo) go to shared memory: the argmin pixel of every output lies inside
    // the tile, so the data-dependent gather of the reference (:168-175) becomes a shared-memory read.  Interior tiles
    // are copied with cp.async (every copy of the tile in flight at once: the kernel waits for global memory once per
    // tile, not once per output row as the gather version did); border tiles clamp pixel by pixel.
    constexpr int kRowF = L::MW * 3;
    static_assert(kRowF <= kThreads, "one thread per float of a blurred tile row");
    const bool interior = X0 - r >= 0 && X0 - r + L::MW <= W;
    if (interior) {
        const int wrow = tid >> 5, lane_ = tid & 31;
        for (int ey = wrow; ey < L::MH; ey += kThreads / 32) {       // rtv: a warp per row
            const float* g = rtv + static_cast<size_t>(rowmap(Y0 - r + ey)) * W + (X0 - r);
            for (int ex = lane_; ex < L::MW; ex += 32) cp_async4(tile + ey * L::TP + ex, g + ex);
        }
#pragma unroll 4
        for (int ey = 0; ey < L::MH; ++ey) {                          // blurred: the CTA per row
            const float* g = blurred + (static_cast<size_t>(rowmap(Y0 - r + ey)) * W + (X0 - r)) * 3;
            if (tid < kRowF) cp_async4(bl + ey * kRowF + tid, g + tid);
        }
        cp_async_wait_all();
    } else {
        for (int e = tid; e < L::MW * L::MH; e += kThreads) {
            const int ey = e / L::MW, ex = e - ey * L::MW;
            const int gx = clampi(X0 - r + ex, 0, W - 1);
            tile[ey * L::TP + ex] = __ldg(rtv + static_cast<size_t>(rowmap(Y0 - r + ey)) * W + gx);
        }
        for (int e = tid; e < L::MH * kRowF; e += kThreads) {
            const int ey = e / kRowF, t = e - ey * kRowF;
            const int ex = t / 3, c = t - ex * 3;
            bl[e] = __ldg(blurred + (static_cast<size_t>(rowmap(Y0 - r + ey)) * W + clampi(X0 - r + ex, 0, W - 1)) * 3 + c);
        }
    }
    __syncthreads();

    // Both passes of the separable argmin use the same scheme: a thread holds the K+7 inputs of 8 adjacent outputs and
    // builds "first minimum of a power-of-two block" by doubling -- the later half replaces the earlier one only when it is
    // STRICTLY smaller, which is the reference's strict '>' scan (:152-165) made associative -- and a K-window is two
    // overlapping blocks.  A window without any value below 1e10f keeps (1e10f, -1) like the reference's initial state.
    constexpr int N = K + 7;
    constexpr int Q = K >= 8 ? 8 : (K >= 4 ? 4 : (K >= 2 ? 2 : 1));
    // horizontal pass: a thread = 8 adjacent columns of one row -> (min value, smallest kx attaining it)
    for (int item = tid; item < L::MH * (kTW / 8); item += kThreads) {
        const int cg = item & 7, row = item >> 3;
        const float* tr = tile + row * L::TP + cg * 8;
        float v[N];
        int id[N];
#pragma unroll
        for (int j = 0; j < N; ++j) {
            v[j] = tr[j];
            id[j] = j;
        }
#pragma unroll
        for (int d = 1; d < Q; d <<= 1) {
#pragma unroll
            for (int j = 0; j + d < N; ++j) {
                const bool take = v[j + d] < v[j];
                v[j] = take ? v[j + d] : v[j];
                id[j] = take ? id[j + d] : id[j];
            }
        }
        float2* hrow = hs + row * kTW + cg * 8;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const bool take = v[i + K - Q] < v[i];
            float best = take ? v[i + K - Q] : v[i];
            int bj = (take ? id[i + K - Q] : id[i]) - i;
            if (!(1e10f > best)) {  // :152
                best = 1e10f;
                bj = -1;
            }
            hrow[(i + cg) & 7] = make_float2(best, __int_as_float(bj));   // rotated: conflict-free 8-byte stores
        }
    }
    __syncthreads();

    // vertical pass: lane <-> column, 8 consecutive rows per thread -> smallest ky among equal row minima
    const int col = (warp & 1) * 32 + lane;
    const int r0 = (warp >> 1) * kRows;
    const float sigma_alpha = __fdiv_rn(1.f, static_cast<float>(5 * K));  // :121
    float best[kRows];
    int bj[kRows], bky[kRows];
    {
        static_assert(kRows == 8, "8 outputs per thread");
        const int hcol = (col & ~7) + ((col + (col >> 3)) & 7);
        float v[N];
        int xj[N], yk[N];
#pragma unroll
        for (int R = 0; R < N; ++R) {
            const float2 h = hs[(r0 + R) * kTW + hcol];
            v[R] = h.x;
            xj[R] = __float_as_int(h.y);
            yk[R] = R;
        }
#pragma unroll
        for (int d = 1; d < Q; d <<= 1) {
#pragma unroll
            for (int R = 0; R + d < N; ++R) {
                const bool take = v[R + d] < v[R];
                v[R] = take ? v[R + d] : v[R];
                xj[R] = take ? xj[R + d] : xj[R];
                yk[R] = take ? yk[R + d] : yk[R];
            }
        }
#pragma unroll
        for (int i = 0; i < kRows; ++i) {
            const bool take = v[i + K - Q] < v[i];
            best[i] = take ? v[i + K - Q] : v[i];
            bj[i] = take ? xj[i + K - Q] : xj[i];
            bky[i] = (take ? yk[i + K - Q] : yk[i]) - i;
        }
    }
#pragma unroll
    for (int i = 0; i < kRows; ++i) {
        const int y = min(Y0 + r0 + i, H - 1);
        const float rp = tile[(r0 + i + r) * L::TP + col + r];
        const float* bpp = bl + ((r0 + i + r) * L::MW + col + r) * 3;
        const float bp[3] = {bpp[0], bpp[1], bpp[2]};
        float rmin = best[i];
        float bq[3];
        if (bj[i] >= 0) {
            // tile element (col + kx, row + ky) holds the CLAMPED pixel (:159-163), exactly what the reference gathers
            const float* bqp = bl + ((r0 + i + bky[i]) * L::MW + col + bj[i]) * 3;
            bq[0] = bqp[0], bq[1] = bqp[1], bq[2] = bqp[2];
        } else {
            // nothing below 1e10f: the reference keeps its initial coordinates (0, 0)
            rmin = __ldg(rtv + static_cast<size_t>(rowmap(0)) * W);
            const float* g0 = blurred + static_cast<size_t>(rowmap(0)) * W * 3;
            bq[0] = __ldg(g0), bq[1] = __ldg(g0 + 1), bq[2] = __ldg(g0 + 2);
        }
        guide_pixel(rp, rmin, sigma_alpha, bq, bp, obuf + ((r0 + i) * kTW + col) * 3);
    }
    __syncthreads();

    // store the tile: rows of up to 192 bytes, as 32-bit words when the row start is aligned
    const int cols = min(kTW, W - X0);
    const int rows = min(kTH, out_y0 + out_rows - Y0);
    const int row_bytes = cols * 3;
    for (int row = warp; row < rows; row += kThreads / 32) {
        std::uint8_t* g = guide + (static_cast<size_t>(Y0 + row - dst_y0) * W + X0) * 3;
        const std::uint8_t* s = obuf + row * kTW * 3;
        if ((reinterpret_cast<std::uintptr_t>(g) & 3u) == 0) {
            const int words = row_bytes >> 2;
            for (int i = lane; i < words; i += 32)
                reinterpret_cast<std::uint32_t*>(g)[i] = reinterpret_cast<const std::uint32_t*>(s)[i];
            for (int i = (words << 2) + lane; i < row_bytes; i += 32) g[i] = s[i];
        } else {
            for (int i = lane; i < row_bytes; i += 32) g[i] = s[i];
        }
    }
}

__global__ void guide_generic_kernel(const float* __restrict__ blurred, const float* __restrict__ rtv,
                                     std::uint8_t* __restrict__ guide, const int W, const int H, const int k,
                                     const int src_y0, const int src_rows, const int dst_y0, const int out_y0,
                                     const int out_rows) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = out_y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= out_y0 + out_rows) return;
    const RowMap rowmap{H, src_y0, src_rows};
    const int r = k / 2;
    float rmin = 1e10f;
    int mx = 0, my = 0;
    for (int ky = -r; ky <= r; ++ky) {
        const int cy = clampi(y + ky, 0, H - 1);
        for (int kx = -r; kx <= r; ++kx) {
            const int cx = clampi(x + kx, 0, W - 1);
            const float v = __ldg(rtv + static_cast<size_t>(rowmap(cy)) * W + cx);
            if (rmin > v) {
                rmin = v;
                mx = cx;
                my = cy;
            }
        }
    }
    const float rq = __ldg(rtv + static_cast<size_t>(rowmap(my)) * W + mx);
    const float rp = __ldg(rtv + static_cast<size_t>(rowmap(y)) * W + x);
    std::uint8_t out[3];
    const float* gq = blurred + (static_cast<size_t>(rowmap(my)) * W + mx) * 3;
    const float* gp = blurred + (static_cast<size_t>(rowmap(y)) * W + x) * 3;
    const float bq[3] = {__ldg(gq), __ldg(gq + 1), __ldg(gq + 2)};
    const float bp[3] = {__ldg(gp), __ldg(gp + 1), __ldg(gp + 2)};
    guide_pixel(rp, rq, __fdiv_rn(1.f, static_cast<float>(5 * k)), bq, bp, out);
    std::uint8_t* g = guide + (static_cast<size_t>(y - dst_y0) * W + x) * 3;
    g[0] = out[0];
    g[1] = out[1];
    g[2] = out[2];
}

template <int K>
cudaError_t launch_blur_rtv_k(const std::uint8_t* image, const float* magnitude, float* blurred, float* rtv, int W, int H,
                              int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream) {
    using L = BlurLayout<K>;
    const dim3 grid((W + kTW - 1) / kTW, (out_rows + kTH - 1) / kTH);
    cudaError_t e;
    if (magnitude) {
        e = cudaFuncSetAttribute(blur_rtv_kernel<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::total);
        if (e != cudaSuccess) return e;
        e = launch_pdl(blur_rtv_kernel<K, false>, grid, dim3(kThreads), L::total, stream, image, magnitude, blurred, rtv, W, H,
                       src_y0, src_rows, dst_y0, out_y0, out_rows);
    } else {
        e = cudaFuncSetAttribute(blur_rtv_kernel<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::total);
        if (e != cudaSuccess) return e;
        e = launch_pdl(blur_rtv_kernel<K, true>, grid, dim3(kThreads), L::total, stream, image,
                       static_cast<const float*>(nullptr), blurred, rtv, W, H, src_y0, src_rows, dst_y0, out_y0, out_rows);
    }
    return e;
}

template <int K>
cudaError_t launch_guide_k(const float* blurred, const float* rtv, std::uint8_t* guide, int W, int H, int src_y0,
                           int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream) {
    using L = GuideLayout<K>;
    const dim3 grid((W + kTW - 1) / kTW, (out_rows + kTH - 1) / kTH);
    cudaError_t e = cudaFuncSetAttribute(guide_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::total);
    if (e != cudaSuccess) return e;
    return launch_pdl(guide_kernel<K>, grid, dim3(kThreads), L::total, stream, blurred, rtv, guide, W, H, src_y0, src_rows,
                      dst_y0, out_y0, out_rows);
}

}  // namespace

// magnitude == nullptr: the gradient magnitude of `image` is computed on the fly (fused gradient stage).
cudaError_t launch_blur_rtv(const std::uint8_t* image, const float* magnitude, float* blurred, float* rtv, int width,
                            int height, int ksize, int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows,
                            cudaStream_t stream) {
    if (out_rows <= 0 || width <= 0) return cudaSuccess;
    count_launch();
#define VIP_BLUR_CASE(K)                                                                                         \
    case K:                                                                                                      \
        return launch_blur_rtv_k<K>(image, magnitude, blurred, rtv, width, height, src_y0, src_rows, dst_y0, out_y0, \
                                    out_rows, stream);
    switch (ksize) {
        VIP_BLUR_CASE(3)
        VIP_BLUR_CASE(5)
        VIP_BLUR_CASE(7)
        VIP_BLUR_CASE(9)
        VIP_BLUR_CASE(11)
        VIP_BLUR_CASE(13)
        VIP_BLUR_CASE(15)
        default:
            break;
    }
#undef VIP_BLUR_CASE
    const dim3 block(32, 8);
    const dim3 grid((width + 31) / 32, (out_rows + 7) / 8);
    blur_rtv_generic_kernel<<<grid, block, 0, stream>>>(image, magnitude, blurred, rtv, width, height, ksize, src_y0,
                                                       src_rows, dst_y0, out_y0, out_rows);
    return cudaGetLastError();
}

cudaError_t launch_guide(const float* blurred, const float* rtv, std::uint8_t* guide, int width, int height, int ksize,
                         int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream) {
    if (out_rows <= 0 || width <= 0) return cudaSuccess;
    count_launch();
#define VIP_GUIDE_CASE(K)                                                                                            \
    case K:                                                                                                          \
        return launch_guide_k<K>(blurred, rtv, guide, width, height, src_y0, src_rows, dst_y0, out_y0, out_rows, stream);
    switch (ksize) {
        VIP_GUIDE_CASE(3)
        VIP_GUIDE_CASE(5)
        VIP_GUIDE_CASE(7)
        VIP_GUIDE_CASE(9)
        VIP_GUIDE_CASE(11)
        VIP_GUIDE_CASE(13)
        VIP_GUIDE_CASE(15)
        default:
            break;
    }
#undef VIP_GUIDE_CASE
    const dim3 block(32, 8);
    const dim3 grid((width + 31) / 32, (out_rows + 7) / 8);
    guide_generic_kernel<<<grid, block, 0, stream>>>(blurred, rtv, guide, width, height, ksize, src_y0, src_rows, dst_y0,
                                                    out_y0, out_rows);
    return cudaGetLastError();
}

}  // namespace vip
