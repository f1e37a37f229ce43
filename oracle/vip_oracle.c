/*
 * TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or called from the product library.
 *
 * Plain-C CPU restatement of the reference's edge-preserving-filter hot path, written from the
 * reference's behaviour (no code copied).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against
 *   (1) the reference's own CPU headers compiled unmodified (oracle/_ref/libvip_ref_cpu.so, built by
 *       oracle/Makefile from /root/reference/include/cpp in place), bit-for-bit in "cpu" arithmetic mode,
 *   (2) the reference's known-answer test: mt19937(42) 50x50 fixture vs cv2.bilateralFilter
 *       (/root/reference/test/bilateral_filter.cu:35-63), committed as tests/golden/*.npz,
 *   (3) on the GPU box, the reference's own CUDA kernels (oracle/_ref/libvip_ref_cuda.so) in "cuda" mode.
 *
 * Arithmetic modes (the reference's CPU and CUDA paths are not bit-identical to each other):
 *   VIPO_ARITH_CUDA (0): float LUT coefficients + expf, fused multiply-add in the weighted sums,
 *                        double epsilon / fp64 divide in the RTV           (follows /root/reference/src/ *.cu)
 *   VIPO_ARITH_CPU  (1): double LUT coefficients + exp, separate multiply and add, float epsilon
 *                                                                          (follows /root/reference/include/cpp)
 * All images are packed interleaved, pitch = width*channels, borders replicate (clamp).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define VIPO_ARITH_CUDA 0
#define VIPO_ARITH_CPU 1

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* LUTs.  CUDA: src/bilateral_filter_impl.cu:204-239 (float coeff, float exp);
 *        CPU : include/cpp/bilateral_filter.hpp:12-37 (double coeff, double exp, rounded to float).
 * kernel_space is ksize*ksize with a circular mask (0 where kx^2+ky^2 > r^2). */
void vipo_build_luts(int ksize, float sigma_space, float sigma_color, int table_len, int arith, float* kernel_space,
                     float* color_table) {
    const int radius = ksize / 2;
    memset(kernel_space, 0, sizeof(float) * (size_t)ksize * ksize);
    if (arith == VIPO_ARITH_CUDA) {
        const float cc = -1.f / (2 * sigma_color * sigma_color);
        const float cs = -1.f / (2 * sigma_space * sigma_space);
        for (int ky = -radius; ky <= radius; ky++)
            for (int kx = -radius; kx <= radius; kx++) {
                const int r2 = kx * kx + ky * ky;
                if (r2 > radius * radius) continue;
                kernel_space[(ky + radius) * ksize + (kx + radius)] = expf(r2 * cs);
            }
        for (int i = 0; i < table_len; i++) color_table[i] = expf((i * i) * cc);
    } else {
        const double cc = -1. / (2 * sigma_color * sigma_color);
        const double cs = -1. / (2 * sigma_space * sigma_space);
        for (int ky = -radius; ky <= radius; ky++)
            for (int kx = -radius; kx <= radius; kx++) {
                const int r2 = kx * kx + ky * ky;
                if (r2 > radius * radius) continue;
                kernel_space[(ky + radius) * ksize + (kx + radius)] = (float)exp(r2 * cs);
            }
        for (int i = 0; i < table_len; i++) color_table[i] = (float)exp((i * i) * cc);
    }
}

static inline float mad(float a, float b, float c, int arith) {
    if (arith == VIPO_ARITH_CUDA) return fmaf(a, b, c);
    volatile float p = a * b; /* keep the product rounded to fp32 */
    return p + c;
}

/* Bilateral filter: src/bilateral_filter_impl.cu:7-96 ; include/cpp/bilateral_filter.hpp:41-124.
 * Rows [y_begin, y_begin+rows) of the output are produced; src is the FULL image. */
void vipo_joint_bilateral_filter_rows(const uint8_t* src, const uint8_t* guide, uint8_t* dst, int w, int h, int ksize,
                                      float sigma_space, float sigma_color, int arith, int y_begin, int rows) {
    const int radius = ksize / 2;
    float* ks = (float*)malloc(sizeof(float) * (size_t)ksize * ksize);
    float lut[768];
    vipo_build_luts(ksize, sigma_space, sigma_color, 768, arith, ks, lut);
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = y_begin; y < y_begin + rows; y++) {
        for (int x = 0; x < w; x++) {
            const uint8_t* c = guide + ((size_t)y * w + x) * 3;
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, sk = 0.f;
            for (int ky = -radius; ky <= radius; ky++) {
                const int yy = clampi(y + ky, 0, h - 1);
                for (int kx = -radius; kx <= radius; kx++) {
                    const int xx = clampi(x + kx, 0, w - 1);
                    const uint8_t* g = guide + ((size_t)yy * w + xx) * 3;
                    const uint8_t* p = src + ((size_t)yy * w + xx) * 3;
                    const int dist = abs((int)c[0] - (int)g[0]) + abs((int)c[1] - (int)g[1]) + abs((int)c[2] - (int)g[2]);
                    const float k = ks[(ky + radius) * ksize + (kx + radius)] * lut[dist];
                    s0 = mad((float)p[0], k, s0, arith);
                    s1 = mad((float)p[1], k, s1, arith);
                    s2 = mad((float)p[2], k, s2, arith);
                    sk += k;
                }
            }
            uint8_t* o = dst + ((size_t)y * w + x) * 3;
            o[0] = (uint8_t)(s0 / sk + 0.5f);
            o[1] = (uint8_t)(s1 / sk + 0.5f);
            o[2] = (uint8_t)(s2 / sk + 0.5f);
        }
    }
    free(ks);
}

void vipo_bilateral_filter(const uint8_t* src, uint8_t* dst, int w, int h, int ksize, float sigma_space,
                           float sigma_color, int arith) {
    vipo_joint_bilateral_filter_rows(src, src, dst, w, h, ksize, sigma_space, sigma_color, arith, 0, h);
}

/* src/bilateral_filter_impl.cu:98-202 ; include/cpp/bilateral_filter.hpp:126-207 */
void vipo_joint_bilateral_filter(const uint8_t* src, const uint8_t* guide, uint8_t* dst, int w, int h, int ksize,
                                 float sigma_space, float sigma_color, int arith) {
    vipo_joint_bilateral_filter_rows(src, guide, dst, w, h, ksize, sigma_space, sigma_color, arith, 0, h);
}

/* Border-replicated integral image: include/cpp/border_replicated_integral_image.hpp:6-85.
 * table is (h+2r+1) x (w+2r+1) x 3, 32-bit; unsigned arithmetic (wrap-around is well defined and gives
 * the right box sums for any size). */
void vipo_integral_u8c3(const uint8_t* src, int w, int h, int radius, uint32_t* table) {
    const int tw = w + 2 * radius + 1, th = h + 2 * radius + 1;
    memset(table, 0, sizeof(uint32_t) * 3 * (size_t)tw * th);
    for (int Y = 1; Y < th; Y++) {
        const int sy = clampi(Y - 1 - radius, 0, h - 1);
        uint32_t run[3] = {0, 0, 0};
        for (int X = 1; X < tw; X++) {
            const int sx = clampi(X - 1 - radius, 0, w - 1);
            const uint8_t* p = src + ((size_t)sy * w + sx) * 3;
            uint32_t* t = table + ((size_t)Y * tw + X) * 3;
            const uint32_t* up = table + ((size_t)(Y - 1) * tw + X) * 3;
            for (int c = 0; c < 3; c++) {
                run[c] += p[c];
                t[c] = up[c] + run[c];
            }
        }
    }
}

/* get(x0,y0,x1,y1): inclusive clamped box sum, border_replicated_integral_image.hpp:34-47 */
void vipo_integral_get(const uint32_t* table, int w, int radius, int x0, int y0, int x1, int y1, uint32_t* out3) {
    const int tw = w + 2 * radius + 1;
    const uint32_t* a = table + ((size_t)(y1 + radius + 1) * tw + (x1 + radius + 1)) * 3;
    const uint32_t* b = table + ((size_t)(y1 + radius + 1) * tw + (x0 + radius)) * 3;
    const uint32_t* c = table + ((size_t)(y0 + radius) * tw + (x1 + radius + 1)) * 3;
    const uint32_t* d = table + ((size_t)(y0 + radius) * tw + (x0 + radius)) * 3;
    for (int ch = 0; ch < 3; ch++) out3[ch] = a[ch] - b[ch] - c[ch] + d[ch];
}

/* Adaptive bilateral filter: src/adaptive_bilateral_filter_impl.cu:7-115 ;
 * include/cpp/adaptive_bilateral_filter.hpp:13-104.  The box sum is an exact integer either way. */
void vipo_adaptive_bilateral_filter_rows(const uint8_t* src, uint8_t* dst, int w, int h, int ksize, float sigma_space,
                                         float sigma_color, int arith, int y_begin, int rows) {
    const int radius = ksize / 2;
    float* ks = (float*)malloc(sizeof(float) * (size_t)ksize * ksize);
    float* lut = (float*)malloc(sizeof(float) * 1536);
    vipo_build_luts(ksize, sigma_space, sigma_color, 1536, arith, ks, lut);
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = y_begin; y < y_begin + rows; y++) {
        for (int x = 0; x < w; x++) {
            const uint8_t* c = src + ((size_t)y * w + x) * 3;
            int b0 = 0, b1 = 0, b2 = 0;
            for (int ky = -radius; ky <= radius; ky++) {
                const int yy = clampi(y + ky, 0, h - 1);
                for (int kx = -radius; kx <= radius; kx++) {
                    const uint8_t* p = src + ((size_t)yy * w + clampi(x + kx, 0, w - 1)) * 3;
                    b0 += p[0]; b1 += p[1]; b2 += p[2];
                }
            }
            const float off0 = (float)c[0] - (float)b0 / (float)(ksize * ksize);
            const float off1 = (float)c[1] - (float)b1 / (float)(ksize * ksize);
            const float off2 = (float)c[2] - (float)b2 / (float)(ksize * ksize);
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, sk = 0.f;
            for (int ky = -radius; ky <= radius; ky++) {
                const int yy = clampi(y + ky, 0, h - 1);
                for (int kx = -radius; kx <= radius; kx++) {
                    const uint8_t* p = src + ((size_t)yy * w + clampi(x + kx, 0, w - 1)) * 3;
                    const float d0 = (float)((int)p[0] - (int)c[0]) - off0;
                    const float d1 = (float)((int)p[1] - (int)c[1]) - off1;
                    const float d2 = (float)((int)p[2] - (int)c[2]) - off2;
                    const float dist = fabsf(d0) + fabsf(d1) + fabsf(d2);
                    const float k = ks[(ky + radius) * ksize + (kx + radius)] * lut[(int)dist];
                    s0 = mad((float)p[0], k, s0, arith);
                    s1 = mad((float)p[1], k, s1, arith);
                    s2 = mad((float)p[2], k, s2, arith);
                    sk += k;
                }
            }
            uint8_t* o = dst + ((size_t)y * w + x) * 3;
            o[0] = (uint8_t)(s0 / sk + 0.5f);
            o[1] = (uint8_t)(s1 / sk + 0.5f);
            o[2] = (uint8_t)(s2 / sk + 0.5f);
        }
    }
    free(ks);
    free(lut);
}

void vipo_adaptive_bilateral_filter(const uint8_t* src, uint8_t* dst, int w, int h, int ksize, float sigma_space,
                                    float sigma_color, int arith) {
    vipo_adaptive_bilateral_filter_rows(src, dst, w, h, ksize, sigma_space, sigma_color, arith, 0, h);
}

/* Gradient magnitude: src/gradient_impl.cu:7-66 ; include/cpp/gradient.hpp:13-85 ; test/gradient.cu:9-34.
 * Central differences with clamped coordinates, no 1/2 factor. */
void vipo_gradient_u8(const uint8_t* src, float* dst, int w, int h, int ch) {
#pragma omp parallel for
    for (int y = 0; y < h; y++) {
        const int ym = clampi(y - 1, 0, h - 1), yp = clampi(y + 1, 0, h - 1);
        for (int x = 0; x < w; x++) {
            const int xm = clampi(x - 1, 0, w - 1), xp = clampi(x + 1, 0, w - 1);
            int acc = 0; /* <= 6*255^2, exact in fp32 too */
            for (int c = 0; c < ch; c++) {
                const int dx = (int)src[((size_t)y * w + xp) * ch + c] - (int)src[((size_t)y * w + xm) * ch + c];
                const int dy = (int)src[((size_t)yp * w + x) * ch + c] - (int)src[((size_t)ym * w + x) * ch + c];
                acc += dx * dx + dy * dy;
            }
            dst[(size_t)y * w + x] = sqrtf((float)acc);
        }
    }
}

/* f32 input.  CUDA (src/gradient_impl.cu:52-65): del_x = sum_c fma(d,d,acc), del_y likewise, sqrtf(del_x+del_y).
 * CPU (include/cpp/gradient.hpp:44-53): sum += h*h + v*v per channel. */
void vipo_gradient_f32(const float* src, float* dst, int w, int h, int ch, int arith) {
#pragma omp parallel for
    for (int y = 0; y < h; y++) {
        const int ym = clampi(y - 1, 0, h - 1), yp = clampi(y + 1, 0, h - 1);
        for (int x = 0; x < w; x++) {
            const int xm = clampi(x - 1, 0, w - 1), xp = clampi(x + 1, 0, w - 1);
            if (arith == VIPO_ARITH_CUDA) {
                float gx = 0.f, gy = 0.f;
                for (int c = 0; c < ch; c++) {
                    const float dx = src[((size_t)y * w + xp) * ch + c] - src[((size_t)y * w + xm) * ch + c];
                    gx = fmaf(dx, dx, gx);
                }
                for (int c = 0; c < ch; c++) {
                    const float dy = src[((size_t)yp * w + x) * ch + c] - src[((size_t)ym * w + x) * ch + c];
                    gy = fmaf(dy, dy, gy);
                }
                dst[(size_t)y * w + x] = sqrtf(gx + gy);
            } else {
                float sum = 0.f;
                for (int c = 0; c < ch; c++) {
                    const float dx = src[((size_t)y * w + xp) * ch + c] - src[((size_t)y * w + xm) * ch + c];
                    const float dy = src[((size_t)yp * w + x) * ch + c] - src[((size_t)ym * w + x) * ch + c];
                    volatile float a = dx * dx;
                    volatile float b = dy * dy;
                    volatile float s = a + b;
                    sum += s;
                }
                dst[(size_t)y * w + x] = sqrtf(sum);
            }
        }
    }
}

/* Box blur + relative total variation: src/bilateral_texture_filter_impl.cu:10-104 ;
 * include/cpp/bilateral_texture_filter.hpp:17-83.  Sum of magnitudes is accumulated in fp32 in raster order. */
void vipo_blur_rtv(const uint8_t* image, const float* magnitude, float* blurred, float* rtv, int w, int h, int ksize,
                   int arith) {
    const int radius = ksize / 2;
#pragma omp parallel for
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            float s0 = 0.f, s1 = 0.f, s2 = 0.f;
            float imax = 0.f, imin = 256.f, mmax = 0.f, msum = 0.f;
            for (int ky = -radius; ky <= radius; ky++) {
                const int yy = clampi(y + ky, 0, h - 1);
                for (int kx = -radius; kx <= radius; kx++) {
                    const int xx = clampi(x + kx, 0, w - 1);
                    const uint8_t* p = image + ((size_t)yy * w + xx) * 3;
                    s0 += p[0]; s1 += p[1]; s2 += p[2];
                    const float inten = (float)(p[0] + p[1] + p[2]) / 3.f;
                    imax = fmaxf(imax, inten);
                    imin = fminf(imin, inten);
                    const float m = magnitude[(size_t)yy * w + xx];
                    mmax = fmaxf(mmax, m);
                    msum += m;
                }
            }
            float* b = blurred + ((size_t)y * w + x) * 3;
            b[0] = s0 / (float)(ksize * ksize);
            b[1] = s1 / (float)(ksize * ksize);
            b[2] = s2 / (float)(ksize * ksize);
            const float num = (imax - imin) * mmax;
            if (arith == VIPO_ARITH_CUDA) rtv[(size_t)y * w + x] = (float)((double)num / ((double)msum + 1e-9));
            else rtv[(size_t)y * w + x] = num / (msum + 1e-9f);
        }
    }
}

/* Guide image: src/bilateral_texture_filter_impl.cu:106-177 ; include/cpp/bilateral_texture_filter.hpp:85-145.
 * argmin over the k x k window, first in raster order wins; alpha in (-1,0]; blend, +0.5, truncate, clamp.
 * CUDA contracts the blend as fma(alpha, Bq, (1-alpha)*Bp) (SASS of the reference kernel). */
void vipo_guide(const float* blurred, const float* rtv, uint8_t* guide, int w, int h, int ksize, int arith) {
    const int radius = ksize / 2;
    const float sigma_alpha = 1.f / (float)(5 * ksize);
#pragma omp parallel for
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            float rmin = (arith == VIPO_ARITH_CUDA) ? 1e10f : 3.402823466e+38f;
            int mx = 0, my = 0;
            for (int ky = -radius; ky <= radius; ky++) {
                const int yy = clampi(y + ky, 0, h - 1);
                for (int kx = -radius; kx <= radius; kx++) {
                    const int xx = clampi(x + kx, 0, w - 1);
                    const float v = rtv[(size_t)yy * w + xx];
                    if (rmin > v) { rmin = v; mx = xx; my = yy; }
                }
            }
            const float alpha = 2 / (1 + expf(sigma_alpha * (rtv[(size_t)y * w + x] - rtv[(size_t)my * w + mx]))) - 1.f;
            const float* bq = blurred + ((size_t)my * w + mx) * 3;
            const float* bp = blurred + ((size_t)y * w + x) * 3;
            uint8_t* g = guide + ((size_t)y * w + x) * 3;
            for (int c = 0; c < 3; c++) {
                float t;
                if (arith == VIPO_ARITH_CUDA) {
                    t = fmaf(alpha, bq[c], (1 - alpha) * bp[c]) + 0.5f;
                } else {
                    volatile float a = alpha * bq[c];
                    volatile float b = (1 - alpha) * bp[c];
                    volatile float s = a + b;
                    t = s + 0.5f;
                }
                g[c] = (uint8_t)clampi((int)t, 0, 255);
            }
        }
    }
}

/* Bilateral texture filter: src/bilateral_texture_filter_impl.cu:199-214 ;
 * include/cpp/bilateral_texture_filter.hpp:151-171 with the JBF step = the reference's own joint bilateral
 * (ksize 2k-1, sigma_space k-1, sigma_color sqrt(3)), as in the reference's CUDA path (.cu:188,210). */
void vipo_bilateral_texture_filter(const uint8_t* src, uint8_t* dst, int w, int h, int ksize, int nitr, int arith) {
    const size_t n = (size_t)w * h;
    uint8_t* src_n = (uint8_t*)malloc(n * 3);
    uint8_t* guide = (uint8_t*)malloc(n * 3);
    float* mag = (float*)malloc(n * sizeof(float));
    float* rtv = (float*)malloc(n * sizeof(float));
    float* blurred = (float*)malloc(n * 3 * sizeof(float));
    const float sigma_color = (arith == VIPO_ARITH_CUDA) ? 1.73205080757f : (float)sqrt(3.0);
    memcpy(dst, src, n * 3);
    for (int it = 0; it < nitr; it++) {
        memcpy(src_n, dst, n * 3);
        vipo_gradient_u8(src_n, mag, w, h, 3);
        vipo_blur_rtv(src_n, mag, blurred, rtv, w, h, ksize, arith);
        vipo_guide(blurred, rtv, guide, w, h, ksize, arith);
        vipo_joint_bilateral_filter(src_n, guide, dst, w, h, 2 * ksize - 1, (float)(ksize - 1), sigma_color, arith);
    }
    free(src_n); free(guide); free(mag); free(rtv); free(blurred);
}

#ifdef _OPENMP
#include <omp.h>
void vipo_set_num_threads(int n) { omp_set_num_threads(n > 0 ? n : omp_get_num_procs()); }
int vipo_get_num_threads(void) { return omp_get_max_threads(); }
#else
void vipo_set_num_threads(int n) { (void)n; }
int vipo_get_num_threads(void) { return 1; }
#endif
