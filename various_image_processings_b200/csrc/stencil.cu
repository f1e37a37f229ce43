// Bilateral-family stencil kernels for sm_100a: bilateral (BF), joint bilateral (JBF) and adaptive bilateral
// (ABF) filters.  Replaces the reference's bilateral_filter_kernel / joint_bilateral_filter_kernel
// (src/bilateral_filter_impl.cu:7-96, 98-202) and adaptive_bilateral_filter_kernel
// (src/adaptive_bilateral_filter_impl.cu:7-115).  New design, no code shared with the reference:
//
//  * one persistent CTA per SM of up to three 256-thread warp groups; every group has its own tile buffer and named
//    barrier and pulls 8P x 32 output tiles from a counter in shared memory, all groups share one LUT.  A group stages
//    its tile plus the radius halo ONCE as packed 32-bit pixels (B | G<<8 | R<<16; JBF: {guide, src} pairs), four
//    pixels per thread from aligned 32-bit loads, so a tap is one VABSDIFF4.U8.ACC instead of six byte loads and
//    three abs;
//  * a lane owns one image ROW of the warp's P x 32 sub-tile and P horizontally adjacent outputs: with an odd
//    tile pitch all 32 lanes hit distinct banks, and one neighbour load + unpack feeds up to P taps;
//  * the range LUT is replicated 32x in shared memory (entry i of lane l at word 32*i + l) so the
//    data-dependent gather is bank-conflict free for ANY image content (random data would otherwise
//    serialise ~3.5x);
//  * the zero-weight taps of the circular mask are never issued (bit-neutral: the reference adds +0.0f for
//    them): a window row of 2*hw+1 taps is cut into runs of 8 taps plus one odd remainder
//    (7/5/3/1).  Only those five run bodies exist as code (~30 KB for BF), whatever the radius: the first
//    version unrolled every half width separately and spent more time waiting for instruction fetch than for
//    anything else (ncu: stall_no_instruction 2.96 per issue at r=15, profiles/r01_*);
//  * accumulation order and rounding are those of the reference kernel: raster order, w = ks * lut (one
//    rounding), fmaf(I, w, sum) per channel, sumk += w, IEEE divide, +0.5f, truncate.  Hence the output is
//    bit-identical to the reference's CUDA build, not merely within 1 LSB.
//
// ABF: the k x k box mean comes from exact integer sliding sums over the shared tile (the reference brute
// forces k^2 float adds of integers, which is the same value); the range index is
// (int)(|d0|+|d1|+|d2|), d = (q - p) - offset, in the reference's fp32 operation order.  The truncation is done
// on the FMA pipe (add 2^23 with round-toward-zero; the mantissa then holds floor(dist) exactly) instead of
// F2I on the quarter-rate XU pipe.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <type_traits>

#include "vip_internal.h"

namespace vip {
namespace {

constexpr int kThreads = 256;            // threads of one warp group (one tile in flight per group)
constexpr int kWarps = kThreads / 32;
constexpr int kMaxGroups = 4;            // most warp groups any mode runs per CTA (scheduler slots, mbarriers)
// Warp groups per CTA and hence registers per thread (one CTA per SM): BF / JBF 3 groups = 24 warps at <= 80 registers;
// ABF 4 groups = 32 warps at <= 64 registers.  Measured on B200 (profiles/r02f_groups_ab.txt, all bit-exact): 4 groups
// make ABF 4K r=7 1.4 % faster (0.794 -> 0.783 ms) but BF 8K r=15 4 % slower (6.855 -> 7.142 ms; the 64-register run
// body needs 4.5 % more instructions) and JBF 1 % slower.
__host__ __device__ constexpr int max_groups(const int mode) { return mode == kABF ? 4 : 3; }
constexpr int kSmemBudget = 232448 - 1024;  // opt-in maximum of dynamic shared memory per CTA (227 KB) minus the driver's 1 KB
constexpr int kTileRows = 32;  // one lane per output row
constexpr int kMaxTiledRadius = 15;

template <int MODE>
struct Elem {
    using type = std::uint32_t;
};
template <>
struct Elem<kJBF> {
    using type = uint2;  // x = guide pixel, y = src pixel
};

// Packed fp32x2 (sm_100 FFMA2 / FADD2).  {s0,s1} of one output accumulate {f0,f1} * w with the weight as the broadcast
// scalar operand (FFMA2 Rd, Ra.F32x2, Rb.F32, Rd); s2 and sk are paired over two adjacent OUTPUTS instead:
// {s2_i, s2_i+1} += {w_i, w_i+1} * f2 (f2 broadcast) and {sk_i, sk_i+1} += {w_i, w_i+1} (fma(w, 1, sk) rounds like
// sk + w; ptxas emits FADD2).  The FMUL results land in an aligned register pair, so no packing MOVs are needed (the
// first version paired {s2, sk} with the multiplicand {f2, 1.0f} and spent ~1 MOV per tap building that pair).
// Measured on B200 (tools/kbench.py): a packed instruction occupies the dispatch port like two scalar ones, so packing
// pays through the instructions it removes: BF 8K r=15 7.39 -> 7.00 ms, JBF 17x17 0.596 -> 0.565 ms.  ABF packs
// everything across two adjacent outputs (see Acc<kABF>): 4K r=7 0.867 -> 0.833 ms.
__device__ __forceinline__ unsigned long long pack2(const float lo, const float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(const unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void fma2(unsigned long long& acc, const unsigned long long a, const unsigned long long b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}

__device__ __forceinline__ unsigned long long add2(const unsigned long long a, const unsigned long long b) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

template <int MODE, int P>
struct Acc {
    unsigned long long s01[P];      // {s0, s1} of output i: multiplicand {f0, f1}, weight as the broadcast scalar
    unsigned long long s2[P / 2];   // {s2 of output 2p, s2 of output 2p+1}: multiplicand {w_2p, w_2p+1}, f2 broadcast
    unsigned long long sk[P / 2];   // {sk of output 2p, sk of output 2p+1}: + {w_2p, w_2p+1}
    std::uint32_t c[P];             // packed centre (range source)
};

// ABF: every quantity is a pair over two horizontally adjacent outputs (2p, 2p+1), so that one packed fp32x2
// instruction serves two taps: the neighbour value is the broadcast scalar operand, the per-output values
// (-centre, -offset, weights, sums) are the packed ones.
template <int P>
struct Acc<kABF, P> {
    unsigned long long s[4][P / 2];     // s0, s1, s2, sk
    unsigned long long nfp[3][P / 2];   // -(centre as float), per channel
    unsigned long long noff[3][P / 2];  // -(centre - box mean)
};

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

__device__ __forceinline__ std::uint32_t load_packed(const std::uint8_t* __restrict__ p) {
    return static_cast<std::uint32_t>(__ldg(p)) | (static_cast<std::uint32_t>(__ldg(p + 1)) << 8) |
           (static_cast<std::uint32_t>(__ldg(p + 2)) << 16);
}

// Four consecutive packed BGR pixels (12 bytes at ANY byte alignment) -> four 32-bit B|G<<8|R<<16 words, from aligned
// 32-bit loads: the 16-byte window that contains them is read as 3 (aligned case) or 4 words, realigned with three
// funnel shifts and cut with four more.  The 4th word is only touched when the data straddle it.
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

// ---- TMA (bulk asynchronous copy) tile staging, BF only, opt-in (VIP_STENCIL_TMA=1; DESIGN.md 3.3 has the A/B) ----------
// One bulk copy per tile row moves the row's raw bytes (16-byte aligned superset) global -> shared through the TMA unit,
// completion counted on an mbarrier; the group then repacks bytes -> 32-bit pixels from shared memory.
__device__ __forceinline__ void mbar_init(const std::uint32_t mbar, const int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(const std::uint32_t mbar, const std::uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(const std::uint32_t mbar, const std::uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}" ::"r"(mbar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(const std::uint32_t smem_dst, const void* gmem_src, const std::uint32_t bytes,
                                         const std::uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst),
                 "l"(gmem_src), "r"(bytes), "r"(mbar)
                 : "memory");
}
// Four packed BGR pixels (12 bytes at any byte offset) from SHARED memory: the twin of load_packed4
__device__ __forceinline__ void load_packed4_shared(const std::uint32_t addr, std::uint32_t out[4]) {
    const unsigned m = addr & 3u;
    const std::uint32_t a0 = addr - m;
    std::uint32_t w0, w1, w2, w3 = 0u;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(a0));
    asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(a0));
    asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(w2) : "r"(a0));
    if (m) asm volatile("ld.shared.u32 %0, [%1+12];" : "=r"(w3) : "r"(a0));
    const unsigned sh = 8u * m;
    const std::uint32_t v0 = __funnelshift_r(w0, w1, sh), v1 = __funnelshift_r(w1, w2, sh), v2 = __funnelshift_r(w2, w3, sh);
    out[0] = v0 & 0x00ffffffu;
    out[1] = __funnelshift_r(v0, v1, 24) & 0x00ffffffu;
    out[2] = __funnelshift_r(v1, v2, 16) & 0x00ffffffu;
    out[3] = v2 >> 8;
}

// How the range table sits in shared memory (template parameter LK of the kernels):
//   kLutFull     every entry 32 times (128 B): lane l reads copy l, any 32 indices are conflict-free; index not clamped
//   kLutClamp    same, index clamped to lut_n-1 (a zero entry standing for the cut all-zero tail)
//   kLutNarrow   every entry 8 times (32 B), clamped: for tables too long for 32 copies next to the tile buffers (ABF
//                with a wide sigma_color: 1536 entries x 128 B = 192 KB); lanes l and l+8k share a copy, so two
//                different indices can meet in a bank (<= 4-way, ~1.5 on average) -- only the LDS of the tap replays
//   kLutPremul   BF / JBF with a SHORT table (sigma_color of a few levels: the joint bilateral step of the texture filter has
//                sqrt 3, 27 entries): one 32x replicated, clamped row PER DISTINCT SPATIAL WEIGHT, holding RN(ks * lut[i])
//                -- the very float the tap's FMUL would produce -- so the FMUL disappears from the tap (7 -> 6 issue slots);
//                the weight rows then hold the byte offset of each tap's table row instead of its weight
enum LutKind { kLutFull = 0, kLutClamp = 1, kLutNarrow = 2, kLutPremul = 3 };
__host__ __device__ constexpr bool lut_clamped(const int lk) { return lk != kLutFull; }
__host__ __device__ constexpr int lut_shift(const int lk) { return lk == kLutNarrow ? 5 : 7; }  // log2(bytes per entry)
// bits(2^23 + i) << shift, modulo 2^32, for i = 0: what abf_weight's shifted float bits carry on top of (i << shift)
__host__ __device__ constexpr std::uint32_t abf_index_bias(const int lk) { return 0x4b000000u << lut_shift(lk); }

__device__ __forceinline__ float lds_f32(const std::uint32_t addr) {
    float v;
    asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

// u8 -> float: I2F.U8 with a byte selector (one XU-pipe instruction).  The XU-free alternative (PRMT to the bits of
// 2^23 + byte, then FADD -2^23) costs a second issue slot and measured 3-7 % slower (tools/kbench.py, B200).
template <int BYTE>
__device__ __forceinline__ float byte_to_float(const std::uint32_t v) {
    return static_cast<float>((v >> (8 * BYTE)) & 0xffu);
}

// A PIECE of one window row for the P outputs of this thread: the NN consecutive tile elements t[0 .. NN) ("neighbours")
// and, of the taps they form with the outputs -- neighbour n is tap j = n - i of output i --, those with JLO <= j < JHI.
// t points at the tile element of (first output, tap 0 of the piece), wp at the weight of tap 0 (16-byte aligned; taps
// down to JLO lie before it).  Per accumulator the taps are visited in increasing j, i.e. increasing kx.
//   a whole short row of L taps (or a run of 8 of the old decomposition): NN = P + L - 1, JLO = 0, JHI = L
//   sliding decomposition of a long row (do_row): first 8 neighbours <8, 0, 8>, then steady blocks of 8 neighbours in
//   which every output has a tap <8, -(P-1), 8>, then the last r taps and the P - 1 neighbours only the later outputs
//   still see <r + P - 1, -(P-1), r>.  Every tile element of the row is then loaded and converted ONCE (the runs of 8
//   re-read the 7 elements two runs share: 0.25 instead of 0.16 elements per tap at r = 15).
__host__ __device__ constexpr int floor4(const int v) { return v >= 0 ? (v & ~3) : -((-v + 3) & ~3); }

template <int MODE, int P, int NN, int JLO, int JHI, int LK>
__device__ __forceinline__ void run_taps(const typename Elem<MODE>::type* __restrict__ t, const float* __restrict__ wp,
                                         const std::uint32_t lut_addr, const std::uint32_t clamp_i, const float clamp_f,
                                         Acc<MODE, P>& A) {
    constexpr int WOFF = floor4(JLO), NW = (JHI - WOFF + 3) & ~3;   // weights w[j - WOFF] of taps WOFF .. WOFF + NW
    float w[NW];
#pragma unroll
    for (int j = 0; j < NW / 4; ++j) {
        const float4 v = reinterpret_cast<const float4*>(wp + WOFF)[j];
        w[4 * j + 0] = v.x;
        w[4 * j + 1] = v.y;
        w[4 * j + 2] = v.z;
        w[4 * j + 3] = v.w;
    }
    const unsigned long long one2 = pack2(1.f, 1.f);
    if constexpr (LK == kLutPremul) {
        // w[] carries the byte offset of each tap's table row; fold this lane's copy address in once per piece
#pragma unroll
        for (int j = JLO; j < JHI; ++j) w[j - WOFF] = __uint_as_float(__float_as_uint(w[j - WOFF]) + lut_addr);
    }
    auto weight = [&](const std::uint32_t q, const std::uint32_t c, const float ks) {
        std::uint32_t sad = __vsadu4(q, c);
        if constexpr (lut_clamped(LK)) sad = min(sad, clamp_i);
        if constexpr (LK == kLutPremul)
            return lds_f32(__float_as_uint(ks) + (sad << lut_shift(LK)));   // the row holds RN(ks * lut[sad]) already
        else
            return __fmul_rn(ks, lds_f32(lut_addr + (sad << lut_shift(LK))));
    };
#pragma unroll
    for (int n = 0; n < NN; ++n) {
        std::uint32_t q, v;
        if constexpr (MODE == kJBF) {
            const uint2 e = t[n];
            q = e.x;
            v = e.y;
        } else {
            q = t[n];
            v = q;
        }
        const float f0 = byte_to_float<0>(v);
        const float f1 = byte_to_float<1>(v);
        const float f2 = byte_to_float<2>(v);
        const unsigned long long f01 = pack2(f0, f1);
#pragma unroll
        for (int p = 0; p < P / 2; ++p) {
            const int j = n - 2 * p;  // tap of output 2p; output 2p+1 sees tap j-1
            const bool lo = j >= JLO && j < JHI, hi = j - 1 >= JLO && j - 1 < JHI;
            float wl = 0.f, wh = 0.f;
            // VIP_PACK (A/B builds): 2 = everything packed (default), 1 = {s0,s1} packed but s2 / sk scalar, 0 = all scalar
#ifndef VIP_PACK
#define VIP_PACK 2
#endif
            auto acc01 = [&](unsigned long long& s01, const float wt) {
                if constexpr (VIP_PACK >= 1) {
                    fma2(s01, f01, pack2(wt, wt));
                } else {
                    float a0, a1;
                    unpack2(s01, a0, a1);
                    a0 = __fmaf_rn(f0, wt, a0);
                    a1 = __fmaf_rn(f1, wt, a1);
                    s01 = pack2(a0, a1);
                }
            };
            if (lo) {
                wl = weight(q, A.c[2 * p], w[lo ? j - WOFF : 0]);
                acc01(A.s01[2 * p], wl);
            }
            if (hi) {
                wh = weight(q, A.c[2 * p + 1], w[hi ? j - 1 - WOFF : 0]);
                acc01(A.s01[2 * p + 1], wh);
            }
            if (VIP_PACK < 2 && lo && hi) {
                float x2, y2, xk, yk;
                unpack2(A.s2[p], x2, y2);
                unpack2(A.sk[p], xk, yk);
                x2 = __fmaf_rn(f2, wl, x2);
                y2 = __fmaf_rn(f2, wh, y2);
                xk = __fadd_rn(xk, wl);
                yk = __fadd_rn(yk, wh);
                A.s2[p] = pack2(x2, y2);
                A.sk[p] = pack2(xk, yk);
            } else if (lo && hi) {
                const unsigned long long ww = pack2(wl, wh);
                fma2(A.s2[p], ww, pack2(f2, f2));
                fma2(A.sk[p], ww, one2);
            } else if (lo || hi) {
                float x2, y2, xk, yk;
                unpack2(A.s2[p], x2, y2);
                unpack2(A.sk[p], xk, yk);
                if (lo) {
                    x2 = __fmaf_rn(f2, wl, x2);
                    xk = __fadd_rn(xk, wl);
                } else {
                    y2 = __fmaf_rn(f2, wh, y2);
                    yk = __fadd_rn(yk, wh);
                }
                A.s2[p] = pack2(x2, y2);
                A.sk[p] = pack2(xk, yk);
            }
        }
    }
}

// ABF weight of one tap from its three shifted differences (src/adaptive_bilateral_filter_impl.cu:67-73, 103-105)
template <int LK>
__device__ __forceinline__ float abf_weight(const float d0, const float d1, const float d2, const float ks,
                                            const std::uint32_t lut_addr, const float clamp_f) {
    float dist = __fadd_rn(__fadd_rn(fabsf(d0), fabsf(d1)), fabsf(d2));
    if constexpr (lut_clamped(LK)) dist = fminf(dist, clamp_f);
    // (int)dist: bits(dist +rz 2^23) = 0x4b000000 + floor(dist); << 7 leaves 0x80000000 + 128*floor(dist),
    // and lut_addr already carries the -0x80000000 (32-bit wrap-around; abf_index_bias for the other entry sizes)
    const std::uint32_t addr = lut_addr + (__float_as_uint(__fadd_rz(dist, 8388608.f)) << lut_shift(LK));
    return __fmul_rn(ks, lds_f32(addr));
}

// The same for the two outputs of a pair: the truncating add of 2^23 is ONE packed add.rz.f32x2 (each half rounds like
// the scalar instruction), everything else as in abf_weight.  -DVIP_ABF_RZ2=0 builds the scalar form (A/B).
#ifndef VIP_ABF_RZ2
#define VIP_ABF_RZ2 1
#endif
template <int LK>
__device__ __forceinline__ void abf_weight2(const float d0l, const float d1l, const float d2l, const float d0h,
                                            const float d1h, const float d2h, const float ksl, const float ksh,
                                            const std::uint32_t lut_addr, const float clamp_f, float& wl, float& wh) {
    float distl = __fadd_rn(__fadd_rn(fabsf(d0l), fabsf(d1l)), fabsf(d2l));
    float disth = __fadd_rn(__fadd_rn(fabsf(d0h), fabsf(d1h)), fabsf(d2h));
    if constexpr (lut_clamped(LK)) {
        distl = fminf(distl, clamp_f);
        disth = fminf(disth, clamp_f);
    }
    unsigned long long t;
    asm("add.rz.f32x2 %0, %1, %2;" : "=l"(t) : "l"(pack2(distl, disth)), "l"(pack2(8388608.f, 8388608.f)));
    float tl, th;
    unpack2(t, tl, th);
    wl = __fmul_rn(ksl, lds_f32(lut_addr + (__float_as_uint(tl) << lut_shift(LK))));
    wh = __fmul_rn(ksh, lds_f32(lut_addr + (__float_as_uint(th) << lut_shift(LK))));
}
// ABF run: outputs are handled in pairs (2p, 2p+1).  Neighbour n is tap j = n - 2p of the first and tap j - 1 of
// the second output; when both are inside the run the differences and the four accumulations are packed fp32x2
// operations (same IEEE roundings per half as the scalar ones), otherwise the one valid output goes scalar.
template <int P, int NN, int JLO, int JHI, int LK>
__device__ __forceinline__ void run_taps_abf(const std::uint32_t* __restrict__ t, const float* __restrict__ wp,
                                             const std::uint32_t lut_addr, const float clamp_f, Acc<kABF, P>& A) {
    constexpr int WOFF = floor4(JLO), NW = (JHI - WOFF + 3) & ~3;
    float w[NW];
#pragma unroll
    for (int j = 0; j < NW / 4; ++j) {
        const float4 v = reinterpret_cast<const float4*>(wp + WOFF)[j];
        w[4 * j + 0] = v.x;
        w[4 * j + 1] = v.y;
        w[4 * j + 2] = v.z;
        w[4 * j + 3] = v.w;
    }
    const unsigned long long one2 = pack2(1.f, 1.f);
#pragma unroll
    for (int n = 0; n < NN; ++n) {
        const std::uint32_t v = t[n];
        const float f0 = byte_to_float<0>(v);
        const float f1 = byte_to_float<1>(v);
        const float f2 = byte_to_float<2>(v);
#pragma unroll
        for (int p = 0; p < P / 2; ++p) {
            const int j = n - 2 * p;  // tap of output 2p; output 2p+1 sees tap j-1
            const bool lo = j >= JLO && j < JHI, hi = j - 1 >= JLO && j - 1 < JHI;
            if (lo && hi) {
                const unsigned long long d0 = add2(add2(pack2(f0, f0), A.nfp[0][p]), A.noff[0][p]);
                const unsigned long long d1 = add2(add2(pack2(f1, f1), A.nfp[1][p]), A.noff[1][p]);
                const unsigned long long d2 = add2(add2(pack2(f2, f2), A.nfp[2][p]), A.noff[2][p]);
                float d0l, d0h, d1l, d1h, d2l, d2h;
                unpack2(d0, d0l, d0h);
                unpack2(d1, d1l, d1h);
                unpack2(d2, d2l, d2h);
                float wl, wh;
                if constexpr (VIP_ABF_RZ2) {
                    abf_weight2<LK>(d0l, d1l, d2l, d0h, d1h, d2h, w[j - WOFF], w[j - 1 - WOFF], lut_addr, clamp_f, wl, wh);
                } else {
                    wl = abf_weight<LK>(d0l, d1l, d2l, w[j - WOFF], lut_addr, clamp_f);
                    wh = abf_weight<LK>(d0h, d1h, d2h, w[j - 1 - WOFF], lut_addr, clamp_f);
                }
                const unsigned long long ww = pack2(wl, wh);
                fma2(A.s[0][p], ww, pack2(f0, f0));
                fma2(A.s[1][p], ww, pack2(f1, f1));
                fma2(A.s[2][p], ww, pack2(f2, f2));
                fma2(A.s[3][p], ww, one2);
            } else if (lo || hi) {
                const int jj = lo ? j : j - 1;
                float nf[3], no[3], a[4], other[4];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float x, y;
                    unpack2(A.nfp[c][p], x, y);
                    nf[c] = lo ? x : y;
                    unpack2(A.noff[c][p], x, y);
                    no[c] = lo ? x : y;
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float x, y;
                    unpack2(A.s[c][p], x, y);
                    a[c] = lo ? x : y;
                    other[c] = lo ? y : x;
                }
                const float d0 = __fadd_rn(__fadd_rn(f0, nf[0]), no[0]);
                const float d1 = __fadd_rn(__fadd_rn(f1, nf[1]), no[1]);
                const float d2 = __fadd_rn(__fadd_rn(f2, nf[2]), no[2]);
                const float wt = abf_weight<LK>(d0, d1, d2, w[jj - WOFF], lut_addr, clamp_f);
                a[0] = __fmaf_rn(f0, wt, a[0]);
                a[1] = __fmaf_rn(f1, wt, a[1]);
                a[2] = __fmaf_rn(f2, wt, a[2]);
                a[3] = __fadd_rn(a[3], wt);
#pragma unroll
                for (int c = 0; c < 4; ++c) A.s[c][p] = lo ? pack2(a[c], other[c]) : pack2(other[c], a[c]);
            }
        }
    }
}

// One window row of len = 2*hw+1 taps (odd).  Rows of up to 7 taps are one piece.  Longer rows are either cut into runs
// of 8 taps plus the odd remainder, or SLIDE along the tile row (see run_taps) so that every tile element is loaded and
// converted once.  Measured on B200 (profiles/r02c_sliding_ab.txt, all bit-exact): sliding executes 3.6 % fewer
// instructions at BF 8K r=15 (XU pipe 58 -> 38 %) for the SAME 6.85 ms -- the tap loop is bound by the 2-cycle packed
// fp32x2 instructions on the FMA pipe (ncu: FFMA2 / FADD2 carry 40 % of the stall samples, reason math-pipe throttle),
// not by the instruction count -- and its 1.7x larger code shows up as instruction-fetch stalls (0.49 per issue); it
// wins 1.4 % for ABF (4 outputs per thread: the re-read share is largest), loses 3 % at 1080p BF k=9.  Hence: sliding
// for ABF, runs of 8 for BF / JBF.  -DVIP_SLIDING=0 / =1 forces one decomposition everywhere (A/B builds).
template <int MODE>
__host__ __device__ constexpr bool sliding_rows() {
#if defined(VIP_SLIDING)
    return VIP_SLIDING != 0;
#else
    return MODE == kABF;
#endif
}
template <int MODE, int P, int LK>
__device__ __forceinline__ void do_row(int len, const typename Elem<MODE>::type* __restrict__ t,
                                       const float* __restrict__ wp, const std::uint32_t lut_addr,
                                       const std::uint32_t clamp_i, const float clamp_f, Acc<MODE, P>& A) {
    constexpr bool kSlide = sliding_rows<MODE>();
    auto piece = [&](auto nn, auto jlo, auto jhi) {
        constexpr int NN = decltype(nn)::value, JLO = decltype(jlo)::value, JHI = decltype(jhi)::value;
        if constexpr (MODE == kABF) {
            run_taps_abf<P, NN, JLO, JHI, LK>(t, wp, lut_addr, clamp_f, A);
        } else {
            run_taps<MODE, P, NN, JLO, JHI, LK>(t, wp, lut_addr, clamp_i, clamp_f, A);
        }
    };
    auto run = [&](auto lc) {   // a whole row / run of L taps
        constexpr int L = decltype(lc)::value;
        piece(std::integral_constant<int, P + L - 1>{}, std::integral_constant<int, 0>{}, std::integral_constant<int, L>{});
    };
    auto tail = [&](auto rc) {  // the last r taps of a sliding row + the P - 1 elements only the later outputs still see
        constexpr int R = decltype(rc)::value;
        piece(std::integral_constant<int, R + P - 1>{}, std::integral_constant<int, -(P - 1)>{}, std::integral_constant<int, R>{});
    };
    if (kSlide && len >= 9) {
        piece(std::integral_constant<int, 8>{}, std::integral_constant<int, 0>{}, std::integral_constant<int, 8>{});
        len -= 8, t += 8, wp += 8;
#pragma unroll 1
        for (; len >= 8; len -= 8, t += 8, wp += 8)
            piece(std::integral_constant<int, 8>{}, std::integral_constant<int, -(P - 1)>{}, std::integral_constant<int, 8>{});
        switch (len) {
            case 7: tail(std::integral_constant<int, 7>{}); break;
            case 5: tail(std::integral_constant<int, 5>{}); break;
            case 3: tail(std::integral_constant<int, 3>{}); break;
            case 1: tail(std::integral_constant<int, 1>{}); break;
            default: break;
        }
        return;
    }
    if constexpr (!kSlide) {
#pragma unroll 1
        for (; len >= 8; len -= 8, t += 8, wp += 8) run(std::integral_constant<int, 8>{});
    }
    switch (len) {
        case 7: run(std::integral_constant<int, 7>{}); break;
        case 5: run(std::integral_constant<int, 5>{}); break;
        case 3: run(std::integral_constant<int, 3>{}); break;
        case 1: run(std::integral_constant<int, 1>{}); break;
        default: break;
    }
}

struct SmemLayout {
    int lut_off, w_off, hw_off, sched_off, tile_off, hs_off, bs_off, total;   // tile_off / hs_off / bs_off: group 0; group g adds g * group_stride
    int tile_pitch, tile_w, tile_h, hs_pitch, w_pitch, group_stride;
    int stage_off, stage_pitch, mbar_off;   // TMA staging (BF): raw byte rows of the tile + one mbarrier per group
};

// Raw-byte staging buffer of the TMA tile load: rows of 3 * tile_w bytes plus up to 15 bytes of alignment lead-in,
// rounded to the 16 bytes a bulk copy moves
__host__ __device__ constexpr int stage_pitch_bytes(int P, int radius) { return (3 * (kWarps * P + 2 * radius) + 15 + 15) & ~15; }

// lut_n: entries of the table as replicated into shared memory (kLutPremul: rows x entries per row)
__host__ __device__ inline SmemLayout make_layout(int mode, int P, int radius, int lut_n, int lut_kind, int n_groups,
                                                  bool tma = false) {
    SmemLayout L;
    const int k = 2 * radius + 1;
    L.tile_w = kWarps * P + 2 * radius;
    L.tile_h = kTileRows + 2 * radius;
    L.tile_pitch = L.tile_w | 1;  // odd: 32 lanes on 32 consecutive rows hit 32 distinct banks
    L.hs_pitch = (kWarps * P) | 1;
    int off = 0;
    L.lut_off = off;
    off += lut_n << lut_shift(lut_kind);
    L.w_pitch = (k + 7) & ~7;  // row ky: its 2*hw+1 non-zero weights from index 0, zero padded (= stencil_w_pitch)
    L.w_off = off;
    off += k * L.w_pitch * 4;
    L.hw_off = off;
    off += k * 4;
    off = (off + 15) & ~15;
    L.sched_off = off;  // tile counter of the CTA + the tile each group is working on (1 + kMaxGroups ints)
    off += 32;
    static_assert(1 + kMaxGroups <= 8, "scheduler slots");
    L.mbar_off = off;   // 8 bytes per group (always reserved: 32 bytes)
    off += 8 * 4;
    L.tile_off = off;
    off += L.tile_pitch * L.tile_h * (mode == kJBF ? 8 : 4);
    off = (off + 15) & ~15;
    L.hs_off = off;
    if (mode == kABF) off += L.hs_pitch * L.tile_h * 8;
    off = (off + 15) & ~15;
    L.bs_off = off;  // ABF: k x k box sums of the tile's outputs, uint4 {B, G, R, -} per output
    if (mode == kABF) off += L.hs_pitch * kTileRows * 16;
    off = (off + 15) & ~15;
    L.stage_off = off;
    L.stage_pitch = stage_pitch_bytes(P, radius);
    if (tma) off += L.stage_pitch * L.tile_h;
    off = (off + 15) & ~15;
    L.group_stride = off - L.tile_off;
    L.total = L.tile_off + n_groups * L.group_stride;
    return L;
}

// Prologue of the tiled kernels: fills the replicated range table, the spatial weight rows and the half widths.
// Every global load is issued before anything depends on it: a thread loads ONE table entry (coalesced), then each warp
// broadcasts its 32 entries by shuffle and writes their copies with conflict-free 128-byte stores.  (The first version
// loaded entry i>>5 once per copy: 32 dependent L2 round trips, ~6 us of every launch.)
template <int LK>
__device__ __forceinline__ void fill_tables(const float* __restrict__ lut, const int lut_n,
                                         const float* __restrict__ w_rows, const int* __restrict__ hw, const int k,
                                         float* __restrict__ lut_s, float* __restrict__ w_s, int* __restrict__ hw_s,
                                         const int n_w) {
    constexpr int kRep = 1 << (lut_shift(LK) - 2), kPerStore = 32 / kRep;  // copies per entry; entries per warp-wide store
    const int lane_all = threadIdx.x & 31;
    const int wbase = threadIdx.x & ~31;
    float wv = 0.f;
    int hv = 0;
    if (threadIdx.x < n_w) wv = __ldg(w_rows + threadIdx.x);
    if (threadIdx.x < k) hv = __ldg(hw + threadIdx.x);
    for (int base = 0; base < lut_n; base += blockDim.x) {
        const int e = base + threadIdx.x;
        const float v = e < lut_n ? __ldg(lut + e) : 0.f;
        const int e0 = base + wbase;  // first entry of this warp
#pragma unroll
        for (int j = 0; j < 32; j += kPerStore) {
            const int src = j + lane_all / kRep;
            const float vj = __shfl_sync(0xffffffffu, v, src);
            if (e0 + src < lut_n) lut_s[(e0 + src) * kRep + (lane_all & (kRep - 1))] = vj;
        }
    }
    if (threadIdx.x < n_w) w_s[threadIdx.x] = wv;
    for (int i = threadIdx.x + blockDim.x; i < n_w; i += blockDim.x) w_s[i] = __ldg(w_rows + i);
    if (threadIdx.x < k) hw_s[threadIdx.x] = hv;
}

// Out-of-line copy for the BF kernels: ptxas' register allocation of the tap loops depends on whether the prologue is
// inlined (measured on B200, same source otherwise: BF 8K r=15 6.83 ms out of line vs 6.89 ms inlined, while ABF and JBF
// are 0.5 - 1 % faster with it inlined).
template <int LK>
__device__ __noinline__ void fill_tables_call(const float* __restrict__ lut, const int lut_n,
                                              const float* __restrict__ w_rows, const int* __restrict__ hw, const int k,
                                              float* __restrict__ lut_s, float* __restrict__ w_s, int* __restrict__ hw_s,
                                              const int n_w) {
    fill_tables<LK>(lut, lut_n, w_rows, hw, k, lut_s, w_s, hw_s, n_w);
}

// PEER: halo rows that the strip buffer does not hold are staged straight from the neighbouring ranks' device memory
// (StencilArgs::above / below: peer or IPC-mapped pointers, i.e. NVLink loads inside the tile staging) -- the halo
// exchange of a row-strip split fused into the filter, no halo copy and no halo rows in the local buffer at all.
template <int MODE, int P, int LK, bool PEER, bool TMA = false>
__global__ void __launch_bounds__(kThreads * max_groups(MODE), 1)
    stencil_kernel(const StencilArgs a, const int tiles_x, const int n_tiles) {
    static_assert(!TMA || (MODE == kBF && !PEER), "TMA staging is built for the plain bilateral kernel only");
    using E = typename Elem<MODE>::type;
    extern __shared__ __align__(16) unsigned char smem[];
    const int radius = a.radius;
    const int k = 2 * radius + 1;
    const int n_groups = blockDim.x / kThreads;
    const int group = threadIdx.x / kThreads;  // warp group: its own tile buffer, named barrier and tile sequence
    const int lut_total = a.lut_n * a.lut_rows;   // lut_rows = 1 unless LK == kLutPremul
    const SmemLayout L = make_layout(MODE, P, radius, lut_total, LK, n_groups, TMA);
    float* const lut_s = reinterpret_cast<float*>(smem + L.lut_off);
    float* const w_s = reinterpret_cast<float*>(smem + L.w_off);
    int* const hw_s = reinterpret_cast<int*>(smem + L.hw_off);
    int* const sched = reinterpret_cast<int*>(smem + L.sched_off);
    E* const tile = reinterpret_cast<E*>(smem + L.tile_off + group * L.group_stride);
    uint2* const hs = reinterpret_cast<uint2*>(smem + L.hs_off + group * L.group_stride);
    uint4* const bs = reinterpret_cast<uint4*>(smem + L.bs_off + group * L.group_stride);

    const int tid = threadIdx.x - group * kThreads;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    auto group_sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "n"(kThreads) : "memory"); };

    // one-time per CTA (shared by all groups): replicated LUT, spatial weight rows, half widths
    if constexpr (MODE == kBF)
        fill_tables_call<LK>(a.lut, lut_total, a.w_rows, a.hw, k, lut_s, w_s, hw_s, k * L.w_pitch);
    else
        fill_tables<LK>(a.lut, lut_total, a.w_rows, a.hw, k, lut_s, w_s, hw_s, k * L.w_pitch);
    constexpr int kRepLog2 = lut_shift(LK) - 2;  // copies per entry (log2)
    if (threadIdx.x == 0) sched[0] = 0;
    [[maybe_unused]] const std::uint32_t mbar =
        static_cast<std::uint32_t>(__cvta_generic_to_shared(smem + L.mbar_off + 8 * group));
    [[maybe_unused]] const std::uint32_t stage =
        static_cast<std::uint32_t>(__cvta_generic_to_shared(smem + L.stage_off + group * L.group_stride));
    [[maybe_unused]] std::uint32_t tma_phase = 0;
    if constexpr (TMA) {
        if (tid == 0) {
            mbar_init(mbar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    // Programmatic dependent launch: the prologue above touches only this filter's constant tables, so it may run while
    // the previous kernel of the stream is still draining; everything below reads / writes images and waits for it.
    // The next kernel may be scheduled as soon as SMs free up (no-ops when launched without the attribute).
    pdl_wait();
    pdl_release();
    __syncthreads();

    // byte address of this lane's copy of entry 0; ABF folds the 2^23 bias of its index trick in (see run_taps)
    const std::uint32_t lut_addr =
        static_cast<std::uint32_t>(__cvta_generic_to_shared(lut_s + (lane & ((1 << kRepLog2) - 1)))) -
        (MODE == kABF ? abf_index_bias(LK) : 0u);
    const std::uint32_t clamp_i = static_cast<std::uint32_t>(a.lut_n - 1);
    const float clamp_f = static_cast<float>(a.lut_n - 1);
    const int W = a.width, H = a.height;
    const int tw = L.tile_w, th = L.tile_h, tp = L.tile_pitch;

    // Tiles blockIdx.x, blockIdx.x + gridDim.x, ... belong to this CTA (every SM gets the same number +-1); its warp
    // groups take them in order from a shared-memory counter, so a group that finishes early picks up the next tile.
    for (;;) {
        group_sync();  // previous tile of this group fully consumed (tile buffer and sched slot reusable)
        if (tid == 0) sched[1 + group] = atomicAdd(&sched[0], 1);
        group_sync();
        const int t = blockIdx.x + sched[1 + group] * gridDim.x;
        if (t >= n_tiles) break;
        const int tile_x = t % tiles_x, tile_y = t / tiles_x;
        const int X0 = tile_x * (kWarps * P);
        const int Y0 = a.out_y0 + tile_y * kTileRows;

        // clamp to the image, then (rows no valid output depends on) to the rows the buffers hold; byte offset of the
        // row inside a.src / a.guide -- or, for a PEER launch, relative to a.src although the row lives in a neighbour's
        // buffer (all three are device pointers of one address space, the difference is just an integer)
        auto row_offset = [&](const int ey) -> long long {
            const int y = clampi(Y0 - radius + ey, 0, H - 1);
            const long long pitch = static_cast<long long>(W) * 3;
            if constexpr (PEER) {
                if (y < a.src_y0 && a.above)
                    return (a.above - a.src) + (clampi(y, a.above_y0, a.above_y0 + a.above_rows - 1) - a.above_y0) * pitch;
                if (y >= a.src_y0 + a.src_rows && a.below)
                    return (a.below - a.src) + (clampi(y, a.below_y0, a.below_y0 + a.below_rows - 1) - a.below_y0) * pitch;
            }
            return (clampi(y, a.src_y0, a.src_y0 + a.src_rows - 1) - a.src_y0) * pitch;
        };
        // Tiles whose columns (plus one spare pixel to the right, which the 4th word of load_packed4 may touch) lie
        // inside the image are staged 4 pixels at a time from aligned 32-bit loads; border tiles and the tw % 4
        // leftover columns go pixel by pixel with clamped coordinates.
        const bool interior = X0 - radius >= 0 && X0 - radius + tw < W;
        const int vec_cols = interior ? (tw & ~3) : 0;
        if constexpr (TMA) {
            if (interior) {
                // the host checked that the image rows start 16-byte aligned: the lead-in only depends on the column
                const std::uint32_t lead = static_cast<std::uint32_t>((X0 - radius) * 3) & 15u;
                const std::uint32_t row_bytes = (lead + 3u * tw + 15u) & ~15u;
                if (warp == 0) {
                    // the previous tile's repack (generic-proxy reads of the staging rows) is complete (group_sync above);
                    // order it before the asynchronous-proxy writes of the new copies
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    if (lane == 0) mbar_expect_tx(mbar, row_bytes * th);
                    __syncwarp();
                    for (int ey = lane; ey < th; ey += 32)
                        bulk_g2s(stage + ey * L.stage_pitch, a.src + row_offset(ey) + (X0 - radius) * 3 - lead, row_bytes, mbar);
                }
                mbar_wait(mbar, tma_phase);
                tma_phase ^= 1u;
                const int gpr = tw >> 2;
                for (int item = tid; item < th * gpr; item += kThreads) {
                    const int ey = item / gpr, g = item - ey * gpr;
                    std::uint32_t v[4];
                    load_packed4_shared(stage + ey * L.stage_pitch + lead + 12 * g, v);
                    E* const trow = tile + ey * tp + 4 * g;
#pragma unroll
                    for (int j = 0; j < 4; ++j) trow[j] = v[j];
                }
            }
        }
        if (interior && !TMA) {
            const int gpr = tw >> 2;
            for (int item = tid; item < th * gpr; item += kThreads) {
                const int ey = item / gpr, g = item - ey * gpr;
                const long long o = row_offset(ey) + (X0 - radius + 4 * g) * 3;
                std::uint32_t v[4];
                load_packed4(a.src + o, v);
                E* const trow = tile + ey * tp + 4 * g;
                if constexpr (MODE == kJBF) {
                    std::uint32_t q[4];
                    load_packed4(a.guide + o, q);
#pragma unroll
                    for (int j = 0; j < 4; ++j) trow[j] = make_uint2(q[j], v[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) trow[j] = v[j];
                }
            }
        }
        {
            const int rem = tw - vec_cols;  // columns [vec_cols, tw) of every row
            for (int item = tid; item < th * rem; item += kThreads) {
                const int ey = item / rem, ex = vec_cols + (item - ey * rem);
                const int gx = clampi(X0 - radius + ex, 0, W - 1);
                const long long o = row_offset(ey) + gx * 3;
                if constexpr (MODE == kJBF) {
                    tile[ey * tp + ex] = make_uint2(load_packed(a.guide + o), load_packed(a.src + o));
                } else {
                    tile[ey * tp + ex] = load_packed(a.src + o);
                }
            }
        }
        group_sync();

        if constexpr (MODE == kABF) {
            // horizontal k-sums of every tile row for the 8P output columns: hs[row][col] = {B | R<<16, G}
            const int groups = kWarps * P / 8;
            for (int item = tid; item < th * groups; item += kThreads) {
                const int row = item % th, col0 = (item / th) * 8;
                const std::uint32_t* tr = reinterpret_cast<const std::uint32_t*>(tile) + row * tp + col0;
                std::uint32_t br = 0, g = 0;
                for (int j = 0; j < k; ++j) {
                    const std::uint32_t p = tr[j];
                    br += p & 0x00ff00ffu;
                    g += (p >> 8) & 0xffu;
                }
                uint2* hr = hs + row * L.hs_pitch + col0;
                hr[0] = make_uint2(br, g);
#pragma unroll
                for (int j = 1; j < 8; ++j) {
                    const std::uint32_t pn = tr[j + k - 1], po = tr[j - 1];
                    br += (pn & 0x00ff00ffu) - (po & 0x00ff00ffu);
                    g += ((pn >> 8) & 0xffu) - ((po >> 8) & 0xffu);
                    hr[j] = make_uint2(br, g);
                }
            }
            group_sync();
            // vertical k-sums of the horizontal sums: a thread owns one output column and kTileRows / nseg consecutive
            // rows; the window slides down (add the entering row, subtract the leaving one) -- all exact integers
            constexpr int ncols = kWarps * P, nseg = kThreads / ncols, rows_per = kTileRows / nseg;
            static_assert(nseg * ncols == kThreads && rows_per * nseg == kTileRows, "column / segment split");
            {
                const int col = tid % ncols, y0 = (tid / ncols) * rows_per;
                const uint2* hc = hs + y0 * L.hs_pitch + col;
                std::uint32_t sb = 0, sg = 0, sr = 0;
                for (int j = 0; j < k; ++j) {
                    const uint2 hv = hc[j * L.hs_pitch];
                    sb += hv.x & 0xffffu;
                    sr += hv.x >> 16;
                    sg += hv.y;
                }
                bs[y0 * L.hs_pitch + col] = make_uint4(sb, sg, sr, 0u);
#pragma unroll
                for (int y = 1; y < rows_per; ++y) {
                    const uint2 hn = hc[(y + k - 1) * L.hs_pitch], ho = hc[(y - 1) * L.hs_pitch];
                    sb += (hn.x & 0xffffu) - (ho.x & 0xffffu);
                    sr += (hn.x >> 16) - (ho.x >> 16);
                    sg += hn.y - ho.y;
                    bs[(y0 + y) * L.hs_pitch + col] = make_uint4(sb, sg, sr, 0u);
                }
            }
            group_sync();
        }

        Acc<MODE, P> A;
        const E* const tcentre = tile + (lane + radius) * tp + (warp * P + radius);
        if constexpr (MODE == kABF) {
            float nfp[3][P], noff[3][P];
#pragma unroll
            for (int i = 0; i < P; ++i) {
                const std::uint32_t c = tcentre[i];
                const uint4 box = bs[lane * L.hs_pitch + warp * P + i];
                const std::uint32_t sb = box.x, sg = box.y, sr = box.z;
                const float kk = static_cast<float>(k * k);
                const float c0 = static_cast<float>(c & 0xffu), c1 = static_cast<float>((c >> 8) & 0xffu),
                            c2 = static_cast<float>((c >> 16) & 0xffu);
                nfp[0][i] = -c0;
                nfp[1][i] = -c1;
                nfp[2][i] = -c2;
                // src/adaptive_bilateral_filter_impl.cu:88-93: centre - sum / (ksize*ksize); kept negated (x - y == x + (-y))
                noff[0][i] = -__fsub_rn(c0, __fdiv_rn(static_cast<float>(sb), kk));
                noff[1][i] = -__fsub_rn(c1, __fdiv_rn(static_cast<float>(sg), kk));
                noff[2][i] = -__fsub_rn(c2, __fdiv_rn(static_cast<float>(sr), kk));
            }
#pragma unroll
            for (int p2 = 0; p2 < P / 2; ++p2) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    A.nfp[c][p2] = pack2(nfp[c][2 * p2], nfp[c][2 * p2 + 1]);
                    A.noff[c][p2] = pack2(noff[c][2 * p2], noff[c][2 * p2 + 1]);
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) A.s[c][p2] = 0ull;
            }
        } else {
#pragma unroll
            for (int i = 0; i < P; ++i) {
                A.s01[i] = 0ull;
                if constexpr (MODE == kJBF) {
                    A.c[i] = tcentre[i].x;
                } else {
                    A.c[i] = tcentre[i];
                }
            }
#pragma unroll
            for (int p2 = 0; p2 < P / 2; ++p2) {
                A.s2[p2] = 0ull;
                A.sk[p2] = 0ull;
            }
        }

#pragma unroll 1
        for (int ky = -radius; ky <= radius; ++ky) {
            const int hw = hw_s[ky + radius];
            do_row<MODE, P, LK>(2 * hw + 1, tcentre + ky * tp - hw, w_s + (ky + radius) * L.w_pitch, lut_addr, clamp_i,
                                   clamp_f, A);
        }

        // epilogue: src/bilateral_filter_impl.cu:93-95
        const int y = Y0 + lane;
        const int x0 = X0 + warp * P;
        if (y < a.out_y0 + a.out_rows && x0 < W) {
            std::uint8_t o[3 * P];
#pragma unroll
            for (int i = 0; i < P; ++i) {
                float s0, s1, s2, sk;
                if constexpr (MODE == kABF) {
                    float x[4], y[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) unpack2(A.s[c][i / 2], x[c], y[c]);
                    s0 = (i & 1) ? y[0] : x[0], s1 = (i & 1) ? y[1] : x[1], s2 = (i & 1) ? y[2] : x[2],
                    sk = (i & 1) ? y[3] : x[3];
                } else {
                    float x2, y2, xk, yk;
                    unpack2(A.s01[i], s0, s1);
                    unpack2(A.s2[i / 2], x2, y2);
                    unpack2(A.sk[i / 2], xk, yk);
                    s2 = (i & 1) ? y2 : x2;
                    sk = (i & 1) ? yk : xk;
                }
                o[3 * i + 0] = static_cast<std::uint8_t>(__fadd_rn(__fdiv_rn(s0, sk), 0.5f));
                o[3 * i + 1] = static_cast<std::uint8_t>(__fadd_rn(__fdiv_rn(s1, sk), 0.5f));
                o[3 * i + 2] = static_cast<std::uint8_t>(__fadd_rn(__fdiv_rn(s2, sk), 0.5f));
            }
            std::uint8_t* const drow = a.dst + (static_cast<size_t>(y - a.dst_y0) * W + x0) * 3;
            if (x0 + P <= W && (reinterpret_cast<std::uintptr_t>(drow) & 3u) == 0) {
#pragma unroll
                for (int j = 0; j < 3 * P / 4; ++j) {
                    const std::uint32_t wv = o[4 * j] | (o[4 * j + 1] << 8) | (o[4 * j + 2] << 16) |
                                             (static_cast<std::uint32_t>(o[4 * j + 3]) << 24);
                    reinterpret_cast<std::uint32_t*>(drow)[j] = wv;
                }
            } else {
#pragma unroll
                for (int i = 0; i < P; ++i) {
                    if (x0 + i < W) {
                        drow[3 * i + 0] = o[3 * i + 0];
                        drow[3 * i + 1] = o[3 * i + 1];
                        drow[3 * i + 2] = o[3 * i + 2];
                    }
                }
            }
        }
    }
}

// Any radius, any alignment: one thread per output pixel, straight from global memory.  Same arithmetic
// and order as the tiled kernels; only used beyond kMaxTiledRadius.
template <int MODE>
__global__ void stencil_generic_kernel(const StencilArgs a, const int lut_len) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = a.out_y0 + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= a.width || y >= a.out_y0 + a.out_rows) return;
    const int W = a.width, H = a.height, r = a.radius, k = 2 * r + 1;
    const std::uint8_t* range = (MODE == kJBF) ? a.guide : a.src;
    // byte offset (relative to a.src / a.guide) of clamped image row yc; peer halo rows as in the tiled kernels
    auto row_off = [&](const int yc) -> long long {
        const long long pitch = static_cast<long long>(W) * 3;
        if (MODE != kJBF) {
            if (yc < a.src_y0 && a.above)
                return (a.above - a.src) + (clampi(yc, a.above_y0, a.above_y0 + a.above_rows - 1) - a.above_y0) * pitch;
            if (yc >= a.src_y0 + a.src_rows && a.below)
                return (a.below - a.src) + (clampi(yc, a.below_y0, a.below_y0 + a.below_rows - 1) - a.below_y0) * pitch;
        }
        return (clampi(yc, a.src_y0, a.src_y0 + a.src_rows - 1) - a.src_y0) * pitch;
    };
    const std::uint8_t* cp = range + (static_cast<size_t>(y - a.src_y0) * W + x) * 3;
    const int c0 = cp[0], c1 = cp[1], c2 = cp[2];
    float off0 = 0.f, off1 = 0.f, off2 = 0.f;
    if (MODE == kABF) {
        int b0 = 0, b1 = 0, b2 = 0;
        for (int ky = -r; ky <= r; ++ky) {
            const long long yo = row_off(clampi(y + ky, 0, H - 1));
            for (int kx = -r; kx <= r; ++kx) {
                const std::uint8_t* p = a.src + yo + clampi(x + kx, 0, W - 1) * 3;
                b0 += p[0];
                b1 += p[1];
                b2 += p[2];
            }
        }
        const float kk = static_cast<float>(k * k);
        off0 = __fsub_rn(static_cast<float>(c0), __fdiv_rn(static_cast<float>(b0), kk));
        off1 = __fsub_rn(static_cast<float>(c1), __fdiv_rn(static_cast<float>(b1), kk));
        off2 = __fsub_rn(static_cast<float>(c2), __fdiv_rn(static_cast<float>(b2), kk));
    }
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, sk = 0.f;
    for (int ky = -r; ky <= r; ++ky) {
        const long long yo = row_off(clampi(y + ky, 0, H - 1));
        for (int kx = -r; kx <= r; ++kx) {
            const float ks = __ldg(a.ks_full + (ky + r) * k + (kx + r));
            if (ks == 0.f) continue;  // masked corner: the reference adds +0.0f
            const long long o = yo + clampi(x + kx, 0, W - 1) * 3;
            const std::uint8_t* g = range + o;
            const std::uint8_t* p = a.src + o;
            int idx;
            if (MODE == kABF) {
                const float d0 = __fsub_rn(static_cast<float>(static_cast<int>(g[0]) - c0), off0);
                const float d1 = __fsub_rn(static_cast<float>(static_cast<int>(g[1]) - c1), off1);
                const float d2 = __fsub_rn(static_cast<float>(static_cast<int>(g[2]) - c2), off2);
                idx = __float2int_rz(__fadd_rn(__fadd_rn(fabsf(d0), fabsf(d1)), fabsf(d2)));
            } else {
                idx = abs(c0 - g[0]) + abs(c1 - g[1]) + abs(c2 - g[2]);
            }
            idx = min(idx, lut_len - 1);
            const float wt = __fmul_rn(ks, __ldg(a.lut + idx));
            s0 = __fmaf_rn(static_cast<float>(p[0]), wt, s0);
            s1 = __fmaf_rn(static_cast<float>(p[1]), wt, s1);
            s2 = __fmaf_rn(static_cast<float>(p[2]), wt, s2);
            sk = __fadd_rn(sk, wt);
        }
    }
    std::uint8_t* d = a.dst + (static_cast<size_t>(y - a.dst_y0) * W + x) * 3;
    d[0] = static_cast<std::uint8_t>(__fadd_rn(__fdiv_rn(s0, sk), 0.5f));
    d[1] = static_cast<std::uint8_t>(__fadd_rn(__fdiv_rn(s1, sk), 0.5f));
    d[2] = static_cast<std::uint8_t>(__fadd_rn(__fdiv_rn(s2, sk), 0.5f));
}

// outputs per thread: ABF keeps 6 more floats per output alive (centre, offset), so it runs 4 wide to stay within the
// 80 registers that 24 warps per SM allow
constexpr int outputs_per_thread(int mode) { return mode == kABF ? 4 : 8; }

// Warp groups per CTA that fit the shared-memory budget with `lut_n` table entries of kind `lut_kind` (0: none).
int groups_that_fit(int mode, int radius, int lut_n, int lut_kind) {
    for (int g = max_groups(mode); g >= 1; --g)
        if (make_layout(mode, outputs_per_thread(mode), radius, lut_n, lut_kind, g).total <= kSmemBudget) return g;
    return 0;
}

template <int MODE, int LK, bool PEER = false>
cudaError_t launch_tiled(const StencilArgs& a, int sm_count, cudaStream_t stream) {
    constexpr int P = outputs_per_thread(MODE);
    if constexpr (!PEER && MODE != kJBF) {
        if (a.above || a.below) return launch_tiled<MODE, LK, true>(a, sm_count, stream);
    }
    const int n_groups = groups_that_fit(MODE, a.radius, a.lut_n * a.lut_rows, LK);
    if (n_groups < 1) return cudaErrorInvalidConfiguration;  // stencil_pick_lut never hands out such a table
    SmemLayout L = make_layout(MODE, P, a.radius, a.lut_n * a.lut_rows, LK, n_groups);
    auto kern = stencil_kernel<MODE, P, LK, PEER>;
    if constexpr (MODE == kBF && !PEER && LK != kLutPremul) {
        // TMA tile staging (opt-in): image rows 16-byte aligned and the staging rows fit next to the same warp groups
        static const bool tma_on = [] {
            const char* v = std::getenv("VIP_STENCIL_TMA");
            return v && v[0] == '1';
        }();
        const SmemLayout Lt = make_layout(MODE, P, a.radius, a.lut_n * a.lut_rows, LK, n_groups, true);
        if (tma_on && (reinterpret_cast<std::uintptr_t>(a.src) & 15u) == 0 && (a.width * 3) % 16 == 0 && Lt.total <= kSmemBudget) {
            L = Lt;
            kern = stencil_kernel<MODE, P, LK, false, true>;
        }
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total);
    if (e != cudaSuccess) return e;
    const int tiles_x = (a.width + kWarps * P - 1) / (kWarps * P);
    const int tiles_y = (a.out_rows + kTileRows - 1) / kTileRows;
    const int n_tiles = tiles_x * tiles_y;
    // one persistent CTA per SM
    int grid = sm_count;
    if (grid > n_tiles) grid = n_tiles;
    e = launch_pdl(kern, dim3(grid), dim3(kThreads * n_groups), L.total, stream, a, tiles_x, n_tiles);
    count_launch();
    return e;
}

template <int MODE>
cudaError_t launch_tiled_kind(const StencilArgs& a, int sm_count, cudaStream_t stream) {
    switch (a.lut_kind) {
        case kLutFull: return launch_tiled<MODE, kLutFull>(a, sm_count, stream);
        case kLutClamp: return launch_tiled<MODE, kLutClamp>(a, sm_count, stream);
        case kLutPremul:
            if constexpr (MODE != kABF) return launch_tiled<MODE, kLutPremul>(a, sm_count, stream);
            return cudaErrorInvalidConfiguration;   // stencil_pick_lut never picks it for ABF
        default: return launch_tiled<MODE, kLutNarrow>(a, sm_count, stream);
    }
}

template <int MODE>
cudaError_t launch_generic(const StencilArgs& a, cudaStream_t stream) {
    const dim3 block(32, 8);
    const dim3 grid((a.width + 31) / 32, (a.out_rows + 7) / 8);
    stencil_generic_kernel<MODE><<<grid, block, 0, stream>>>(a, MODE == kABF ? 1536 : 768);
    count_launch();
    return cudaGetLastError();
}

}  // namespace

int stencil_max_tiled_radius() { return kMaxTiledRadius; }

int stencil_w_pitch(int radius) { return (2 * radius + 1 + 7) & ~7; }

void stencil_describe(StencilMode mode, int radius, int lut_n, int lut_rows, int lut_kind, int* copies, int* clamped,
                      int* groups, int* smem_bytes) {
    lut_n *= lut_rows;
    const int g = groups_that_fit(mode, radius, lut_n, lut_kind);
    *copies = 1 << (lut_shift(lut_kind) - 2);
    *clamped = lut_clamped(lut_kind) ? 1 : 0;
    *groups = g;
    *smem_bytes = g > 0 ? make_layout(mode, outputs_per_thread(mode), radius, lut_n, lut_kind, g).total : 0;
}

void stencil_pick_lut(StencilMode mode, int radius, int table_len, int last_nz, int n_ks, int* lut_n, int* lut_kind) {
    // The table is shared by the warp groups of the one CTA an SM runs.  Candidates, in order of preference at an equal
    // number of warp groups: the whole table 32x replicated (no clamp instruction); the table cut after its last
    // non-zero entry -- every index > last_nz has weight exactly 0, so clamping the index to last_nz+1 (a zero entry)
    // is bit-neutral; the cut table 8x replicated.  More resident warp groups win over a conflict-free table: a single
    // 256-thread group per SM cannot hide the LDS / I2F latencies of the tap loop.
    *lut_n = table_len;
    *lut_kind = kLutFull;
    if (radius > kMaxTiledRadius) return;
    const int cut_n = last_nz + 2 < table_len ? last_nz + 2 : table_len;
    const int g_full = groups_that_fit(mode, radius, table_len, kLutFull);
    const int g_cut = groups_that_fit(mode, radius, cut_n, kLutClamp);
    const int g_narrow = groups_that_fit(mode, radius, cut_n, kLutNarrow);
    int best = g_full;
    if (g_cut > best) {
        best = g_cut;
        *lut_n = cut_n;
        *lut_kind = kLutClamp;
    }
    if (g_narrow > best) {
        best = g_narrow;
        *lut_n = cut_n;
        *lut_kind = kLutNarrow;
    }
    // One premultiplied row per distinct spatial weight (n_ks of them) removes the FMUL of every tap (8.5 -> 7.7
    // instructions per tap for the texture filter's 17 x 17 joint bilateral step, bit-exact).  Measured on B200
    // (profiles/r02b_premul_ab.txt): 0.559 vs 0.562 ms -- nothing, for the reason given at do_row -- so it is OFF unless
    // VIP_STENCIL_PREMUL=1 asks for it; it then is taken whenever it keeps as many warp groups as the best plain layout.
    static const bool premul_on = [] {
        const char* v = std::getenv("VIP_STENCIL_PREMUL");
        return v && v[0] == '1';
    }();
    if (premul_on && mode != kABF && n_ks > 0 && static_cast<long long>(cut_n) * n_ks <= 16384 &&
        groups_that_fit(mode, radius, cut_n * n_ks, kLutPremul) >= best) {
        *lut_n = cut_n;
        *lut_kind = kLutPremul;
    }
}

cudaError_t launch_stencil(StencilMode mode, const StencilArgs& a, int sm_count, cudaStream_t stream) {
    if (a.out_rows <= 0 || a.width <= 0) return cudaSuccess;
    if (a.radius > kMaxTiledRadius) {
        switch (mode) {
            case kBF: return launch_generic<kBF>(a, stream);
            case kJBF: return launch_generic<kJBF>(a, stream);
            default: return launch_generic<kABF>(a, stream);
        }
    }
    switch (mode) {
        case kBF: return launch_tiled_kind<kBF>(a, sm_count, stream);
        case kJBF: return launch_tiled_kind<kJBF>(a, sm_count, stream);
        default: return launch_tiled_kind<kABF>(a, sm_count, stream);
    }
}

}  // namespace vip
