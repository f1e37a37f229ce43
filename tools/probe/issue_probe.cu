// Measured "relevant roofline" of the bilateral stencils ("measure with an FFMA micro-kernel on the box").
// The stencils are bound by instruction issue, not by HBM or tensor cores, so their roof is the rate at which an SM
// can issue the instruction kinds of the tap loop.  Each probe is a persistent grid (one 1024-thread CTA per SM, 8
// independent dependency chains per thread) that issues ONE instruction kind back to back:
//   0  FFMA            scalar fp32 fused multiply-add                      -> thread-instructions / s
//   1  FFMA2           packed fma.rn.f32x2 (two fp32 FMAs per instruction)   -> packed instructions / s (x2 = FMAs)
//   2  FFMA + VABSDIFF4  alternating fp32 (FMA pipe) and the packed-byte SAD of the range index (ALU pipe): do they overlap?
//   3  LDS.32          conflict-free shared-memory loads (lane l reads word l of a 32-word row)
//   4  I2F.U8          byte -> float conversion (the XU-pipe instruction of the tile unpack)
//   5  FFMA2 + VABSDIFF4 alternating: does a packed FMA block the dispatch port for the ALU pipe too, or only the FMA pipe?
//   6  FFMA2 + LDS.32    alternating: the same question for the LSU
//   7  FFMA + LOP3       alternating (FMA pipe + a plain ALU instruction)
//   8  FFMA2 + FFMA      alternating (packed and scalar FMAs sharing the FMA pipe)
// Not part of the product library: built by __graft_entry__.build() into tools/probe/libvip_probe.so and called by
// bench.py (key "issue.measured_peak") and tools/peaks.py through ctypes.
#include <cuda_runtime.h>

#include <cstdint>

namespace {

constexpr int kThreads = 1024;
constexpr int kChains = 8;
constexpr int kUnroll = 16;  // instructions per chain per loop trip

template <int KIND>
__global__ void __launch_bounds__(kThreads, 1) probe_kernel(float* __restrict__ sink, const int trips, const float seed) {
    __shared__ float lds_src[32 * 33];
    for (int i = threadIdx.x; i < 32 * 33; i += kThreads) lds_src[i] = seed * i;
    __syncthreads();
    const std::uint32_t lane_addr =
        static_cast<std::uint32_t>(__cvta_generic_to_shared(lds_src + (threadIdx.x & 31)));
    float a[kChains];
    unsigned long long p[kChains];
    std::uint32_t u[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
        a[c] = seed + c + threadIdx.x;
        asm("mov.b64 %0, {%1, %1};" : "=l"(p[c]) : "f"(a[c]));
        u[c] = (KIND == 3 || (KIND == 6 && !(c & 1))) ? lane_addr + (((threadIdx.x >> 5) * 5u + c) & 31u) * 132u : threadIdx.x * 2654435761u + c;
    }
    const float m = 0.999f + seed * 1e-9f, b = seed * 1e-3f;
    unsigned long long m2, b2;
    asm("mov.b64 %0, {%1, %1};" : "=l"(m2) : "f"(m));
    asm("mov.b64 %0, {%1, %1};" : "=l"(b2) : "f"(b));
    for (int t = 0; t < trips; ++t) {
#pragma unroll
        for (int k = 0; k < kUnroll; ++k) {
#pragma unroll
            for (int c = 0; c < kChains; ++c) {
                if constexpr (KIND == 0) {
                    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(m), "f"(b));
                } else if constexpr (KIND == 1) {
                    asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[c]) : "l"(m2), "l"(b2));
                } else if constexpr (KIND == 2) {
                    if (c & 1)
                        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(m), "f"(b));
                    else
                        asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %0;" : "+r"(u[c]) : "r"(u[(c + 2) % kChains]));
                } else if constexpr (KIND == 5) {
                    if (c & 1)
                        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[c]) : "l"(m2), "l"(b2));
                    else
                        asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %0;" : "+r"(u[c]) : "r"(u[(c + 2) % kChains]));
                } else if constexpr (KIND == 6) {
                    if (c & 1) {
                        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[c]) : "l"(m2), "l"(b2));
                    } else {
                        float v;
                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(u[c]));
                        u[c] += __float_as_uint(v) >> 31;
                    }
                } else if constexpr (KIND == 7) {
                    if (c & 1)
                        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(m), "f"(b));
                    else
                        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[c]) : "r"(u[(c + 2) % kChains]), "r"(0x5bd1e995u));
                } else if constexpr (KIND == 8) {
                    if (c & 1)
                        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[c]) : "f"(m), "f"(b));
                    else
                        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[c]) : "l"(m2), "l"(b2));
                } else if constexpr (KIND == 3) {
                    // the row is warp-uniform (pitch 33 words: lane l hits bank (row + l) % 32, conflict-free); the next
                    // address depends on the loaded value's sign bit (always 0), which keeps the chain honest
                    float v;
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(u[c]));
                    u[c] += __float_as_uint(v) >> 31;
                } else {
                    float v;
                    asm volatile("cvt.rn.f32.u8 %0, %1;" : "=f"(v) : "r"(u[c]));
                    u[c] = __float_as_uint(v);
                }
            }
        }
    }
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[c]));
        acc += a[c] + lo + hi + __uint_as_float(u[c] & 0x3fffffffu) + ((KIND == 3 || KIND == 6) ? lds_src[u[c] & 31u] : 0.f);
    }
    if (acc == 123.456f) sink[0] = acc;  // never true in practice: keeps the chains alive
}

template <int KIND>
cudaError_t run_kind(float* sink, int grid, int trips, cudaStream_t s) {
    probe_kernel<KIND><<<grid, kThreads, 0, s>>>(sink, trips, 1.25f);
    return cudaGetLastError();
}

}  // namespace

// Runs probe `kind` for `trips` loop trips; returns 0 and the measured rate in instructions per second (thread-level:
// warp instructions x 32) plus the elapsed milliseconds, or a cudaError_t.
extern "C" __attribute__((visibility("default"))) int vip_probe_issue_rate(int kind, int trips, double* instr_per_s,
                                                                           float* elapsed_ms) {
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    float* sink = nullptr;
    if ((e = cudaMalloc(&sink, 4)) != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    auto launch = [&](int n) -> cudaError_t {
        switch (kind) {
            case 0: return run_kind<0>(sink, sms, n, nullptr);
            case 1: return run_kind<1>(sink, sms, n, nullptr);
            case 2: return run_kind<2>(sink, sms, n, nullptr);
            case 3: return run_kind<3>(sink, sms, n, nullptr);
            case 4: return run_kind<4>(sink, sms, n, nullptr);
            case 5: return run_kind<5>(sink, sms, n, nullptr);
            case 6: return run_kind<6>(sink, sms, n, nullptr);
            case 7: return run_kind<7>(sink, sms, n, nullptr);
            case 8: return run_kind<8>(sink, sms, n, nullptr);
            default: return cudaErrorInvalidValue;
        }
    };
    e = launch(trips / 8 + 1);  // warm-up (clock ramp, instruction cache)
    if (e == cudaSuccess) {
        cudaEventRecord(e0, nullptr);
        e = launch(trips);
        cudaEventRecord(e1, nullptr);
    }
    if (e == cudaSuccess) e = cudaEventSynchronize(e1);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    if (e != cudaSuccess) return e;
    // KIND 3 / 4 (and the LDS half of 6) carry one dependent integer instruction per probe instruction; it is not counted
    const double n_instr = static_cast<double>(sms) * kThreads * static_cast<double>(trips) * kUnroll * kChains;
    if (instr_per_s) *instr_per_s = n_instr / (ms * 1e-3);
    if (elapsed_ms) *elapsed_ms = ms;
    return 0;
}
