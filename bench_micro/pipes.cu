// Micro-benchmarks of the sm_100a per-SM instruction rates the stencil kernels are bound by (issue
// roofline): warp-instructions / clk / SM for FFMA, FADD, packed FFMA2, F2I, I2F, VABSDIFF4, LDS.  Standalone:
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu && ./pipes
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

constexpr int kIters = 1024;
constexpr int kThreads = 1024;

#define CHAINS 8
#define REPS 8

template <int OP>
__global__ void __launch_bounds__(kThreads) k(float* out, long long* clocks, int flag) {
    __shared__ float sm[32 * 64];
    for (int i = threadIdx.x; i < 32 * 64; i += blockDim.x) sm[i] = (float)(i & 63);
    __syncthreads();
    float a[CHAINS], b[CHAINS], c2[CHAINS];
    unsigned u[CHAINS];
    for (int i = 0; i < CHAINS; ++i) {
        a[i] = threadIdx.x * 0.001f + i;
        b[i] = 1.0f + i * 1e-3f; c2[i] = 0.5f * i;
        u[i] = threadIdx.x * 2654435761u + i;
    }
    const float c = flag ? 1.000001f : 0.999999f;
    const float d = flag ? 0.5f : 0.25f;
    const unsigned lane = threadIdx.x & 31;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int rep = 0; rep < REPS; ++rep)
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) {
            if (OP == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));
            if (OP == 1) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(d));
            if (OP == 2) {
                unsigned long long p, q, r;
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p) : "f"(a[i]), "f"(b[i]));
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(q) : "f"(c), "f"(c));
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(d), "f"(d));
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p) : "l"(q), "l"(r));
                asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(b[i]) : "l"(p));
            }
            if (OP == 3) { int r; asm volatile("cvt.rzi.s32.f32 %0, %1;" : "=r"(r) : "f"(a[i])); u[i] += r; }
            if (OP == 4) { float r; asm volatile("cvt.rn.f32.u32 %0, %1;" : "=f"(r) : "r"(__float_as_uint(a[i]) & 0xffu)); a[i] += r; }
            if (OP == 5) asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %0;" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]));
            if (OP == 6) { a[i] += sm[((u[i] + it) & 63) * 32 + lane]; }
            if (OP == 7) asm volatile("add.rz.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(d));
            if (OP == 8) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d)); asm volatile("add.u32 %0, %0, %1;" : "+r"(u[i]) : "r"(u[(i+1)%CHAINS])); }
            if (OP == 9) { float r; asm volatile("{.reg .b32 t; prmt.b32 t, %1, 0x4b000000, 0x7440; mov.b32 %0, t;}" : "=f"(r) : "r"(__float_as_uint(a[i]))); a[i] += r; }
            if (OP == 14) { unsigned long long p, q, r;
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p) : "f"(a[i]), "f"(b[i]));
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(q) : "f"(c), "f"(c));
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(d), "f"(d));
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p) : "l"(q), "l"(r));
                asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(b[i]) : "l"(p));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]), "r"(lane)); }
            if (OP == 15) { unsigned long long p, q, r;
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p) : "f"(a[i]), "f"(b[i]));
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(q) : "f"(c), "f"(c));
                asm volatile("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(d), "f"(d));
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p) : "l"(q), "l"(r));
                asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(b[i]) : "l"(p));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(c2[i]) : "f"(c), "f"(d)); }
            if (OP == 16) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c), "f"(d));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(b[i]) : "f"(c), "f"(d));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]), "r"(lane)); }
            if (OP == 10) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c));
            if (OP == 11) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]), "r"(lane));
            if (OP == 12) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]), "r"(lane));
            if (OP == 13) asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b[i]));
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < CHAINS; ++i) s += a[i] + b[i] + u[i] + c2[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int per_iter_instr, int sms) {
    float* out;
    long long* clk;
    cudaMalloc(&out, sizeof(float) * sms * kThreads);
    cudaMalloc(&clk, sizeof(long long) * sms);
    k<OP><<<sms, kThreads>>>(out, clk, 1);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OP><<<sms, kThreads>>>(out, clk, 1);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[1024];
    cudaMemcpy(h, clk, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; ++i) mean += h[i];
    mean /= sms;
    const double winstr = (double)kIters * CHAINS * REPS * per_iter_instr * (kThreads / 32);
    printf("%-28s %8.3f warp-instr/clk/SM  (%.0f clk, %.3f ms, err=%d)\n", name, winstr / mean, mean, ms, (int)cudaGetLastError());
    cudaFree(out);
    cudaFree(clk);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs, clock %d kHz\n", p.name, sms, p.clockRate);
    run<0>("FFMA (3-reg)", 1, sms);
    run<1>("FADD", 1, sms);
    run<10>("FMUL", 1, sms);
    run<2>("FFMA2 (f32x2)", 1, sms);
    run<3>("F2I.TRUNC (+IADD)", 2, sms);
    run<4>("I2F.U32 (+LOP,+FADD)", 3, sms);
    run<9>("PRMT u8->f32 bits (+FADD)", 2, sms);
    run<5>("VABSDIFF4.ACC", 1, sms);
    run<6>("LDS gather replicated(+3)", 4, sms);
    run<7>("FADD.RZ", 1, sms);
    run<8>("FFMA + IADD pair", 2, sms);
    run<11>("LOP3", 1, sms);
    run<12>("IMAD", 1, sms);
    run<13>("FMNMX", 1, sms);
    run<14>("FFMA2 + LOP3", 2, sms);
    run<15>("FFMA2 + FFMA", 2, sms);
    run<16>("2 FFMA + LOP3", 3, sms);
    return 0;
}
