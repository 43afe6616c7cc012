// throughput of the epilogue's instruction kinds: warp-instructions per cycle per SMSP
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#define ITER 2048
template <int OP> __global__ void k(uint32_t* out, long long* cyc, uint32_t seed)
{
    uint32_t r[8];
    for (int i = 0; i < 8; ++i) r[i] = seed + threadIdx.x * 8 + i;
    const uint32_t c1 = 0x2c002c00u + seed, c2 = 0x30003000u + seed;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2));
            if (OP == 1) asm volatile("mul.rn.f16x2 %0, %0, %0;" : "+r"(r[i]));
            if (OP == 2) asm volatile("{ .reg .b16 a,b; mov.b32 {a,b}, %0; tanh.approx.f16 a, a; mov.b32 %0, {a,b}; }" : "+r"(r[i]));
            if (OP == 3) asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(r[i]));
            if (OP == 4) asm volatile("fma.rn.f16x2 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(0x2c002c00u));   // immediate operand
            if (OP == 5) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c1));
            if (OP == 6) { float a = __uint_as_float(r[i]), b = __uint_as_float(r[(i + 1) & 7]); asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(a), "f"(b)); }
            if (OP == 7) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2));
            if (OP == 8) asm volatile("tanh.approx.f32 %0, %0;" : "+r"(r[i]));
            if (OP == 9) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(r[i]));
            if (OP == 10) { asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[(i + 4) & 7]) : "r"(c1), "r"(c2)); }
            if (OP == 11) { asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2)); if (i < 2) asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(r[7 - i])); }
        }
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 8; ++i) s ^= r[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int per_iter)
{
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 148 * 8);
    k<OP><<<148, 512>>>(out, cyc, 1); k<OP><<<148, 512>>>(out, cyc, 1);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = (double)h[0];
    // 4 warps per SMSP, per_iter warp-instr per loop iteration each
    printf("%-34s %8.0f cycles  -> %.3f warp-instr/clk/SMSP  (%.2f clk per warp-instr)\n", name, c, 4.0 * per_iter * ITER / c, c / (4.0 * per_iter * ITER));
    cudaFree(out); cudaFree(cyc);
}
int main()
{
    run<0>("HFMA2 reg", 8); run<4>("HFMA2 imm", 8); run<1>("HMUL2", 8); run<5>("HMNMX2", 8); run<6>("F2FP pack", 8);
    run<7>("FFMA", 8); run<2>("MUFU.TANH.F16 (1 half)", 8); run<3>("tanh.f16x2 (2 MUFU + PRMT)", 8); run<8>("MUFU.TANH f32", 8);
    run<9>("MUFU.EX2 f32", 8); run<10>("HFMA2 + LOP3 pairs", 16); run<11>("8 HFMA2 + 2 tanh.f16x2", 10);
    return 0;
}
