// Microbenchmark: tcgen05.mma issue/execution rate on B200 for the operand layouts the decoder kernel could use.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/ubench_mma profiles/ubench_mma.cu && /tmp/ubench_mma
// One CTA per SM; one thread issues `iters` MMAs (M=128, K=16, fp16 -> fp32 TMEM accumulator) back to back from fixed
// shared-memory operands, commits, waits; prints clocks per MMA.  Variants: N, swizzle mode of the smem descriptors,
// A operand from tensor memory, and a second group of warps hammering shared memory with 16-byte stores (the epilogue's
// operand writes) at the same time.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include "../gaussian_gan_decoder_b200/csrc/tc_ptx.cuh"
using namespace ggd;

__device__ __forceinline__ uint64_t desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t mode)
{
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) |
           (1ull << 46) | ((uint64_t)mode << 61);
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc)
{
    asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p; }"
                 ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

__global__ void __launch_bounds__(256, 1) k(int N, int mode, int a_tmem, int hammer, int iters, int ksteps, long long* out)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::mbar_fence_init(); }
    if (warp == 0) ptx::tmem_alloc<512>(&slot);
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    const uint32_t sa = ptx::smem_u32(smem), sb = sa + 64 * 1024;
    if (threadIdx.x == 0) {
        const uint32_t idesc = ptx::idesc_f16(128, N);
        const uint32_t lbo = mode == 0 ? 128 : 16, sbo = mode == 0 ? 16 * 16 * ksteps : 1024;
        const uint32_t step = mode == 0 ? 256 : 32;
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it)
            for (int s = 0; s < ksteps; ++s) {
                if (a_tmem) mma_ts(tm, tm + 256 + 8 * s, desc(sb + s * step, lbo, sbo, mode), idesc, (it | s) ? 1u : 0u);
                else ptx::mma_f16_ss(tm, desc(sa + s * step, lbo, sbo, mode), desc(sb + s * step, lbo, sbo, mode), idesc, (it | s) ? 1u : 0u);
            }
        const long long t1 = clock64();
        ptx::mma_commit(&bar);
        ptx::mbar_wait(&bar, 0);
        const long long t2 = clock64();
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    } else if (hammer && warp >= 4) {
        // 4 warps writing 16 bytes per thread into a 36 KB region, like the hidden-layer epilogue does
        uint4 v = make_uint4(threadIdx.x, 1, 2, 3);
        uint8_t* base = smem + 128 * 1024;
        for (int it = 0; it < iters * ksteps * hammer; ++it) {
            *reinterpret_cast<uint4*>(base + ((threadIdx.x - 128) * 16 + (it & 15) * 2048)) = v;
            v.x += it;
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tm);
}

// ---- issue-path variants (N = 128, no swizzle): how fast can ONE warp feed the tensor core? ----
__device__ __forceinline__ void mma_elect(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc)
{   // executed by the whole (converged) warp: operands stay warp-uniform, one elected lane issues
    asm volatile("{ .reg .pred p, e; setp.ne.b32 p, %4, 0; elect.sync _|e, 0xffffffff;\n"
                 "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p; }"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
template <int V>
__global__ void __launch_bounds__(576, 1) k2(int iters, long long* out, int hammer_warps, int hammer_gap, int N = 128)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::mbar_fence_init(); }
    if (warp == 0) ptx::tmem_alloc<512>(&slot);
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    const uint32_t sa = ptx::smem_u32(smem), sb = sa + 64 * 1024;
    const uint32_t idesc = ptx::idesc_f16(128, N);
    const uint64_t a0 = desc(sa, 128, 16 * 16 * 9, 0), b0 = desc(sb, 128, 16 * 16 * 9, 0);
    if (V == 0 && threadIdx.x == 0) {            // one thread, descriptors advanced by adding to the low word, unrolled
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int s = 0; s < 9; ++s) ptx::mma_f16_ss(tm, a0 + 16 * s, b0 + 16 * s, idesc, (it | s) ? 1u : 0u);
        }
        const long long t1 = clock64();
        ptx::mma_commit(&bar);
        ptx::mbar_wait(&bar, 0);
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = clock64() - t0; }
    }
    if (warp >= 2 && warp < 2 + hammer_warps) {  // epilogue-like shared-memory stores next to the MMAs: 16 B per thread,
        uint4 v = make_uint4(threadIdx.x, 1, 2, 3);   // conflict-free, `hammer_gap` dependent FMAs between two stores
        uint8_t* base = smem + 128 * 1024 + (warp - 2) * 2048;
        float f = threadIdx.x;
        for (int it = 0; it < iters * 9 * 4; ++it) {
            *reinterpret_cast<uint4*>(base + (threadIdx.x & 31) * 16 + (it & 3) * 512) = v;
            if (hammer_gap >= 0) { for (int g = 0; g < hammer_gap; ++g) f = f * 1.0001f + 0.5f; }
            else {   // issue-bound arithmetic: 8 independent half2 chains, like the GELU epilogue
                __half2 h0 = __float2half2_rn(f), h1 = h0, h2 = h0, h3 = h0, h4 = h0, h5 = h0, h6 = h0, h7 = h0;
                const __half2 c = __float2half2_rn(0.999f);
                for (int g = 0; g < -hammer_gap; ++g) {
                    h0 = __hfma2(h0, c, h1); h1 = __hfma2(h1, c, h2); h2 = __hfma2(h2, c, h3); h3 = __hfma2(h3, c, h4);
                    h4 = __hfma2(h4, c, h5); h5 = __hfma2(h5, c, h6); h6 = __hfma2(h6, c, h7); h7 = __hfma2(h7, c, h0);
                }
                f += __low2float(h0) + __low2float(h3) + __low2float(h5) + __low2float(h7);
            }
            v.x = __float_as_uint(f);
        }
        if (f == 12345.f) out[3] = 1;
    }
    if (V == 1 && warp == 0) {                   // whole warp, uniform operands, elected issuer
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int s = 0; s < 9; ++s) mma_elect(tm, a0 + 16 * s, b0 + 16 * s, idesc, (it | s) ? 1u : 0u);
        }
        const long long t1 = clock64();
        if (threadIdx.x == 0) {
            ptx::mma_commit(&bar);
            ptx::mbar_wait(&bar, 0);
            if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = clock64() - t0; }
        }
        __syncwarp();
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tm);
}

// ---- one tile-layer at a time: 9 MMAs, commit, wait for completion — what a dependent step of the decoder costs ----
template <int kSteps>
__global__ void __launch_bounds__(640, 1) k3(int iters, long long* out, int N, int ldtm_warps = 0, int f16acc = 0)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::mbar_fence_init(); }
    if (warp == 0) ptx::tmem_alloc<512>(&slot);
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    const uint32_t sa = ptx::smem_u32(smem), sb = sa + 64 * 1024;
    const uint32_t idesc = f16acc ? (ptx::idesc_f16(128, N) & ~(3u << 4)) : ptx::idesc_f16(128, N);   // bits [4,6): 0 = f16 accumulator
    const uint64_t a0 = desc(sa, 128, 16 * 16 * 9, 0), b0 = desc(sb, 128, 16 * 16 * 9, 0);
    if (threadIdx.x == 0) {
        long long t_issue = 0, t_total = 0;
        for (int it = 0; it < iters; ++it) {
            const long long t0 = clock64();
#pragma unroll
            for (int s = 0; s < kSteps; ++s) ptx::mma_f16_ss(tm, a0 + 16 * s, b0 + 16 * s, idesc, s ? 1u : 0u);
            ptx::mma_commit(&bar);
            const long long t1 = clock64();
            ptx::mbar_wait(&bar, it & 1);
            const long long t2 = clock64();
            t_issue += t1 - t0; t_total += t2 - t0;
        }
        if (blockIdx.x == 0) { out[0] = t_issue; out[1] = t_total; }
    } else if (warp >= 4 && warp < 4 + ldtm_warps) {
        // epilogue-like tensor-memory reads of ANOTHER accumulator (columns 256..383) while the MMAs write columns 0..127
        const uint32_t trow = tm + ((uint32_t)((warp & 3) * 32) << 16) + 256 + ((warp >> 2) & 3) * 32;
        uint32_t acc = 0;
        for (int it = 0; it < iters * 6; ++it) {
            if (f16acc) {     // 32 columns of 16-bit accumulators packed two per register
                uint32_t v[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                               "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(trow) : "memory");
                ptx::tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 16; ++q) acc += v[q];
            } else {
                uint32_t v[32];
                ptx::tmem_ld32(trow, v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 32; ++q) acc += v[q];
            }
        }
        if (acc == 0x12345u) out[3] = acc;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tm);
}

int main()
{
    {
        long long* d; cudaMalloc(&d, 64);
        long long h[2];
        cudaFuncSetAttribute(k3<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(k3<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        for (int n : {128, 16}) {
            k3<9><<<148, 256, 200 * 1024>>>(400, d, n);
            if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k3 failed\n"); return 1; }
            printf("burst of 9 MMAs N=%3d + commit + wait: issue %.0f clk, until complete %.0f clk per burst\n", n, h[0] / 400.0, h[1] / 400.0);
        }
        for (int lw : {4, 16}) {
            k3<9><<<148, 640, 200 * 1024>>>(400, d, 128, lw);
            if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k3 ldtm failed\n"); return 1; }
            printf("burst of 9 MMAs N=128 with %2d warps reading tensor memory: issue %.0f clk, until complete %.0f clk per burst\n", lw, h[0] / 400.0, h[1] / 400.0);
        }
        for (int lw : {0, 16}) {
            k3<9><<<148, 640, 200 * 1024>>>(400, d, 128, lw, 1);
            if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k3 f16 failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
            printf("burst of 9 MMAs N=128, f16 accumulator, %2d warps reading it packed: issue %.0f clk, until complete %.0f clk per burst\n", lw, h[0] / 400.0, h[1] / 400.0);
        }
        k3<3><<<148, 256, 200 * 1024>>>(400, d, 128);
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("burst of 3 MMAs N=128 + commit + wait: issue %.0f clk, until complete %.0f clk per burst\n", h[0] / 400.0, h[1] / 400.0);
    }
    {
        long long* d; cudaMalloc(&d, 64);
        long long h[2];
        cudaFuncSetAttribute(k2<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(k2<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        for (int hw : {4, 8, 16})
            for (int gap : {0, 8, 32}) {
                k2<0><<<148, 576, 200 * 1024>>>(200, d, hw, gap);
                if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k2 hammer failed\n"); return 1; }
                printf("one thread unrolled + %2d warps storing (gap %2d FMAs) issue %.1f clk/MMA, complete %.1f clk/MMA\n", hw, gap, h[0] / 1800.0, h[1] / 1800.0);
            }
        for (int n : {16, 128})
            for (int hw : {0, 16})
                for (int gap : {-8}) {
                    k2<0><<<148, 576, 200 * 1024>>>(200, d, hw, gap, n);
                    if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k2 hfma failed\n"); return 1; }
                    printf("N=%3d one thread unrolled + %2d warps of HFMA2 streams: issue %.1f clk/MMA, complete %.1f clk/MMA\n", n, hw, h[0] / 1800.0, h[1] / 1800.0);
                }
        k2<0><<<148, 256, 200 * 1024>>>(200, d, 0, 0);
        if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k2<0> failed\n"); return 1; }
        printf("%-48s issue %.1f clk/MMA, complete %.1f clk/MMA\n", "one thread, unrolled, desc += 16", h[0] / 1800.0, h[1] / 1800.0);
        for (int n : {16, 128})
            for (int hw : {0, 16}) {
                k2<1><<<148, 576, 200 * 1024>>>(200, d, hw, -8, n);
                if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k2<1> n failed\n"); return 1; }
                printf("N=%3d whole warp + elect + %2d warps of HFMA2 streams: issue %.1f clk/MMA, complete %.1f clk/MMA\n", n, hw, h[0] / 1800.0, h[1] / 1800.0);
            }
        k2<1><<<148, 256, 200 * 1024>>>(200, d, 0, 0);
        if (cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("k2<1> failed\n"); return 1; }
        printf("%-48s issue %.1f clk/MMA, complete %.1f clk/MMA\n", "whole warp + elect.sync, unrolled", h[0] / 1800.0, h[1] / 1800.0);
    }
    long long* d; cudaMalloc(&d, 16);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    struct { int N, mode, a_tmem, hammer; const char* name; } cases[] = {
        {128, 0, 0, 0, "N=128 no-swizzle SS (decoder today)"}, {128, 2, 0, 0, "N=128 swizzle-128B SS"},
        {256, 0, 0, 0, "N=256 no-swizzle SS"}, {256, 2, 0, 0, "N=256 swizzle-128B SS"},
        {128, 0, 1, 0, "N=128 no-swizzle, A in TMEM"}, {128, 2, 1, 0, "N=128 swizzle-128B, A in TMEM"},
        {128, 0, 0, 4, "N=128 no-swizzle SS + shared-memory stores"}, {128, 2, 0, 4, "N=128 swizzle-128B SS + stores"},
        {16, 0, 0, 0, "N=16 no-swizzle SS (last layer)"}, {64, 0, 0, 0, "N=64 no-swizzle SS"},
    };
    for (auto& c : cases)
        for (int ksteps : {9, 4}) {
            const int iters = 200;
            k<<<148, 256, 200 * 1024>>>(c.N, c.mode, c.a_tmem, c.hammer, iters, ksteps, d);
            long long h[2];
            cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { printf("%s: %s\n", c.name, cudaGetErrorString(e)); return 1; }
            printf("%-48s ksteps %d: issue %.1f clk/MMA, complete %.1f clk/MMA\n", c.name, ksteps,
                   (double)h[0] / (iters * ksteps), (double)h[1] / (iters * ksteps));
        }
    return 0;
}
