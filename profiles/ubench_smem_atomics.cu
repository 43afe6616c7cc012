// Throughput of shared-memory atomics / scattered stores as the per-tile sorts use them:
// one CTA of 1024 threads per SM, every thread issues ITER operations on pseudo-random words of a table.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o ubench_smem_atomics ubench_smem_atomics.cu
#include <cstdio>
#include <cstdint>
#define ITER 4096
template <int OP, int WORDS> __global__ void __launch_bounds__(1024, 1) k(uint32_t* out, long long* cyc)
{
    extern __shared__ uint32_t s[];
    for (int i = threadIdx.x; i < WORDS; i += 1024) s[i] = 0;
    uint32_t x = threadIdx.x * 2654435761u + 12345u, acc = 0;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 4
    for (int it = 0; it < ITER; ++it) {
        x = x * 1664525u + 1013904223u;
        const uint32_t a = (x >> 8) % WORDS;
        if (OP == 0) atomicAdd(&s[a], 1u);                       // result unused: RED-style
        if (OP == 1) acc += atomicAdd(&s[a], 1u);                // result used (rank hand-out)
        if (OP == 2) s[a] = x;                                   // scattered 4-byte store
        if (OP == 3) reinterpret_cast<uint64_t*>(s)[a >> 1] = x; // scattered 8-byte store
        if (OP == 4) acc += s[a];                                // scattered load
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * 1024 + threadIdx.x] = acc + s[threadIdx.x];
}
template <int OP, int WORDS> void run(const char* name)
{
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    cudaFuncSetAttribute(k<OP, WORDS>, cudaFuncAttributeMaxDynamicSharedMemorySize, WORDS * 4);
    k<OP, WORDS><<<148, 1024, WORDS * 4>>>(out, cyc); k<OP, WORDS><<<148, 1024, WORDS * 4>>>(out, cyc);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-44s %9lld clk -> %.2f clk per warp-instruction per SM, %.3f clk per element\n", name, h[0],
           (double)h[0] / (32.0 * ITER), (double)h[0] / (1024.0 * ITER));
    cudaFree(out); cudaFree(cyc);
}
int main()
{
    run<0, 8192>("atomicAdd (unused), 8192 words"); run<1, 8192>("atomicAdd (rank used), 8192 words");
    run<1, 128>("atomicAdd (rank used), 128 words"); run<1, 49152>("atomicAdd (rank used), 49152 words");
    run<2, 8192>("store 4 B scattered, 8192 words"); run<3, 49152>("store 8 B scattered, 24576 slots");
    run<4, 8192>("load 4 B scattered, 8192 words");
    return 0;
}
