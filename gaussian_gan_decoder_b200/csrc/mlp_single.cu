// Stand-alone ``Decoder.forward(triplane_features, gaussian_features)``: mean over the three
// plane samples, concatenate, 4-layer GELU MLP — one head, features supplied by the caller.
// Replaces main/decoder_models/base_decoder.py:21-27 when the module is called directly (the
// wrapper classes use the fused sample+decode kernels instead).  fp32 CUDA-core path; weights
// are read in their nn.Linear layout, no packing step.
#include "decode.cuh"

namespace ggd {
namespace {

constexpr int kPts = 128, kThreads = 256, kKC = 32;

__device__ __forceinline__ float gelu_erf(float x) { return x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f)); }

// dst[n][pt] = gelu(b[n] + sum_k src[k][pt] * W[n][k]),  W: [128][in] row-major
__device__ __forceinline__ void dense(const float* __restrict__ src, int in, const float* __restrict__ W,
                                      const float* __restrict__ b, float* __restrict__ dst, float* __restrict__ wbuf)
{
    const int pt = threadIdx.x & (kPts - 1), half = threadIdx.x >> 7;
    float acc[64];
#pragma unroll
    for (int j = 0; j < 64; ++j) acc[j] = __ldg(b + half * 64 + j);
    for (int k0 = 0; k0 < in; k0 += kKC) {
        __syncthreads();
        for (int i = threadIdx.x; i < kKC * kHidden; i += kThreads) {
            const int n = i >> 5, kk = i & 31;                       // coalesced along k
            wbuf[kk * kHidden + n] = (k0 + kk < in) ? __ldg(W + (size_t)n * in + k0 + kk) : 0.f;
        }
        __syncthreads();
        const int kmax = min(kKC, in - k0);
        for (int kk = 0; kk < kmax; ++kk) {
            const float x = src[(k0 + kk) * kPts + pt];
            const float4* w4 = reinterpret_cast<const float4*>(wbuf + kk * kHidden + half * 64);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float4 w = w4[j];
                acc[4 * j + 0] = fmaf(x, w.x, acc[4 * j + 0]); acc[4 * j + 1] = fmaf(x, w.y, acc[4 * j + 1]);
                acc[4 * j + 2] = fmaf(x, w.z, acc[4 * j + 2]); acc[4 * j + 3] = fmaf(x, w.w, acc[4 * j + 3]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 64; ++j) dst[(half * 64 + j) * kPts + pt] = gelu_erf(acc[j]);
}

__global__ void __launch_bounds__(kThreads, 1)
mlp_single_kernel(const float* __restrict__ tri /*[3][P][32]*/, const float* __restrict__ gf /*[P][k]*/, int k, int P,
                  const float* W0, const float* b0, const float* W1, const float* b1, const float* W2,
                  const float* b2, const float* W3, const float* b3, int n_out, float* __restrict__ out)
{
    extern __shared__ __align__(16) float smem[];
    float* x0 = smem;                       // [in <= 128][128]
    float* hA = x0 + kMaxIn * kPts;
    float* hB = hA + kHidden * kPts;
    float* wbuf = hB + kHidden * kPts;
    const int in = kFeat + k;
    const int tiles = (P + kPts - 1) / kPts;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int p0 = tile * kPts;
        __syncthreads();
        for (int i = threadIdx.x; i < in * kPts; i += kThreads) {
            const int row = i / kPts, pt = i % kPts, gp = p0 + pt;
            float v = 0.f;
            if (gp < P) {
                if (row < kFeat) {
                    const size_t o = (size_t)gp * kFeat + row;
                    v = (__ldg(tri + o) + __ldg(tri + (size_t)P * kFeat + o) + __ldg(tri + 2 * (size_t)P * kFeat + o)) / 3.f;
                } else {
                    v = __ldg(gf + (size_t)gp * k + (row - kFeat));
                }
            }
            x0[row * kPts + pt] = v;
        }
        __syncthreads();
        dense(x0, in, W0, b0, hA, wbuf);
        dense(hA, kHidden, W1, b1, hB, wbuf);
        dense(hB, kHidden, W2, b2, hA, wbuf);
        __syncthreads();
        for (int i = threadIdx.x; i < n_out * kPts; i += kThreads) {
            const int o = i / kPts, pt = i % kPts, gp = p0 + pt;
            if (gp >= P) continue;
            float acc = __ldg(b3 + o);
            for (int kk = 0; kk < kHidden; ++kk) acc = fmaf(hA[kk * kPts + pt], __ldg(W3 + o * kHidden + kk), acc);
            out[(size_t)gp * n_out + o] = acc;
        }
    }
}

}  // namespace
}  // namespace ggd

using namespace ggd;

extern "C" int ggd_mlp_forward(void* stream, const float* tri_features, const float* gaussian_features, int k, int P,
                               const float* W0, const float* b0, const float* W1, const float* b1, const float* W2,
                               const float* b2, const float* W3, const float* b3, int hidden, int out_features, float* out)
{
    if (P <= 0 || k < 0 || kFeat + k > kMaxIn) { set_error("ggd_mlp_forward: unsupported sizes P=%d k=%d", P, k); return GGD_E_INVALID; }
    if (hidden != kHidden || out_features < 1 || out_features > 128) { set_error("ggd_mlp_forward: hidden must be 128, out_features in [1,128]"); return GGD_E_INVALID; }
    if (!tri_features || (k > 0 && !gaussian_features) || !W0 || !b0 || !W1 || !b1 || !W2 || !b2 || !W3 || !b3 || !out) { set_error("ggd_mlp_forward: null pointer"); return GGD_E_INVALID; }
    const size_t smem = sizeof(float) * (size_t)((kMaxIn + 2 * kHidden) * kPts + kKC * kHidden);
    GGD_CUDA_CHECK(cudaFuncSetAttribute(mlp_single_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, sms = 0;
    GGD_CUDA_CHECK(cudaGetDevice(&dev));
    GGD_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int tiles = (P + kPts - 1) / kPts;
    mlp_single_kernel<<<tiles < sms ? tiles : sms, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        tri_features, gaussian_features, k, P, W0, b0, W1, b1, W2, b2, W3, b3, out_features, out);
    GGD_LAUNCH_CHECK("mlp_single_kernel");
    return GGD_OK;
}
