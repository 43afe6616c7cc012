// Decoder backward, chain kernel: per 64-point tile and per head (in reverse chaining order)
// recompute the head's forward in fp32, propagate dL/d(outputs) back to dL/d(pre-activations)
// of every layer, to the sampled features (-> plane gradient, scatter-add) and to the chained
// inputs (-> earlier heads' output gradients).  The pre-activation gradients dZ_l and the layer
// inputs H_l are written, as fp16 tensor-core operands, to a staging area from which
// decode_bwd_dw.cu accumulates the weight gradients; the tiny last layer is reduced here.
//
// Replaces PyTorch autograd through main/decoder_models/base_decoder.py:21-27 (x5), the head
// post-math of the three wrapper classes, the Embedder (no gradient: positions are inputs) and
// grid_sample backward (eg3d/.../renderer.py:64, PanoHead/.../renderer.py:57).
//
// This is the exact-fp32 CUDA-core flavour (accuracy anchor; same role as decode_fwd.cu).
#include "decode_bwd.cuh"
#include "sampling.cuh"

namespace ggd {
namespace {

constexpr int kPts = kBwdPts;
constexpr int kThreads = 256;
constexpr int kKC = 32;
constexpr int kOutPerThread = kHidden / (kThreads / kPts);   // 32

__device__ __forceinline__ float gelu_f(float x) { return x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_grad(float x)
{
    return 0.5f * (1.0f + erff(x * 0.70710678118654752440f)) + x * 0.3989422804014327f * expf(-0.5f * x * x);
}

// zdst[n][pt] = bias[n] + sum_k src[k][pt] * Wt[k][n]   (Wt: packed [K_pad][128]); hdst = gelu(z)
__device__ __forceinline__ void dense_fwd(const float* __restrict__ src, int K_pad, const float* __restrict__ Wt,
                                          const float* __restrict__ bias, float* __restrict__ zdst,
                                          float* __restrict__ hdst, float* __restrict__ wbuf)
{
    const int pt = threadIdx.x & (kPts - 1), q = threadIdx.x >> 6;
    float acc[kOutPerThread];
#pragma unroll
    for (int j = 0; j < kOutPerThread; ++j) acc[j] = __ldg(bias + q * kOutPerThread + j);
    for (int k0 = 0; k0 < K_pad; k0 += kKC) {
        __syncthreads();
        const float4* g4 = reinterpret_cast<const float4*>(Wt + (size_t)k0 * kHidden);
        float4* w4s = reinterpret_cast<float4*>(wbuf);
#pragma unroll
        for (int i = 0; i < (kKC * kHidden / 4) / kThreads; ++i) w4s[threadIdx.x + i * kThreads] = __ldg(g4 + threadIdx.x + i * kThreads);
        __syncthreads();
#pragma unroll 4
        for (int kk = 0; kk < kKC; ++kk) {
            const float x = src[(k0 + kk) * kPts + pt];
            const float4* w4 = reinterpret_cast<const float4*>(wbuf + kk * kHidden + q * kOutPerThread);
#pragma unroll
            for (int j = 0; j < kOutPerThread / 4; ++j) {
                const float4 w = w4[j];
                acc[4 * j + 0] = fmaf(x, w.x, acc[4 * j + 0]); acc[4 * j + 1] = fmaf(x, w.y, acc[4 * j + 1]);
                acc[4 * j + 2] = fmaf(x, w.z, acc[4 * j + 2]); acc[4 * j + 3] = fmaf(x, w.w, acc[4 * j + 3]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kOutPerThread; ++j) {
        const int n = q * kOutPerThread + j;
        zdst[n * kPts + pt] = acc[j];
        hdst[n * kPts + pt] = gelu_f(acc[j]);
    }
}

// dst[k][pt] = sum_n dz[n][pt] * W[n][k]  (W: nn.Linear layout [128][in], row stride `in`), k < 128 (zero for k >= in)
__device__ __forceinline__ void dense_bwd(const float* __restrict__ dz, const float* __restrict__ W, int in,
                                          float (&acc)[kOutPerThread], float* __restrict__ wbuf)
{
    const int pt = threadIdx.x & (kPts - 1), q = threadIdx.x >> 6;
#pragma unroll
    for (int j = 0; j < kOutPerThread; ++j) acc[j] = 0.f;
    for (int n0 = 0; n0 < kHidden; n0 += kKC) {
        __syncthreads();
        for (int i = threadIdx.x; i < kKC * kHidden; i += kThreads) {
            const int nn = i >> 7, k = i & 127;
            wbuf[i] = k < in ? __ldg(W + (size_t)(n0 + nn) * in + k) : 0.f;
        }
        __syncthreads();
#pragma unroll 4
        for (int nn = 0; nn < kKC; ++nn) {
            const float x = dz[(n0 + nn) * kPts + pt];
            const float4* w4 = reinterpret_cast<const float4*>(wbuf + nn * kHidden + q * kOutPerThread);
#pragma unroll
            for (int j = 0; j < kOutPerThread / 4; ++j) {
                const float4 w = w4[j];
                acc[4 * j + 0] = fmaf(x, w.x, acc[4 * j + 0]); acc[4 * j + 1] = fmaf(x, w.y, acc[4 * j + 1]);
                acc[4 * j + 2] = fmaf(x, w.z, acc[4 * j + 2]); acc[4 * j + 3] = fmaf(x, w.w, acc[4 * j + 3]);
            }
        }
    }
}

// fp16 operand tile [rows][64 points], K-major: thread writes 8 consecutive points (16 bytes) of a row
__device__ __forceinline__ void store_tile(uint8_t* __restrict__ dst, const float* __restrict__ src, int rows_src,
                                           int rows_total, int one_row)
{
    for (int i = threadIdx.x; i < rows_total * (kPts / 8); i += kThreads) {
        const int f = i >> 3, c = i & 7;
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float v0 = 0.f, v1 = 0.f;
            if (f < rows_src) { v0 = src[f * kPts + 8 * c + 2 * e]; v1 = src[f * kPts + 8 * c + 2 * e + 1]; }
            else if (f == one_row) { v0 = 1.f; v1 = 1.f; }
            const __half2 h = __floats2half2_rn(v0, v1);
            pk[e] = *reinterpret_cast<const uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(dst + canon_off(f, 8 * c, kPts)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
}

}  // namespace
}  // namespace ggd

#include "decode_bwd_kernel.inl"
