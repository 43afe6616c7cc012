// fp32 fused sample + decode forward kernel (see decode.cu for what it replaces).
#include "decode_args.cuh"
#include "sampling.cuh"

namespace ggd {



namespace {

constexpr int kPts = 128;       // points per tile
constexpr int kThreads = 256;   // thread = (point, half of the 128 outputs)
constexpr int kKC = 32;         // weight rows staged per step

__device__ __forceinline__ float gelu_erf(float x) { return x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float softplus(float x) { return x > 20.0f ? x : log1pf(expf(x)); }

// dst[n][pt] = act(bias[n] + sum_k src[k][pt] * Wt[k][n]),  n < 128
__device__ __forceinline__ void dense128(const float* __restrict__ src, int K_pad,
                                         const float* __restrict__ Wt, const float* __restrict__ bias,
                                         float* __restrict__ dst, float* __restrict__ wbuf)
{
    const int pt = threadIdx.x & (kPts - 1), half = threadIdx.x >> 7;
    float acc[64];
#pragma unroll
    for (int j = 0; j < 64; ++j) acc[j] = __ldg(bias + half * 64 + j);
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
            const float4* w4 = reinterpret_cast<const float4*>(wbuf + kk * kHidden + half * 64);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float4 w = w4[j];
                acc[4 * j + 0] = fmaf(x, w.x, acc[4 * j + 0]);
                acc[4 * j + 1] = fmaf(x, w.y, acc[4 * j + 1]);
                acc[4 * j + 2] = fmaf(x, w.z, acc[4 * j + 2]);
                acc[4 * j + 3] = fmaf(x, w.w, acc[4 * j + 3]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 64; ++j) dst[(half * 64 + j) * kPts + pt] = gelu_erf(acc[j]);
}

__global__ void __launch_bounds__(kThreads, 1)
decode_fwd_kernel(const DecodeFwdArgs a)
{
    extern __shared__ __align__(16) float smem[];
    float* base = smem;                       // [128 rows][128 pts] : feat | pos | chained outputs
    float* hA = base + kMaxIn * kPts;         // [128][128]
    float* hB = hA + kHidden * kPts;          // [128][128]
    float* wbuf = hB + kHidden * kPts;        // [32][128]
    const DecoderLayout& L = a.L;
    const int tid = threadIdx.x;
    const int tiles = (a.P + kPts - 1) / kPts;
    const size_t plane_stride = (size_t)a.D * a.H * a.W * kFeat;

    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int p0 = tile * kPts;
        __syncthreads();
        for (int i = tid; i < kMaxIn * kPts; i += kThreads) base[i] = 0.f;
        __syncthreads();
        // ---- sampling: 8 lanes per point, 4 channels per lane ----
        {
            const int sub = tid & 7, grp = tid >> 3;
            for (int it = 0; it < kPts / 32; ++it) {
                const int pt = it * 32 + grp, gp = p0 + pt;
                if (gp < a.P) {
                    const float x = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp);
                    const float y = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 1);
                    const float z = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 2);
                    float4 s0, s1, s2;
                    if (a.arch == 0) {   // (x,y) (x,z) (z,x)
                        s0 = sample4(a.planes, 0, 1, a.H, a.W, x, y, 0.f, sub);
                        s1 = sample4(a.planes + plane_stride, 0, 1, a.H, a.W, x, z, 0.f, sub);
                        s2 = sample4(a.planes + 2 * plane_stride, 0, 1, a.H, a.W, z, x, 0.f, sub);
                    } else {             // (x,y,z) (x,z,y) (y,z,x) -> (W,H,D)
                        s0 = sample4(a.planes, 1, a.D, a.H, a.W, x, y, z, sub);
                        s1 = sample4(a.planes + plane_stride, 1, a.D, a.H, a.W, x, z, y, sub);
                        s2 = sample4(a.planes + 2 * plane_stride, 1, a.D, a.H, a.W, y, z, x, sub);
                    }
                    base[(4 * sub + 0) * kPts + pt] = (s0.x + s1.x + s2.x) / 3.f;
                    base[(4 * sub + 1) * kPts + pt] = (s0.y + s1.y + s2.y) / 3.f;
                    base[(4 * sub + 2) * kPts + pt] = (s0.z + s1.z + s2.z) / 3.f;
                    base[(4 * sub + 3) * kPts + pt] = (s0.w + s1.w + s2.w) / 3.f;
                }
            }
        }
        // ---- position rows (identity, then sin/cos per linearly spaced frequency) ----
        float my_xyz[3] = {0.f, 0.f, 0.f};
        if (tid < kPts && p0 + tid < a.P) {
            for (int j = 0; j < 3; ++j) {
                my_xyz[j] = __ldg(a.xyz + 3 * (size_t)(p0 + tid) + j);
                base[(kFeat + j) * kPts + tid] = my_xyz[j];
            }
            for (int f = 0; f < a.pos_freqs; ++f) {
                // torch.linspace(1, 512, 10): start + f*step in fp32
                const float freq = 1.0f + (float)f * (511.0f / 9.0f);
                for (int j = 0; j < 3; ++j) {
                    base[(kFeat + 3 + 6 * f + j) * kPts + tid] = sinf(my_xyz[j] * freq);
                    base[(kFeat + 6 + 6 * f + j) * kPts + tid] = cosf(my_xyz[j] * freq);
                }
            }
        }
        __syncthreads();

        int chain_row = kFeat + L.pos_dim;
        for (int oi = 0; oi < kHeads; ++oi) {
            const int h = L.order[oi];
            const float* pk = a.packed;
            dense128(base, L.in_pad[h], pk + L.pk_w[h][0], pk + L.pk_b[h][0], hA, wbuf);
            dense128(hA, kHidden, pk + L.pk_w[h][1], pk + L.pk_b[h][1], hB, wbuf);
            dense128(hB, kHidden, pk + L.pk_w[h][2], pk + L.pk_b[h][2], hA, wbuf);
            __syncthreads();
            // last layer: [128] -> out (<= 4), weights [128][4]
            for (int i = tid; i < kHidden * 4; i += kThreads) wbuf[i] = __ldg(pk + L.pk_w[h][3] + i);
            __syncthreads();
            if (tid < kPts) {
                const float* b3 = pk + L.pk_b[h][3];
                float o0 = __ldg(b3), o1 = __ldg(b3 + 1), o2 = __ldg(b3 + 2), o3 = __ldg(b3 + 3);
#pragma unroll 8
                for (int k = 0; k < kHidden; ++k) {
                    const float x = hA[k * kPts + tid];
                    const float4 w = reinterpret_cast<const float4*>(wbuf)[k];
                    o0 = fmaf(x, w.x, o0); o1 = fmaf(x, w.y, o1); o2 = fmaf(x, w.z, o2); o3 = fmaf(x, w.w, o3);
                }
                float o[4] = {o0, o1, o2, o3};
                const int gp = p0 + tid;
                const int n = head_out(h);
                if (h == kXyz) for (int j = 0; j < 3; ++j) o[j] = o[j] * 0.01f + my_xyz[j];
                if (h == kScale) for (int j = 0; j < 3; ++j) o[j] = -softplus(o[j] + 5.0f) + a.scale_bias;
                if (gp < a.P) {
                    for (int j = 0; j < n; ++j) a.out[h][(size_t)gp * n + j] = o[j];
                    if (h == kScale && a.act_scale) for (int j = 0; j < 3; ++j) a.act_scale[(size_t)gp * 3 + j] = expf(o[j]);
                    if (h == kOpacity && a.act_opacity) a.act_opacity[gp] = 1.0f / (1.0f + expf(-o[0]));
                    if (h == kRot && a.act_rot) {
                        const float nrm = fmaxf(sqrtf(o[0] * o[0] + o[1] * o[1] + o[2] * o[2] + o[3] * o[3]), 1e-12f);
                        for (int j = 0; j < 4; ++j) a.act_rot[(size_t)gp * 4 + j] = o[j] / nrm;
                    }
                }
                if (L.variant != 0 && oi < kHeads - 1)
                    for (int j = 0; j < n; ++j) base[(chain_row + j) * kPts + tid] = o[j];
            }
            chain_row += head_out(h);
            __syncthreads();
        }
    }
}

}  // namespace

size_t decode_fwd_smem_bytes() { return sizeof(float) * (size_t)(kMaxIn * kPts + 2 * kHidden * kPts + kKC * kHidden); }

int launch_decode_fwd(cudaStream_t s, const DecodeFwdArgs& a, int num_sms)
{
    const size_t smem = decode_fwd_smem_bytes();
    GGD_CUDA_CHECK(cudaFuncSetAttribute(decode_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int tiles = (a.P + kPts - 1) / kPts;
    const int grid = tiles < num_sms ? tiles : num_sms;
    decode_fwd_kernel<<<grid, kThreads, smem, s>>>(a);
    GGD_LAUNCH_CHECK("decode_fwd_kernel");
    return GGD_OK;
}

}  // namespace ggd
