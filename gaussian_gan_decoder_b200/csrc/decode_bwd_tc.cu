// Decoder backward chain on tcgen05 (fp16 operands, fp32 accumulation in TMEM), for decoders
// without the positional embedding (layer-0 K = 48).  Same job as decode_bwd.cu — recompute a
// head's forward, propagate dL/d(outputs) to every layer's pre-activation gradient, to the
// sampled features and to the chained inputs — but every dense contraction is an MMA:
//   forward   z_l  = H_l  W_l^T          (4 MMAs per head, weights from the forward fp16 pack)
//   backward  dH_l = dZ_l W_l            (4 MMAs per head, weights from a transposed fp16 pack)
// The epilogue warps turn z_l into H_{l+1} = gelu(z_l) and G_l = gelu'(z_l) (kept in tensor memory as
// packed fp16, written with tcgen05.st), and dH_l into dZ_{l-1} = dH_l * G_{l-1}.  H_l and dZ_l tiles
// are streamed to the staging area with bulk async stores; decode_bwd_dw_mn.cu turns them into the
// hidden layers' weight gradients.  The feature / chained-input gradient dX_0 accumulates ACROSS heads in
// a dedicated TMEM accumulator, so the chain rule between heads costs nothing; the last layers' weight
// gradients accumulate ACROSS tiles in five more (dW3 += H_3^T dY, both operands MN-major).
// Replaces the same autograd graph as decode_bwd.cu (see there for reference citations).
#include "decode_bwd.cuh"
#include "decode_args.cuh"
#include "sampling.cuh"
#include "tc_ptx.cuh"

namespace ggd {

namespace {

constexpr int kM = 128, kK0 = 48, kKH = 144;
constexpr int kThreadsTc = 576;
constexpr uint32_t kSlotBytes = 128 * 144 * 2;            // 36864: largest weight block
constexpr uint32_t kWT3 = 128 * 16 * 2, kWT12 = 128 * 128 * 2, kWT0 = kK0 * 128 * 2;
constexpr uint32_t kBwdPackHead = kWT3 + 2 * kWT12 + kWT0;
// forward pack geometry (decode_tc.cu): [W0 128x48][W1 128x144][W2 128x144][W3 16x144]
constexpr uint32_t kF0 = 128 * kK0 * 2, kF12 = 128 * kKH * 2, kF3 = 16 * kKH * 2;
constexpr uint32_t kFwdPackHead = kF0 + 2 * kF12 + kF3;
constexpr float kC0 = 0.8f, kC1 = 0.0348f;

__host__ __device__ __forceinline__ uint32_t canon(uint32_t r, uint32_t k, uint32_t K_pad)
{
    return (r >> 3) * (K_pad * 16) + (k >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2;
}

// transposed weights for dH = dZ W:  B[N = input unit k][K = output unit n], K-major
__global__ void pack_weights_bwd_tc_kernel(const DecoderLayout L, const float* __restrict__ flat, uint8_t* __restrict__ packed)
{
    const int h = blockIdx.y / kLayers, l = blockIdx.y % kLayers;
    const int in = l == 0 ? L.in_dim[h] : kHidden, out = l == kLayers - 1 ? head_out(h) : kHidden;
    const int rows = l == 0 ? kK0 : kHidden, Kp = l == kLayers - 1 ? 16 : kHidden;
    const float* W = flat + L.flat_w[h][l];
    const uint32_t off = l == 3 ? 0 : (l == 2 ? kWT3 : (l == 1 ? kWT3 + kWT12 : kWT3 + 2 * kWT12));
    uint8_t* dst = packed + (size_t)h * kBwdPackHead + off;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * Kp; i += gridDim.x * blockDim.x) {
        const int k = i / Kp, n = i % Kp;
        const float v = (k < in && n < out) ? W[(size_t)n * in + k] : 0.f;
        *reinterpret_cast<__half*>(dst + canon(k, n, Kp)) = __float2half_rn(v);
    }
}

__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes)
{
    // The staging area is written once and read once by the weight-gradient kernel: 3.4 GB per step streaming through
    // L2.  Marked evict-first so that it does not push the feature planes (37.7 MB, gathered at random by every tile's
    // forward recompute) out of L2.
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 ::"l"(dst_gmem), "r"(ptx::smem_u32(src_smem)), "r"(bytes), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ float tmem_ld1(uint32_t taddr)
{
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr) : "memory");
    ptx::tmem_ld_wait();
    return __uint_as_float(r);
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16])
{
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, 512;" ::: "memory"); }

// gelu (tanh form fitted to erf-GELU, as in the forward kernel) and its derivative, half2
__device__ __forceinline__ void gelu_and_grad_h2(__half2 x, __half2& g, __half2& dg)
{
    const __half2 c0 = __float2half2_rn(kC0), c1 = __float2half2_rn(kC1), c13 = __float2half2_rn(3.f * kC1);
    const __half2 hf = __float2half2_rn(0.5f), one = __float2half2_rn(1.0f);
    const __half2 x2 = __hmul2(x, x);
    const __half2 th = ptx::tanh_h2(__hmul2(x, __hfma2(x2, c1, c0)));
    const __half2 hx = __hmul2(x, hf);
    g = __hfma2(hx, th, hx);
    const __half2 sech2 = __hfma2(__hneg2(th), th, one);                 // 1 - th^2
    const __half2 du = __hfma2(x2, c13, c0);                              // u'(x)
    dg = __hfma2(__hmul2(hx, du), sech2, __hfma2(hf, th, hf));            // 0.5(1+th) + 0.5 x u' (1-th^2)
}

}  // namespace

size_t bwd_tc_packed_bytes() { return (size_t)kBwdPackHead * kHeads; }

int launch_pack_weights_bwd_tc(cudaStream_t s, const DecoderLayout& L, const float* flat, void* packed)
{
    pack_weights_bwd_tc_kernel<<<dim3(8, kHeads * kLayers), 256, 0, s>>>(L, flat, static_cast<uint8_t*>(packed));
    GGD_LAUNCH_CHECK("pack_weights_bwd_tc_kernel");
    return GGD_OK;
}

}  // namespace ggd

#include "decode_bwd_tc_kernel.inl"
