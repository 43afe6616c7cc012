// Kernel 1, tensor-core flavour: fused tri-plane sampling + five decoder MLPs on
// tcgen05 (fp16 operands, fp32 accumulators in TMEM) for sm_100a.
//
// Replaces the same reference code as decode.cu / decode_fwd.cu (sample_from_planes,
// Embedder, 5 x Decoder, head post-math, GaussianModel activations).
//
// Structure (one persistent CTA per SM, 18 warps):
//   warps 0-15       epilogue: thread = one point (TMEM lane) x one quarter of the 128 accumulator
//                    columns.  They sample the planes into the layer-0 operands, turn an
//                    accumulator into the next layer's fp16 operand (GELU in half2 on the FMA pipe:
//                    gelu_poly_h2) and apply the head post-math (one output component per warp of a
//                    lane quarter).  Three 128-point tiles ("slots", each with its own operand stage
//                    and TMEM accumulator) are in flight, layer by layer: while the warps convert one
//                    slot, the tensor core runs the other slots' MMAs.
//   warp 16          one thread issues tcgen05.mma (M=128, N=128 or 16, K=16 per instruction) from
//                    descriptors built once per tile-layer, K loop unrolled (ptx::mma_f16_ss_steps)
//   warp 17          one thread streams the packed fp16 weights of the current head, layer by
//                    layer, with 1-D bulk async copies into four shared-memory slots; a slot
//                    is refilled with the next head's layer as soon as its last MMA retired.
// A group of three tiles (384 points), one per slot, shares every weight load.
// Operands use the K-major, no-swizzle canonical layout (8x16-byte core matrices); biases are
// folded into the GEMMs: the weights carry a bias column block, whose A-side partner is a
// constant-one column of the layer-0 operand and, for the hidden layers, one 4 KB constant-one
// block shared by all slots (a ninth K step).
// Shared memory (T = 3, no positional embedding): 256 B barriers + 90,624 B weights + 3 x 32,768 B
// operand stages + 4,096 B ones block + 3 x 12,288 B layer-0 operands = 230,144 B of 232,448.
#include "decode_args.cuh"
#include "sampling.cuh"
#include "tc_ptx.cuh"
#include <cstdlib>

namespace ggd {

namespace {

constexpr int kM = 128;                       // points per tile == TMEM lanes
constexpr int kKH = kHidden + 16;             // hidden-layer K incl. the bias column block
constexpr int kNOut = 16;                     // last layer N (>= 4 outputs, MMA minimum for M=128)
constexpr int kAStageBytes = kM * kHidden * 2;    // 32768: one hidden-layer A operand (the bias column block is shared, kOnesBytes)
constexpr int kWHBytes = kHidden * kKH * 2;      // 36864: hidden-layer weights incl. the bias columns
constexpr int kOnesBytes = kM * 16 * 2;          // 4096: [128 rows][16 columns] = 1, 0, .., 0 — the A operand of every bias step
constexpr int kSlots = 3;                        // tiles in flight (operand stage + accumulator each)
constexpr int kW3Bytes = kNOut * kKH * 2;     // 4608
// tensor-memory columns of one slot: hidden accumulator | last-layer accumulator | fp16 activations (A operand)
constexpr uint32_t kColOut = 128, kColSlot = 160;   // 3 slots x 160 columns <= 512
constexpr int kTcThreads = 576;             // 16 epilogue warps + MMA warp + loader warp
constexpr float kGeluC0 = 0.8f, kGeluC1 = 0.0348f;   // tanh-form fit to erf-GELU, max abs err 2.9e-4

// byte offset of element (r, k) in a K-major no-swizzle operand with K_pad columns (fp16)
__host__ __device__ __forceinline__ uint32_t canon(uint32_t r, uint32_t k, uint32_t K_pad)
{
    return (r >> 3) * (K_pad * 16) + (k >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2;
}

struct TcLayout {
    int k0;              // layer-0 K (48, or 112 with the positional embedding)
    size_t head_bytes;   // packed bytes per head
    size_t off[kLayers];  // byte offset of each layer inside a head block
    size_t bytes[kLayers];
};
__host__ __device__ inline TcLayout tc_layout(int pos_dim)
{
    TcLayout t;
    t.k0 = pos_dim == 3 ? 48 : 112;
    t.bytes[0] = (size_t)kM * t.k0 * 2;
    t.bytes[1] = t.bytes[2] = kWHBytes;
    t.bytes[3] = kW3Bytes;
    t.off[0] = 0;
    for (int l = 1; l < kLayers; ++l) t.off[l] = t.off[l - 1] + t.bytes[l - 1];
    t.head_bytes = t.off[3] + t.bytes[3];
    return t;
}

__global__ void pack_weights_tc_kernel(const DecoderLayout L, const float* __restrict__ flat, uint8_t* __restrict__ packed)
{
    const TcLayout t = tc_layout(L.pos_dim);
    const int h = blockIdx.y / kLayers, l = blockIdx.y % kLayers;
    const int in = l == 0 ? L.in_dim[h] : kHidden, out = l == kLayers - 1 ? head_out(h) : kHidden;
    const int K = l == 0 ? t.k0 : kKH, N = l == kLayers - 1 ? kNOut : kHidden;
    const int bias_col = l == 0 ? t.k0 - 1 : kHidden;
    const float* W = flat + L.flat_w[h][l];
    const float* b = flat + L.flat_b[h][l];
    uint8_t* dst = packed + (size_t)h * t.head_bytes + t.off[l];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N * K; i += gridDim.x * blockDim.x) {
        const int n = i / K, k = i % K;
        float v = 0.f;
        if (n < out) v = k < in ? W[(size_t)n * in + k] : (k == bias_col ? b[n] : 0.f);
        *reinterpret_cast<__half*>(dst + canon(n, k, K)) = __float2half_rn(v);
    }
}

__device__ __forceinline__ __half2 gelu_h2(__half2 x)
{
    const __half2 c0 = __float2half2_rn(kGeluC0), c1 = __float2half2_rn(kGeluC1), hf = __float2half2_rn(0.5f);
    const __half2 t = __hfma2(__hmul2(x, x), c1, c0);
    const __half2 th = ptx::tanh_h2(__hmul2(x, t));
    const __half2 hx = __hmul2(x, hf);
    return __hfma2(hx, th, hx);
}
// Same function without the MUFU pipe: gelu(x) = relu(x) - |x| * Phi(-|x|), with
//   Phi(-a) = 0.5 erfc(a / sqrt 2) ~= relu(q(a))^2,   q = quartic fitted over [0, 6] (minimax on a * error, weighted
//   towards small a) with q(0) pinned to the fp16 value next to sqrt(1/2), so that the slope at 0 is right: with a free
//   q(0) the fit had Phi(0) off by 2.6e-3, a 0.5 % gain error on the small activations of a freshly initialised decoder
//   (frame against the reference 96 dB instead of 105).  All five coefficients are fp16 values.
// Max abs error of the resulting GELU 3.1e-4 in exact arithmetic; evaluated in half2 (Horner) the simulated rms error
// against the exact erf-GELU is 2.7e-4 over [-6, 6] and 0.8e-4 over [-0.5, 0.5] (3.1e-4 / 2.8e-4 with the free fit).  Six FMA-pipe instructions per half2 (4 HFMA2 + HMUL2 + HFMA2; round 1
// shipped relu(cubic)^8 = seven) plus one HMNMX2 on the ALU pipe for the clamp.  Measured on B200 (profiles/ubench_pipes.cu):
// HFMA2/HMUL2 issue every 2 clocks per SM sub-partition, MUFU every 8, so the tanh form costs 16 MUFU + 10 FMA-pipe
// clocks per half2 and this form 12 FMA-pipe clocks.  kMask selects the form per half2 of a thread's row segment.
constexpr float kPolyQ0 = 0.70703125f, kPolyQ1 = -0.27783203f, kPolyQ2 = -0.07385254f, kPolyQ3 = 0.0491333f,
                kPolyQ4 = -0.0061454773f, kPolyClamp = 6.0f;
__device__ __forceinline__ __half2 gelu_poly_h2(__half2 x)
{
    const __half2 q0 = __float2half2_rn(kPolyQ0), q1 = __float2half2_rn(kPolyQ1), q2 = __float2half2_rn(kPolyQ2),
                  q3 = __float2half2_rn(kPolyQ3), q4 = __float2half2_rn(kPolyQ4), zero = __float2half2_rn(0.f);
    // The fit holds on [0, 6] and the quartic leaves [0, 1] beyond it (round 1's cubic did too: at |x| ~ 10 its 8th
    // power overflowed and the fp16 path returned NaN for trained-like weights).  Clamp the argument where Phi(-a)
    // has long vanished (6 * Phi(-6) = 6e-9): gelu(x) = relu(x) - O(1e-4) for every |x| > 6.
    const __half2 ax = __hmin2(__habs2(x), __float2half2_rn(kPolyClamp));
    __half2 p = __hfma2(ax, q4, q3);
    p = __hfma2(p, ax, q2);
    p = __hfma2(p, ax, q1);
    p = __hfma2_relu(p, ax, q0);
    p = __hmul2(p, p);
    return __hfma2(__hneg2(ax), p, __hmax2(x, zero));
}
template <uint32_t kMask, int kIdx>
__device__ __forceinline__ __half2 gelu_mix_h2(__half2 x)
{
    if constexpr ((kMask >> kIdx) & 1u) return gelu_poly_h2(x);
    else return gelu_h2(x);
}
__device__ __forceinline__ float softplus_f(float x) { return x > 20.0f ? x : log1pf(expf(x)); }

}  // namespace

size_t tc_packed_bytes(const DecoderLayout& L) { return tc_layout(L.pos_dim).head_bytes * kHeads; }

int launch_pack_weights_tc(cudaStream_t s, const DecoderLayout& L, const float* flat, void* packed)
{
    pack_weights_tc_kernel<<<dim3(8, kHeads * kLayers), 256, 0, s>>>(L, flat, static_cast<uint8_t*>(packed));
    GGD_LAUNCH_CHECK("pack_weights_tc_kernel");
    return GGD_OK;
}

}  // namespace ggd

#include "decode_tc_kernel.inl"
