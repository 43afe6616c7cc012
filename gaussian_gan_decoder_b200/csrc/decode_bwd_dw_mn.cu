// (MN-major variant for the tensor-core chain kernel: operands are [point][feature] tiles of 128 points,
//  i.e. the contraction index K = point is the STRIDED one, so both descriptors are MN-major.)
// Decoder backward, weight gradients of the 15 hidden-layer matrices on tcgen05:
//   dW_l[n][k] = sum over points  dZ_l[p][n] * H_l[p][k]      (k = 128.. is the constant-one row -> db_l)
// Both operands were staged by the chain kernel as fp16, K-major with K = 64 points per tile, so a
// tile is one bulk async copy per operand and four M=128, N=48|112|144, K=16 MMAs.  Each CTA
// accumulates its share of the tiles in TMEM (fp32) and flushes once per matrix with atomics:
// 148 x 190k atomics per step instead of one per (point, weight).
// Replaces the reduction over P that PyTorch autograd performs in the 15 Linear backward GEMMs
// (main/decoder_models/base_decoder.py:11-19).
#include "decode_bwd.cuh"
#include "tc_ptx.cuh"

namespace ggd {
namespace {

constexpr int kStages = 3;
constexpr int kDwThreads = 192;          // warps 0-3 flush, warp 4 MMA, warp 5 loader
constexpr uint32_t kABytes = 128 * 128 * 2, kBBytesMax = 128 * 144 * 2;
constexpr uint32_t kStageBytes = kABytes + kBBytesMax;
constexpr int kTilePts = 128;

__global__ void __launch_bounds__(kDwThreads, 1)
decode_bwd_dw_mn_kernel(const DecoderLayout L, const uint8_t* __restrict__ stage, const size_t P, float* __restrict__ dflat)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);           // [kStages]
    uint64_t* empty = full + kStages;                              // [kStages]
    uint64_t* done = empty + kStages;                              // [1] accumulator complete
    uint64_t* drained = done + 1;                                  // [1] accumulator read out
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 128);
    uint8_t* buf = smem + 1024;
    const BwdStageTc st = bwd_stage_tc(P);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < kStages; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::mbar_init(done, 1);
        ptx::mbar_init(drained, 128);
        ptx::mbar_fence_init();
    }
    if (warp == 4) ptx::tmem_alloc<256>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const size_t tiles = st.tiles;
    const size_t my_tiles = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (warp == 5) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int m = 0; m < kHeads * 3; ++m) {
                const int h = m / 3, l = m % 3;
                for (size_t j = 0; j < my_tiles; ++j, ++it) {
                    const size_t t = blockIdx.x + j * gridDim.x;
                    const uint32_t sidx = it % kStages;
                    if (it >= kStages) ptx::mbar_wait(&empty[sidx], ((it / kStages) - 1) & 1);
                    uint8_t* dst = buf + (size_t)sidx * kStageBytes;
                    const uint32_t nb = (uint32_t)(l == 0 ? st.x0_tile : st.h_tile);
                    const uint8_t* bsrc = l == 0 ? stage + st.x0_off + t * st.x0_tile : stage + st.h_off[h][l - 1] + t * st.h_tile;
                    ptx::mbar_arrive_expect_tx(&full[sidx], kABytes + nb);
                    ptx::bulk_g2s(dst, stage + st.dz_off[h][l] + t * st.dz_tile, kABytes, &full[sidx]);
                    ptx::bulk_g2s(dst + kABytes, bsrc, nb, &full[sidx]);
                }
            }
        }
    } else if (warp == 4) {
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t sbase = ptx::smem_u32(buf);
            for (int m = 0; m < kHeads * 3; ++m) {
                const int l = m % 3;
                const int N = l == 0 ? 48 : kHidden + 16;
                const uint32_t idesc = ptx::idesc_f16(128, N) | (1u << 15) | (1u << 16);   // A and B MN-major
                const uint32_t b_lbo = (uint32_t)N * 16;
                if (m > 0) { ptx::mbar_wait(drained, (m - 1) & 1); ptx::tc_fence_after(); }
                for (size_t j = 0; j < my_tiles; ++j, ++it) {
                    const uint32_t sidx = it % kStages;
                    ptx::mbar_wait(&full[sidx], (it / kStages) & 1);
                    ptx::tc_fence_after();
                    const uint32_t a_addr = sbase + sidx * kStageBytes, b_addr = a_addr + kABytes;
                    // MN-major, no swizzle: leading offset = stride between 8-point (K) blocks, stride offset = 128 B
                    // between 8-feature (M/N) blocks; one K=16 step advances two K blocks
                    ptx::mma_f16_ss_steps<kTilePts / 16>(tmem_base, ptx::smem_desc(a_addr, 2048, 128), ptx::smem_desc(b_addr, b_lbo, 128),
                                                         2 * 2048 / 16, 2 * b_lbo / 16, idesc, j > 0);
                    ptx::mma_commit(&empty[sidx]);
                }
                ptx::mma_commit(done);
            }
        }
    } else {
        // ---- flush: thread = output unit n (TMEM lane); columns = input unit k ----
        const int n = tid;   // 0..127
        const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
        for (int m = 0; m < kHeads * 3; ++m) {
            const int h = m / 3, l = m % 3;
            const int in = l == 0 ? L.in_dim[h] : kHidden;
            const int N = l == 0 ? 48 : kHidden + 16;
            const int bias_col = l == 0 ? 47 : kHidden;
            ptx::mbar_wait(done, m & 1);
            ptx::tc_fence_after();
            if (my_tiles > 0) {
                float* dW = dflat + L.flat_w[h][l] + (size_t)n * in;
                for (int c = 0; c < N; c += 16) {
                    uint32_t v[16];
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        uint32_t r4[4];
                        ptx::tmem_ld4(trow + c + 4 * q4, r4);
                        v[4 * q4] = r4[0]; v[4 * q4 + 1] = r4[1]; v[4 * q4 + 2] = r4[2]; v[4 * q4 + 3] = r4[3];
                    }
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const int k = c + e;
                        const float g = __uint_as_float(v[e]);
                        if (k < in) atomicAdd(dW + k, g);
                        else if (k == bias_col) atomicAdd(dflat + L.flat_b[h][l] + n, g);
                    }
                }
            }
            ptx::tc_fence_before();
            ptx::mbar_arrive(drained);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 4) ptx::tmem_dealloc<256>(tmem_base);
}

}  // namespace

int launch_decode_bwd_dw_mn(cudaStream_t s, const DecoderLayout& L, const uint8_t* stage, size_t P, float* dflat, int num_sms)
{
    const size_t smem = 1024 + (size_t)kStages * kStageBytes;
    GGD_CUDA_CHECK(cudaFuncSetAttribute(decode_bwd_dw_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const size_t tiles = (P + 127) / 128;
    const int grid = (int)(tiles < (size_t)num_sms ? tiles : (size_t)num_sms);
    decode_bwd_dw_mn_kernel<<<grid, kDwThreads, smem, s>>>(L, stage, P, dflat);
    GGD_LAUNCH_CHECK("decode_bwd_dw_mn_kernel");
    return GGD_OK;
}

}  // namespace ggd
