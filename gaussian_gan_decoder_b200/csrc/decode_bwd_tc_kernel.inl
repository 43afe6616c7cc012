// Body of the tcgen05 decoder backward chain kernel (included by decode_bwd_tc.cu).
namespace ggd {
namespace {

// shared memory map (bytes)
constexpr int kWSlots = 3;   // weight blocks in flight: the loader runs up to three MMA steps ahead
constexpr uint32_t kOffX0 = 256, kOffA = kOffX0 + 128 * kK0 * 2, kOffDZ = kOffA + 128 * kKH * 2,
                   kOffDY = kOffDZ + kWT12, kOffDy32 = kOffDY + kWT3, kOffW = kOffDy32 + 128 * 4 * 4,
                   kSmemTotal = kOffW + kWSlots * kSlotBytes;
enum { BW_FULL = 0, BW_EMPTY = kWSlots, B_ACC = 2 * kWSlots, B_READY = 2 * kWSlots + 1, B_NBARS = 2 * kWSlots + 2 };
// TMEM columns: step accumulator | last-layer output | dX_0 (summed over heads) | gelu'(z_l) of the three hidden
// layers as packed fp16 (64 columns each) | last-layer weight gradients (16 columns per head, summed over tiles)
constexpr uint32_t kColAcc = 0, kColY = 128, kColDX = 160, kColG = 256;
__host__ __device__ constexpr uint32_t col_dw3(int hi) { return hi < 4 ? 448u + 16u * (uint32_t)hi : 144u; }

struct StepDesc { uint32_t src_off, bytes; };
__device__ __forceinline__ StepDesc step_weights(int s)
{
    switch (s) {
        case 0: return {0u, kF0};
        case 1: return {kF0, kF12};
        case 2: return {kF0 + kF12, kF12};
        case 3: return {kF0 + 2 * kF12, kF3};
        case 4: return {0u, kWT3};
        case 5: return {kWT3, kWT12};
        case 6: return {kWT3 + kWT12, kWT12};
        default: return {kWT3 + 2 * kWT12, kWT0};
    }
}

__global__ void __launch_bounds__(kThreadsTc, 1)
decode_bwd_chain_tc_kernel(const DecodeBwdTcArgs args)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const DecodeBwdArgs& a = args.b;
    const DecoderLayout& L = a.L;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 128);
    uint8_t* X0 = smem + kOffX0;
    uint8_t* Abuf = smem + kOffA;
    float* dy32 = reinterpret_cast<float*>(smem + kOffDy32);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const BwdStageTc st = bwd_stage_tc((size_t)a.P);
    const int tiles = (int)st.tiles;
    const int chain0 = kFeat + L.pos_dim;

    if (tid == 0) {
        for (int i = 0; i < B_NBARS; ++i) ptx::mbar_init(&bars[i], i == B_READY ? 16 : 1);
        ptx::mbar_fence_init();
    }
    if (warp == 16) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
#ifdef GGD_BWD_PROF
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long t_kernel = clock64();
#define PROF_T(x) const long long x = clock64()
#define PROF_ADD(k, t0) pc[k] += clock64() - (t0)
#else
#define PROF_T(x)
#define PROF_ADD(k, t0)
#endif

    if (warp < 16) {
        // =============================== epilogue warps ===============================
        const int r = tid & 127, colq = tid >> 7;
        const uint32_t trow = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        if (colq == 0) {   // constant-one column block of the hidden operand
            *reinterpret_cast<uint4*>(Abuf + canon(r, kHidden, kKH)) = make_uint4(0x00003c00u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(Abuf + canon(r, kHidden + 8, kKH)) = make_uint4(0u, 0u, 0u, 0u);
        }
        const size_t plane_stride = (size_t)a.D * a.H * a.W * kFeat;
        uint32_t n_acc = 0;
        auto wait_acc = [&]() {
            PROF_T(tw);
            if (lane == 0) ptx::mbar_wait(&bars[B_ACC], n_acc & 1);
            __syncwarp();
            ++n_acc;
            ptx::tc_fence_after();
            PROF_ADD(1, tw);
        };
        auto signal_ready = [&]() {
            ptx::tc_fence_before();
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&bars[B_READY]);
        };
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            const int p0 = tile * kM;
            PROF_T(ts);
            // ---- sampling into the layer-0 operand: the forward's fp16 planes when given (4 lanes per point, as in
            // decode_tc_kernel), else the fp32 planes (8 lanes per point) ----
            if (args.planes16) {
                const int sub = tid & 3, pt = tid >> 2, gps = p0 + pt;
                float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (gps < a.P) {
                    const float x = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gps);
                    const float y = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gps + 1);
                    const float z = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gps + 2);
                    const __half* pl = args.planes16;
                    if (a.arch == 0) {
                        sample8h<1>(pl, 1, a.H, a.W, x, y, 0.f, sub, acc);
                        sample8h<1>(pl + plane_stride, 1, a.H, a.W, x, z, 0.f, sub, acc);
                        sample8h<1>(pl + 2 * plane_stride, 1, a.H, a.W, z, x, 0.f, sub, acc);
                    } else {
                        sample8h<2>(pl, a.D, a.H, a.W, x, y, z, sub, acc);
                        sample8h<2>(pl + plane_stride, a.D, a.H, a.W, x, z, y, sub, acc);
                        sample8h<2>(pl + 2 * plane_stride, a.D, a.H, a.W, y, z, x, sub, acc);
                    }
                }
                uint32_t pk[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const __half2 hv = __floats2half2_rn(acc[2 * q] * (1.0f / 3.0f), acc[2 * q + 1] * (1.0f / 3.0f));   // rounded to fp16 next: no IEEE division
                    pk[q] = *reinterpret_cast<const uint32_t*>(&hv);
                }
                *reinterpret_cast<uint4*>(X0 + canon(pt, 8 * sub, kK0)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            } else {
                const int sub = tid & 7;
                for (int it = 0; it < 2; ++it) {
                    const int pt = it * 64 + (tid >> 3), gp = p0 + pt;
                    float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (gp < a.P) {
                        const float x = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp);
                        const float y = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 1);
                        const float z = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 2);
                        float4 s0, s1, s2;
                        if (a.arch == 0) {
                            s0 = sample4(a.planes, 0, 1, a.H, a.W, x, y, 0.f, sub);
                            s1 = sample4(a.planes + plane_stride, 0, 1, a.H, a.W, x, z, 0.f, sub);
                            s2 = sample4(a.planes + 2 * plane_stride, 0, 1, a.H, a.W, z, x, 0.f, sub);
                        } else {
                            s0 = sample4(a.planes, 1, a.D, a.H, a.W, x, y, z, sub);
                            s1 = sample4(a.planes + plane_stride, 1, a.D, a.H, a.W, x, z, y, sub);
                            s2 = sample4(a.planes + 2 * plane_stride, 1, a.D, a.H, a.W, y, z, x, sub);
                        }
                        m = make_float4((s0.x + s1.x + s2.x) / 3.f, (s0.y + s1.y + s2.y) / 3.f,
                                        (s0.z + s1.z + s2.z) / 3.f, (s0.w + s1.w + s2.w) / 3.f);
                    }
                    const __half2 lo = __floats2half2_rn(m.x, m.y), hi = __floats2half2_rn(m.z, m.w);
                    *reinterpret_cast<uint2*>(X0 + canon(pt, 4 * sub, kK0)) =
                        make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
                }
            }
            {
                if (colq == 0) {   // columns 32..47: position, chained inputs (saved forward results), zero padding, constant one
                    const int gp = p0 + r;
                    float f[16];
#pragma unroll
                    for (int k = 0; k < 16; ++k) f[k] = 0.f;
                    f[15] = 1.0f;
                    if (gp < a.P) {
#pragma unroll
                        for (int j = 0; j < 3; ++j) f[j] = __ldg(a.xyz + 3 * (size_t)gp + j);
                        if (L.variant != 0) {
                            int col = 3;      // chain0 - kFeat
#pragma unroll
                            for (int hi = 0; hi < kHeads - 1; ++hi) {
                                const int h = L.order[hi], n = head_out(h);
                                for (int j = 0; j < n; ++j) {
                                    const float v = __ldg(a.out[h] + (size_t)gp * n + j);
#pragma unroll
                                    for (int k = 3; k < 15; ++k) if (k == col + j) f[k] = v;     // static register indexing
                                }
                                col += n;
                            }
                        }
                    }
                    uint32_t w[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const __half2 hv = __floats2half2_rn(f[2 * q], f[2 * q + 1]);
                        w[q] = *reinterpret_cast<const uint32_t*>(&hv);
                    }
                    *reinterpret_cast<uint4*>(X0 + canon(r, kFeat, kK0)) = make_uint4(w[0], w[1], w[2], w[3]);
                    *reinterpret_cast<uint4*>(X0 + canon(r, kFeat + 8, kK0)) = make_uint4(w[4], w[5], w[6], w[7]);
                }
            }
            signal_ready();
            PROF_ADD(2, ts);

            for (int hi = kHeads - 1; hi >= 0; --hi) {
                const int h = L.order[hi], n_out = head_out(h);
                int my_chain = chain0;
                for (int e = 0; e < hi; ++e) my_chain += head_out(L.order[e]);
                const int gp = p0 + r;
                // ---- forward recompute: layers 0..2 ----
                for (int l = 0; l < 3; ++l) {
                    wait_acc();
                    uint32_t v[32];
                    ptx::tmem_ld32(trow + kColAcc + colq * 32, v);
                    ptx::tmem_ld_wait();
                    uint32_t od[16];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint32_t og[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            __half2 g, dg;
                            gelu_and_grad_h2(__floats2half2_rn(__uint_as_float(v[8 * j + 2 * q]), __uint_as_float(v[8 * j + 2 * q + 1])), g, dg);
                            og[q] = *reinterpret_cast<const uint32_t*>(&g);
                            od[4 * j + q] = *reinterpret_cast<const uint32_t*>(&dg);
                        }
                        *reinterpret_cast<uint4*>(Abuf + canon(r, colq * 32 + 8 * j, kKH)) = make_uint4(og[0], og[1], og[2], og[3]);
                    }
                    // gelu'(z_l) waits in tensor memory for the backward steps (16 packed columns per thread)
                    tmem_st16(trow + kColG + l * 64 + colq * 16, od);
                    tmem_st_wait();
                    signal_ready();
                }
                // ---- last layer: y -> dL/dy (activation derivative, chained gradient from later heads) ----
                wait_acc();
                if (colq == 0) {
                    uint32_t yv[4];
                    ptx::tmem_ld4(trow + kColY, yv);
                    ptx::tmem_ld_wait();
                    float dy[4] = {0.f, 0.f, 0.f, 0.f};
                    for (int j = 0; j < n_out; ++j) {
                        float gsum = 0.f;
                        if (L.variant != 0 && hi < kHeads - 1) gsum = tmem_ld1(trow + kColDX + my_chain + j);
                        if (gp < a.P) {
                            if (a.dout[h]) gsum += __ldg(a.dout[h] + (size_t)gp * n_out + j);
                            float d = 1.f;
                            if (h == kXyz) d = 0.01f;
                            if (h == kScale) d = -1.0f / (1.0f + expf(-(__uint_as_float(yv[j]) + 5.0f)));
                            dy[j] = gsum * d;
                        }
                    }
                    const __half2 d01 = __floats2half2_rn(dy[0], dy[1]), d23 = __floats2half2_rn(dy[2], dy[3]);
                    uint8_t* DY = smem + kOffDY;
                    *reinterpret_cast<uint4*>(DY + canon(r, 0, 16)) =
                        make_uint4(*reinterpret_cast<const uint32_t*>(&d01), *reinterpret_cast<const uint32_t*>(&d23), 0u, 0u);
                    *reinterpret_cast<uint4*>(DY + canon(r, 8, 16)) = make_uint4(0u, 0u, 0u, 0u);
                    *reinterpret_cast<float4*>(dy32 + 4 * r) = make_float4(dy[0], dy[1], dy[2], dy[3]);
                }
                epi_barrier();
                // last-layer bias gradient, reduced over the tile (the weight gradient is an MMA issued with the next
                // step: dW3 += H3^T dY into its own TMEM accumulator, flushed once at the end of the kernel)
                // (warp w < n_out sums component w over the 128 points: 4 rows per lane, then a butterfly — one thread
                // walking the 128 rows took ~4000 clocks per tile-head)
                if (warp < n_out) {
                    float sb = dy32[4 * lane + warp] + dy32[4 * (lane + 32) + warp] + dy32[4 * (lane + 64) + warp] + dy32[4 * (lane + 96) + warp];
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) sb += __shfl_xor_sync(0xffffffffu, sb, o);
                    if (lane == 0) atomicAdd(a.dflat + L.flat_b[h][3] + warp, sb);
                }
                signal_ready();
                // ---- backward: dZ_l = (dZ_{l+1} W_{l+1}) * G_l for l = 2, 1, 0 (in place over G_l) ----
                for (int l = 2; l >= 0; --l) {
                    wait_acc();
                    uint32_t v[32], gw[16];
                    ptx::tmem_ld32(trow + kColAcc + colq * 32, v);
                    tmem_ld16(trow + kColG + l * 64 + colq * 16, gw);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint32_t oz[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const __half2 dh = __floats2half2_rn(__uint_as_float(v[8 * j + 2 * q]), __uint_as_float(v[8 * j + 2 * q + 1]));
                            const __half2 dz = __hmul2(dh, *reinterpret_cast<const __half2*>(&gw[4 * j + q]));
                            oz[q] = *reinterpret_cast<const uint32_t*>(&dz);
                        }
                        *reinterpret_cast<uint4*>(smem + kOffDZ + canon(r, colq * 32 + 8 * j, kHidden)) = make_uint4(oz[0], oz[1], oz[2], oz[3]);
                    }
                    signal_ready();
                }
                // ---- dX_0 accumulates in TMEM across heads; nothing to convert ----
                wait_acc();
                if (hi > 0) signal_ready();
            }
            // ---- feature gradient -> plane gradient (scatter-add), thread = point ----
            if (a.dplanes && colq == 0) {
                uint32_t v[32];
                ptx::tmem_ld32(trow + kColDX, v);
                ptx::tmem_ld_wait();
                const int gp = p0 + r;
                if (gp < a.P) {
                    const float x = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp);
                    const float y = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 1);
                    const float z = a.coord_scale * __ldg(a.xyz + 3 * (size_t)gp + 2);
                    const size_t gplane = (size_t)kFeat * a.D * a.H * a.W;
                    for (int c4 = 0; c4 < 8; ++c4) {
                        const float4 g = make_float4(__uint_as_float(v[4 * c4]) / 3.f, __uint_as_float(v[4 * c4 + 1]) / 3.f,
                                                     __uint_as_float(v[4 * c4 + 2]) / 3.f, __uint_as_float(v[4 * c4 + 3]) / 3.f);
                        if (a.arch == 0) {
                            scatter4(a.dplanes, 0, 1, a.H, a.W, x, y, 0.f, c4, g);
                            scatter4(a.dplanes + gplane, 0, 1, a.H, a.W, x, z, 0.f, c4, g);
                            scatter4(a.dplanes + 2 * gplane, 0, 1, a.H, a.W, z, x, 0.f, c4, g);
                        } else {
                            scatter4(a.dplanes, 1, a.D, a.H, a.W, x, y, z, c4, g);
                            scatter4(a.dplanes + gplane, 1, a.D, a.H, a.W, x, z, y, c4, g);
                            scatter4(a.dplanes + 2 * gplane, 1, a.D, a.H, a.W, y, z, x, c4, g);
                        }
                    }
                }
            }
        }
        // ---- last-layer weight gradients: lane = input unit, 16 columns per head ----
        if (colq == 0) {
            for (int hi = 0; hi < kHeads; ++hi) {
                const int h = L.order[hi], n_out = head_out(h);
                uint32_t wv[4];
                ptx::tmem_ld4(trow + col_dw3(hi), wv);
                ptx::tmem_ld_wait();
                for (int j = 0; j < n_out; ++j) atomicAdd(a.dflat + L.flat_w[h][3] + (size_t)j * kHidden + r, __uint_as_float(wv[j]));
            }
        }
    }
#include "decode_bwd_tc_roles.inl"
#ifdef GGD_BWD_PROF
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 16))
        printf("bwd chain CTA0 warp %d: kernel %lld | wait_acc %lld sampling %lld | mma: wait_w %lld wait_ready %lld bulk_wait %lld issue %lld\n",
               warp, clock64() - t_kernel, pc[1], pc[2], pc[3], pc[4], pc[5], pc[6]);
#endif
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 16) ptx::tmem_dealloc<512>(tmem_base);
}

}  // namespace

int launch_decode_bwd_chain_tc(cudaStream_t s, const DecodeBwdTcArgs& a, int num_sms)
{
    GGD_CUDA_CHECK(cudaFuncSetAttribute(decode_bwd_chain_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemTotal));
    const int tiles = (a.b.P + kM - 1) / kM;
    decode_bwd_chain_tc_kernel<<<tiles < num_sms ? tiles : num_sms, kThreadsTc, kSmemTotal, s>>>(a);
    GGD_LAUNCH_CHECK("decode_bwd_chain_tc_kernel");
    return GGD_OK;
}

}  // namespace ggd
