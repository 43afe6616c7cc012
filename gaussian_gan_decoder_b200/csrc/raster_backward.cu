// Rasteriser backward: blend backward (per-pixel replay, back to front) and the
// per-Gaussian chain conic -> cov2D -> cov3D -> (scale, rotation), mean2D -> mean3D, RGB -> SH.
//
// Replaces BACKWARD::render -> renderCUDA (DGR/cuda_rasterizer/backward.cu:399-557),
// computeCov2DCUDA (:144-274), preprocessCUDA (:346-396), the SH backward (:20-139) and the
// cov3D backward (:278-341; like the reference, no quaternion-normalisation Jacobian).
//
// Gradients are floating-point sums, so parity with the reference is tolerance based (the
// reference itself is order-nondeterministic: 9 fp32 atomics per (pixel, Gaussian) pair).
// What differs structurally:
//  * the nine per-pair contributions are first summed over the 32 pixels of a warp with
//    shuffles and over the 8 warps of the tile in shared memory; global memory sees nine
//    atomics per (tile, Gaussian) instead of nine per (pixel, Gaussian);
//  * warps skip instances whose conservative footprint misses their 8x4 pixel block (footprint.cuh), and
//    the replay starts at the tile's largest n_contrib instead of the end of the list;
//  * the two per-Gaussian kernels of the reference are one kernel.
#include "common.cuh"
#include "footprint.cuh"

namespace ggd {
namespace {

constexpr int kBatch = kTilePixels;

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sums eight per-lane values over the warp with 9 shuffles instead of 40: every exchange also halves the
// number of values a lane still carries.  On return lane L holds, in v[0], the warp total of value
// q = 4 * bit4(L) + 2 * bit3(L) + bit2(L)  (the four lanes sharing L >> 2 hold the same total).
__device__ __forceinline__ float warp_sum8(float (&v)[8], int lane)
{
    {
        const bool up = lane & 16;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float recv = __shfl_xor_sync(0xffffffffu, up ? v[i] : v[i + 4], 16);
            v[i] = (up ? v[i + 4] : v[i]) + recv;
        }
    }
    {
        const bool up = lane & 8;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float recv = __shfl_xor_sync(0xffffffffu, up ? v[i] : v[i + 2], 8);
            v[i] = (up ? v[i + 2] : v[i]) + recv;
        }
    }
    {
        const bool up = lane & 4;
        const float recv = __shfl_xor_sync(0xffffffffu, up ? v[0] : v[1], 4);
        v[0] = (up ? v[1] : v[0]) + recv;
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
    return v[0];
}

__global__ void __launch_bounds__(kTilePixels)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                      const float* __restrict__ bg, const float2* __restrict__ means2D,
                      const float4* __restrict__ conic_opacity, const float* __restrict__ colors,
                      const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpixels, float* __restrict__ dL_dmean2D /*[P][3]*/,
                      float* __restrict__ dL_dconic /*[P][4]*/, float* __restrict__ dL_dopacity,
                      float* __restrict__ dL_dcolors, uint32_t* __restrict__ work /*[2]: next tile, finished CTAs*/,
                      uint32_t num_tiles)
{
    __shared__ uint32_t s_tile_id;
    __shared__ __align__(16) float4 s_rec[3 * kBatch];   // per instance: {x, y, conic.x, conic.y}, {conic.z, opacity, r, g}, {b, -, -, -}
    __shared__ uint32_t s_mask[kBatch];
    uint32_t rec_addr;   // pinned 32-bit shared address, see blend.cu
    asm volatile("mov.u32 %0, %1;" : "=r"(rec_addr) : "r"((uint32_t)__cvta_generic_to_shared(s_rec)));
    __shared__ uint32_t s_id[kBatch];
    __shared__ float s_acc[kBatch][9];  // per staged instance: dcolor[3], dmean2D[2], dconic[3], dopacity
    __shared__ uint32_t s_max;

    const uint32_t tiles_x = (W + kTileX - 1) / kTileX;
    const int tr = threadIdx.x, warp = tr >> 5, lane = tr & 31;
    // persistent CTAs pull tiles from a counter (see blend.cu: tile lists differ a lot in length)
    for (;;) {
    __syncthreads();
    if (tr == 0) s_tile_id = atomicAdd(work, 1u);
    __syncthreads();
    const uint32_t tile_id = s_tile_id;
    if (tile_id >= num_tiles) break;
    const uint32_t tile_bx = tile_id % tiles_x, tile_by = tile_id / tiles_x;
    const uint32_t px = tile_bx * kTileX + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_by * kTileY + (warp >> 1) * 4 + (lane >> 3);
    const uint32_t pix_id = W * py + px;
    const float fx = (float)px, fy = (float)py;
    const float tile_x0 = (float)(tile_bx * kTileX), tile_y0 = (float)(tile_by * kTileY);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint2 range = ranges[tile_id];

    const float T_final = inside ? final_Ts[pix_id] : 0.f;
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
    float dpix[3] = {0.f, 0.f, 0.f};
    if (inside) {
        const size_t plane = (size_t)H * W;
        dpix[0] = dL_dpixels[pix_id]; dpix[1] = dL_dpixels[plane + pix_id]; dpix[2] = dL_dpixels[2 * plane + pix_id];
    }
    const float bg_dot = bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2];
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

    // nothing behind the tile's largest n_contrib can contribute: start the replay there
    if (tr == 0) s_max = 0;
    __syncthreads();
    atomicMax(&s_max, last_contributor);
    __syncthreads();
    const int len = (int)min(s_max, range.y - range.x);
    const int rounds = (len + kBatch - 1) / kBatch;

    float T = T_final, last_alpha = 0.f;
    float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};

    for (int i = 0; i < rounds; ++i) {
        __syncthreads();
        const int progress = i * kBatch + tr;              // distance from the back of the (trimmed) list
        if (progress < len) {
            const uint32_t id = __ldg(point_list + range.x + (len - 1 - progress));
            const float2 xy = __ldg(means2D + id);
            const float4 co = __ldg(conic_opacity + id);
            s_id[tr] = id;
            s_rec[3 * tr] = make_float4(xy.x, xy.y, co.x, co.y);
            s_rec[3 * tr + 1] = make_float4(co.z, co.w, __ldg(colors + 3 * (size_t)id), __ldg(colors + 3 * (size_t)id + 1));
            s_rec[3 * tr + 2].x = __ldg(colors + 3 * (size_t)id + 2);
            s_mask[tr] = warp_block_mask(xy.x, xy.y, co.x, co.y, co.z, co.w, tile_x0, tile_y0);
        }
#pragma unroll
        for (int q = 0; q < 9; ++q) s_acc[tr][q] = 0.f;
        __syncthreads();

        const int lim = min(kBatch, len - i * kBatch);
        for (int j0 = 0; j0 < lim; j0 += 32) {
            const bool hit = (j0 + lane < lim) && ((s_mask[j0 + lane] >> warp) & 1u);
            uint32_t bits = __ballot_sync(0xffffffffu, hit);
            while (bits) {
                const int j = j0 + __ffs(bits) - 1;
                bits &= bits - 1;
                const uint32_t k = (uint32_t)(len - 1 - (i * kBatch + j));   // position in the tile list
                float g[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                bool active = k < last_contributor;
                if (active) {
                    const uint32_t ra = rec_addr + 48u * j;
                    float4 A, B;
                    float cb;
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(A.x), "=f"(A.y), "=f"(A.z), "=f"(A.w) : "r"(ra));
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+16];" : "=f"(B.x), "=f"(B.y), "=f"(B.z), "=f"(B.w) : "r"(ra));
                    asm volatile("ld.shared.f32 %0, [%1+32];" : "=f"(cb) : "r"(ra));
                    const float dx = A.x - fx, dy = A.y - fy;
                    const float power = -0.5f * (A.z * dx * dx + B.x * dy * dy) - A.w * dx * dy;
                    active = !(power > 0.0f);
                    if (active) {
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, B.y * G);
                        active = !(alpha < 1.0f / 255.0f);
                        if (active) {
                            T = T / (1.f - alpha);
                            const float dchannel_dcolor = alpha * T;
                            const float col[3] = {B.z, B.w, cb};
                            float dL_dalpha = 0.f;
#pragma unroll
                            for (int ch = 0; ch < 3; ++ch) {
                                accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                                last_color[ch] = col[ch];
                                dL_dalpha += (col[ch] - accum_rec[ch]) * dpix[ch];
                                g[ch] = dchannel_dcolor * dpix[ch];
                            }
                            dL_dalpha *= T;
                            last_alpha = alpha;
                            dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                            const float dL_dG = B.y * dL_dalpha;
                            const float gdx = G * dx, gdy = G * dy;
                            const float dG_ddelx = -gdx * A.z - gdy * A.w;
                            const float dG_ddely = -gdy * B.x - gdx * A.w;
                            g[3] = dL_dG * dG_ddelx * ddelx_dx;
                            g[4] = dL_dG * dG_ddely * ddely_dy;
                            g[5] = -0.5f * gdx * dx * dL_dG;
                            g[6] = -0.5f * gdx * dy * dL_dG;
                            g[7] = -0.5f * gdy * dy * dL_dG;
                            g[8] = G * dL_dalpha;
                        }
                    }
                }
                if (__any_sync(0xffffffffu, active)) {
                    // g[0..7] through the butterfly (lane L ends up with the total of value q(L)), g[8] the plain way
                    const float last = warp_sum(g[8]);
                    float (&g8)[8] = reinterpret_cast<float (&)[8]>(g);
                    const float tot = warp_sum8(g8, lane);
                    if ((lane & 3) == 0) atomicAdd(&s_acc[j][((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)], tot);
                    if (lane == 1) atomicAdd(&s_acc[j][8], last);
                }
            }
        }
        __syncthreads();
        if (progress < len) {
            const uint32_t id = s_id[tr];
            const float* acc = s_acc[tr];
            if (acc[0] != 0.f || acc[1] != 0.f || acc[2] != 0.f || acc[3] != 0.f || acc[4] != 0.f ||
                acc[5] != 0.f || acc[6] != 0.f || acc[7] != 0.f || acc[8] != 0.f) {
                atomicAdd(dL_dcolors + 3 * (size_t)id, acc[0]);
                atomicAdd(dL_dcolors + 3 * (size_t)id + 1, acc[1]);
                atomicAdd(dL_dcolors + 3 * (size_t)id + 2, acc[2]);
                atomicAdd(dL_dmean2D + 3 * (size_t)id, acc[3]);
                atomicAdd(dL_dmean2D + 3 * (size_t)id + 1, acc[4]);
                atomicAdd(dL_dconic + 4 * (size_t)id, acc[5]);
                atomicAdd(dL_dconic + 4 * (size_t)id + 1, acc[6]);
                atomicAdd(dL_dconic + 4 * (size_t)id + 3, acc[7]);
                atomicAdd(dL_dopacity + id, acc[8]);
            }
        }
    }
    }   // tile loop
    if (tr == 0 && atomicAdd(work + 1, 1u) == gridDim.x - 1) { work[0] = 0; work[1] = 0; }   // leave the queue reset
}

}  // namespace
}  // namespace ggd

#include "raster_backward_geom.inl"
