// Tri-plane / tri-grid texel fetch shared by the decode kernels.
// grid_sample semantics of the reference call sites (bilinear / trilinear, padding_mode="zeros",
// align_corners=False): eg3d/training/volumetric_rendering/renderer.py:64,
// PanoHead/training/volumetric_rendering/renderer.py:57.
#pragma once
#include "decode.cuh"
#include <cuda_fp16.h>

namespace ggd {

// one bilinear / trilinear sample of 4 channels, zeros padding, align_corners=False
__device__ __forceinline__ float4 sample4(const float* __restrict__ plane, int arch, int D, int H, int W,
                                          float gx, float gy, float gz, int c4)
{
    const float ix = ((gx + 1.f) * W - 1.f) * 0.5f, iy = ((gy + 1.f) * H - 1.f) * 0.5f;
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const int x0 = (int)fx0, y0 = (int)fy0;
    const float wx1 = ix - fx0, wy1 = iy - fy0, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
    int z0 = 0;
    float wz0 = 1.f, wz1 = 0.f;
    if (arch == 1) {
        const float iz = ((gz + 1.f) * D - 1.f) * 0.5f, fz0 = floorf(iz);
        z0 = (int)fz0; wz1 = iz - fz0; wz0 = 1.f - wz1;
    }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const int nz = arch == 1 ? 2 : 1;
    for (int dz = 0; dz < nz; ++dz) {
        const int z = z0 + dz;
        if (z < 0 || z >= D) continue;
        const float wz = dz ? wz1 : wz0;
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
            const int y = y0 + dy;
            if (y < 0 || y >= H) continue;
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                const int x = x0 + dx;
                if (x < 0 || x >= W) continue;
                const float w = (dx ? wx1 : wx0) * (dy ? wy1 : wy0) * wz;
                const float4 v = __ldg(reinterpret_cast<const float4*>(plane + (((size_t)z * H + y) * W + x) * kFeat) + c4);
                acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y);
                acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
            }
        }
    }
    return acc;
}

}  // namespace ggd

namespace ggd {

// fp16 channels-last planes, 8 channels (16 bytes) per lane.  All taps of a plane are issued as
// one batch of independent 16-byte loads (out-of-range taps read a clamped texel with weight 0),
// so a thread has up to 8 loads in flight instead of a load->FMA chain.
template <int NZ>   // 1: bilinear plane, 2: trilinear grid
__device__ __forceinline__ void sample8h(const __half* __restrict__ plane, int D, int H, int W,
                                         float gx, float gy, float gz, int c8, float (&acc)[8])
{
    const float ix = ((gx + 1.f) * W - 1.f) * 0.5f, iy = ((gy + 1.f) * H - 1.f) * 0.5f;
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const int x0 = (int)fx0, y0 = (int)fy0;
    const float wx[2] = {1.f - (ix - fx0), ix - fx0}, wy[2] = {1.f - (iy - fy0), iy - fy0};
    int z0 = 0;
    float wz[2] = {1.f, 0.f};
    if (NZ == 2) {
        const float iz = ((gz + 1.f) * D - 1.f) * 0.5f, fz0 = floorf(iz);
        z0 = (int)fz0; wz[1] = iz - fz0; wz[0] = 1.f - wz[1];
    }
    uint4 v[NZ * 4];
    float w[NZ * 4];
#pragma unroll
    for (int dz = 0; dz < NZ; ++dz)
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                const int x = x0 + dx, y = y0 + dy, z = z0 + dz;
                const bool ok = x >= 0 && x < W && y >= 0 && y < H && z >= 0 && z < D;
                const int xc = min(max(x, 0), W - 1), yc = min(max(y, 0), H - 1), zc = min(max(z, 0), D - 1);
                w[(dz * 2 + dy) * 2 + dx] = ok ? wx[dx] * wy[dy] * wz[dz] : 0.f;
                v[(dz * 2 + dy) * 2 + dx] =
                    __ldg(reinterpret_cast<const uint4*>(plane + (((size_t)zc * H + yc) * W + xc) * kFeat) + c8);
            }
#pragma unroll
    for (int t = 0; t < NZ * 4; ++t) {
        const __half2* h = reinterpret_cast<const __half2*>(&v[t]);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float2 f = __half22float2(h[q]);
            acc[2 * q] = fmaf(w[t], f.x, acc[2 * q]);
            acc[2 * q + 1] = fmaf(w[t], f.y, acc[2 * q + 1]);
        }
    }
}

}  // namespace ggd

namespace ggd {

// scatter-add of one sample's feature gradient into the generator-layout plane gradient
__device__ __forceinline__ void scatter4(float* __restrict__ dplane /*[32*D][H][W] of one plane*/, int arch, int D, int H, int W,
                                         float gx, float gy, float gz, int c4, float4 g)
{
    const float ix = ((gx + 1.f) * W - 1.f) * 0.5f, iy = ((gy + 1.f) * H - 1.f) * 0.5f;
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const int x0 = (int)fx0, y0 = (int)fy0;
    const float wx1 = ix - fx0, wy1 = iy - fy0, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
    int z0 = 0;
    float wz0 = 1.f, wz1 = 0.f;
    if (arch == 1) {
        const float iz = ((gz + 1.f) * D - 1.f) * 0.5f, fz0 = floorf(iz);
        z0 = (int)fz0; wz1 = iz - fz0; wz0 = 1.f - wz1;
    }
    const size_t HW = (size_t)H * W;
    const int nz = arch == 1 ? 2 : 1;
    for (int dz = 0; dz < nz; ++dz) {
        const int z = z0 + dz;
        if (z < 0 || z >= D) continue;
        for (int dy = 0; dy < 2; ++dy) {
            const int y = y0 + dy;
            if (y < 0 || y >= H) continue;
            for (int dx = 0; dx < 2; ++dx) {
                const int x = x0 + dx;
                if (x < 0 || x >= W) continue;
                const float w = (dx ? wx1 : wx0) * (dy ? wy1 : wy0) * (dz ? wz1 : wz0);
                // generator layout: channel index = c * D + z  (PanoHead/.../renderer.py:51-52)
                float* p = dplane + ((size_t)(4 * c4) * D + z) * HW + (size_t)y * W + x;
                atomicAdd(p, w * g.x);
                atomicAdd(p + (size_t)D * HW, w * g.y);
                atomicAdd(p + 2 * (size_t)D * HW, w * g.z);
                atomicAdd(p + 3 * (size_t)D * HW, w * g.w);
            }
        }
    }
}

}  // namespace ggd
