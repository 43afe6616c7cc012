// Tri-plane / tri-grid texel fetch shared by the decode kernels.
// grid_sample semantics of the reference call sites (bilinear / trilinear, padding_mode="zeros",
// align_corners=False): eg3d/training/volumetric_rendering/renderer.py:64,
// PanoHead/training/volumetric_rendering/renderer.py:57.
#pragma once
#include "decode.cuh"

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
