// Tri-plane / tri-grid texel fetch shared by the decode kernels.
// grid_sample semantics of the reference call sites (bilinear / trilinear, padding_mode="zeros",
// align_corners=False): eg3d/training/volumetric_rendering/renderer.py:64,
// PanoHead/training/volumetric_rendering/renderer.py:57.
#pragma once
#include "decode.cuh"
#include <cuda_fp16.h>

namespace ggd {

// Index of a sample's lower tap along an axis of n texels.  The conversion saturates for positions far outside the box
// (INT_MAX beyond ~1e9 texels, +inf) and "index + 1" then overflows — the out-of-range tests that follow were compiled on
// the assumption that it cannot, and a huge positive coordinate read outside the planes.  Clamped to [-2, n]: still out of
// range on the side it was, same bits for every position whose taps touch the planes (NaN converts to 0 as before).
__device__ __forceinline__ int tap_index(float floor_of_position, int n)
{
    return min(max((int)floor_of_position, -2), n);
}

// one bilinear / trilinear sample of 4 channels, zeros padding, align_corners=False
__device__ __forceinline__ float4 sample4(const float* __restrict__ plane, int arch, int D, int H, int W,
                                          float gx, float gy, float gz, int c4)
{
    const float ix = ((gx + 1.f) * W - 1.f) * 0.5f, iy = ((gy + 1.f) * H - 1.f) * 0.5f;
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const int x0 = tap_index(fx0, W), y0 = tap_index(fy0, H);
    const float wx1 = ix - fx0, wy1 = iy - fy0, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
    int z0 = 0;
    float wz0 = 1.f, wz1 = 0.f;
    if (arch == 1) {
        const float iz = ((gz + 1.f) * D - 1.f) * 0.5f, fz0 = floorf(iz);
        z0 = tap_index(fz0, D); wz1 = iz - fz0; wz0 = 1.f - wz1;
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

// fp16 channels-last planes, 8 channels (16 bytes) per lane.  Written for instruction count: bounds are folded into
// the per-axis weights — an out-of-range tap reads a clamped texel with weight 0 — and the byte offsets are per-axis
// too, so a tap costs one three-input add, one 16-byte load, a weight product and its 8 conversions + 8 FMAs: ~750
// instructions for the three tri-grid samples of a point quarter (1300 with per-tap bounds tests and 64-bit index
// arithmetic), same bits as before.
// (The mixed-precision FMA of sm_100 — fma.rn.f32.f16, SASS FHFMA — would save the conversions but wants the weight
// in fp16: measured, the frame against the reference loses 10 dB with trained-like weights.  Not used.)
// one axis of grid_sample (align_corners=False, zeros padding): weights of the two taps, byte offsets of their texels
__device__ __forceinline__ void sample_axis(float g, int n, uint32_t stride_bytes, float (&w)[2], uint32_t (&o)[2])
{
    const float i = ((g + 1.f) * n - 1.f) * 0.5f, f0 = floorf(i);
    const int i0 = tap_index(f0, n);
    const float w1 = i - f0, w0 = 1.f - w1;
    w[0] = (unsigned)i0 < (unsigned)n ? w0 : 0.f;
    w[1] = (unsigned)(i0 + 1) < (unsigned)n ? w1 : 0.f;
    o[0] = (uint32_t)min(max(i0, 0), n - 1) * stride_bytes;
    o[1] = (uint32_t)min(max(i0 + 1, 0), n - 1) * stride_bytes;
}
template <int NZ>   // 1: bilinear plane, 2: trilinear grid
__device__ __forceinline__ void sample8h(const __half* __restrict__ plane, int D, int H, int W,
                                         float gx, float gy, float gz, int c8, float (&acc)[8])
{
    const char* base = reinterpret_cast<const char*>(plane) + c8 * 16;
    float wx[2], wy[2], wz[2] = {1.f, 0.f};
    uint32_t ox[2], oy[2], oz[2] = {0u, 0u};
    sample_axis(gx, W, kFeat * 2, wx, ox);
    sample_axis(gy, H, (uint32_t)W * (kFeat * 2), wy, oy);
    if (NZ == 2) sample_axis(gz, D, (uint32_t)H * W * (kFeat * 2), wz, oz);
    uint4 v[NZ * 4];
    float w[NZ * 4];
#pragma unroll
    for (int dz = 0; dz < NZ; ++dz)
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                const int t = (dz * 2 + dy) * 2 + dx;
                w[t] = NZ == 2 ? wx[dx] * wy[dy] * wz[dz] : wx[dx] * wy[dy];
                v[t] = __ldg(reinterpret_cast<const uint4*>(base + (oz[dz] + oy[dy] + ox[dx])));
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

// One x column of a sample, split into "issue the loads" and "use them", so that a caller can keep the next plane's
// loads in flight while it accumulates the current one (the sampling phase of the tensor-core decoder is bound by
// memory latency x loads in flight per thread, not by issue, L1 tag lookups or L2 throughput — measured).
// The two lanes (dx = 0, 1) of a point's channel octet read the two x taps of every (z, y) row — the eight lanes of a
// point fetch 128 contiguous bytes per row; the caller adds the two lanes' sums.
template <int NZ>   // 1: bilinear plane, 2: trilinear grid
struct TapColumn {
    uint4 v[NZ * 2];
    float w[NZ * 2];
    __device__ __forceinline__ void issue(const char* __restrict__ plane_c8 /* plane + 16 * channel octet */, int D, int H, int W,
                                          float gx, float gy, float gz, int dx)
    {
        const float ix = ((gx + 1.f) * W - 1.f) * 0.5f, fx0 = floorf(ix);
        const int xi = tap_index(fx0, W) + dx;
        const float wx1 = ix - fx0, wxl = dx ? wx1 : 1.f - wx1;
        const float wx = (unsigned)xi < (unsigned)W ? wxl : 0.f;
        const char* base = plane_c8 + (uint32_t)min(max(xi, 0), W - 1) * (kFeat * 2);
        float wy[2], wz[2] = {1.f, 0.f};
        uint32_t oy[2], oz[2] = {0u, 0u};
        sample_axis(gy, H, (uint32_t)W * (kFeat * 2), wy, oy);
        if (NZ == 2) sample_axis(gz, D, (uint32_t)H * W * (kFeat * 2), wz, oz);
#pragma unroll
        for (int dz = 0; dz < NZ; ++dz)
#pragma unroll
            for (int dy = 0; dy < 2; ++dy) {
                w[dz * 2 + dy] = NZ == 2 ? wx * wy[dy] * wz[dz] : wx * wy[dy];
                v[dz * 2 + dy] = __ldg(reinterpret_cast<const uint4*>(base + (oz[dz] + oy[dy])));
            }
    }
    __device__ __forceinline__ void accumulate(float (&acc)[8]) const
    {
#pragma unroll
        for (int t = 0; t < NZ * 2; ++t) {
            const __half2* h = reinterpret_cast<const __half2*>(&v[t]);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 f = __half22float2(h[q]);
                acc[2 * q] = fmaf(w[t], f.x, acc[2 * q]);
                acc[2 * q + 1] = fmaf(w[t], f.y, acc[2 * q + 1]);
            }
        }
    }
};

}  // namespace ggd

namespace ggd {

// scatter-add of one sample's feature gradient into the generator-layout plane gradient
__device__ __forceinline__ void scatter4(float* __restrict__ dplane /*[32*D][H][W] of one plane*/, int arch, int D, int H, int W,
                                         float gx, float gy, float gz, int c4, float4 g)
{
    const float ix = ((gx + 1.f) * W - 1.f) * 0.5f, iy = ((gy + 1.f) * H - 1.f) * 0.5f;
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const int x0 = tap_index(fx0, W), y0 = tap_index(fy0, H);
    const float wx1 = ix - fx0, wy1 = iy - fy0, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
    int z0 = 0;
    float wz0 = 1.f, wz1 = 0.f;
    if (arch == 1) {
        const float iz = ((gz + 1.f) * D - 1.f) * 0.5f, fz0 = floorf(iz);
        z0 = tap_index(fz0, D); wz1 = iz - fz0; wz0 = 1.f - wz1;
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
