// Stage 1 of the rasteriser: per-Gaussian projection, EWA covariance, conic,
// screen radius, tile rectangle, SH→RGB, and the inclusive scan of
// tiles_touched (single pass, decoupled look-back).
//
// Replaces DGR/cuda_rasterizer/forward.cu:155-256 (preprocessCUDA and the
// helpers it calls: computeCov3D :118-152, computeCov2D :74-113,
// computeColorFromSH :20-71, auxiliary.h in_frustum :139-164, getRect :46-56,
// ndc2Pix :41-44) and cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:277).
//
// Bit-exactness contract.  radii, tiles_touched and the depth bits feed the
// integer stages, so every floating-point expression below keeps the
// reference's operand order and association (nvcc contracts a*b+c chains the
// same way for the same expression tree, default -fmad=true, no fast-math):
// 3x3 products are summed k = 0,1,2 with the structural zeros left in, the
// NDC→pixel map runs in double, divisions and square roots are IEEE.
// What is NOT taken from the reference is the data movement: the stride-3
// AoS inputs (xyz, scale, SH) and outputs (cov3D, rgb, clamped) go through
// shared memory so that every global access of a warp is a contiguous,
// 16-byte-vectorised segment instead of 12-byte-strided scalars.
#include "common.cuh"

namespace ggd {
namespace {

constexpr int kThreads = 256;
constexpr uint32_t kRectDiffMin = 32;   // tile rectangles from this size on are counted through the difference array

struct M3 {  // column-major 3x3, m[col][row]
    float m[3][3];
};
__device__ __forceinline__ M3 mul(const M3& a, const M3& b)
{
    M3 c;
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i)
            c.m[j][i] = a.m[0][i] * b.m[j][0] + a.m[1][i] * b.m[j][1] + a.m[2][i] * b.m[j][2];
    return c;
}
__device__ __forceinline__ M3 transpose(const M3& a)
{
    M3 t;
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i) t.m[j][i] = a.m[i][j];
    return t;
}

__device__ __forceinline__ float3 xform4x3(const float3& p, const float* __restrict__ m)
{
    return make_float3(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                       m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                       m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]);
}
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* __restrict__ m)
{
    return make_float4(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                       m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                       m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
                       m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]);
}

__device__ __forceinline__ float ndc_to_pix(float v, int S)
{
    return ((v + 1.0) * S - 1.0) * 0.5;  // double on purpose (auxiliary.h:41-44)
}

}  // namespace

namespace {

constexpr float kSH0 = 0.28209479177387814f;
constexpr float kSH1 = 0.4886025119029199f;
__constant__ float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                              -1.0925484305920792f, 0.5462742152960396f};
__constant__ float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                              0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                              -0.5900435899266435f};

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 operator*(float s, V3 v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

// SH evaluation for degree > 0 (not on the decoder path, which is degree 0, but
// on the API surface).  sh points at this Gaussian's [M][3] coefficients.
__device__ V3 sh_color_hi(int deg, V3 result, const float3 pos, const float* __restrict__ campos,
                          const float* __restrict__ shp)
{
    const V3* sh = reinterpret_cast<const V3*>(shp);
    V3 dir = {pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]};
    const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir = {dir.x / len, dir.y / len, dir.z / len};
    const float x = dir.x, y = dir.y, z = dir.z;
    result = result - kSH1 * y * sh[1] + kSH1 * z * sh[2] - kSH1 * x * sh[3];
    if (deg > 1) {
        const float xx = x * x, yy = y * y, zz = z * z;
        const float xy = x * y, yz = y * z, xz = x * z;
        result = result + kSH2[0] * xy * sh[4] + kSH2[1] * yz * sh[5] +
                 kSH2[2] * (2.0f * zz - xx - yy) * sh[6] + kSH2[3] * xz * sh[7] +
                 kSH2[4] * (xx - yy) * sh[8];
        if (deg > 2) {
            result = result + kSH3[0] * y * (3.0f * xx - yy) * sh[9] + kSH3[1] * xy * z * sh[10] +
                     kSH3[2] * y * (4.0f * zz - xx - yy) * sh[11] +
                     kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                     kSH3[4] * x * (4.0f * zz - xx - yy) * sh[13] + kSH3[5] * z * (xx - yy) * sh[14] +
                     kSH3[6] * x * (xx - 3.0f * yy) * sh[15];
        }
    }
    return result;
}

// Coalesced block load of `n_floats` contiguous floats starting at src into
// shared memory (16-byte vectors where alignment allows).
__device__ __forceinline__ void block_load(float* __restrict__ dst, const float* __restrict__ src,
                                           int n_floats)
{
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        const int n4 = n_floats >> 2;
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int i = threadIdx.x; i < n4; i += kThreads) d4[i] = __ldg(s4 + i);
        for (int i = (n4 << 2) + threadIdx.x; i < n_floats; i += kThreads) dst[i] = __ldg(src + i);
    } else {
        for (int i = threadIdx.x; i < n_floats; i += kThreads) dst[i] = __ldg(src + i);
    }
}
__device__ __forceinline__ void block_store(float* __restrict__ dst, const float* __restrict__ src,
                                            int n_floats)
{
    if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        const int n4 = n_floats >> 2;
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int i = threadIdx.x; i < n4; i += kThreads) d4[i] = s4[i];
        for (int i = (n4 << 2) + threadIdx.x; i < n_floats; i += kThreads) dst[i] = src[i];
    } else {
        for (int i = threadIdx.x; i < n_floats; i += kThreads) dst[i] = src[i];
    }
}

__global__ void __launch_bounds__(kThreads)
preprocess_kernel(const PreprocessArgs a, float* __restrict__ depths, int* __restrict__ radii,
                  float2* __restrict__ means2D, float* __restrict__ cov3D_out,
                  float4* __restrict__ conic_opacity, float* __restrict__ rgb_out,
                  uint32_t* __restrict__ tiles_touched, uint8_t* __restrict__ clamped_out)
{
    // staging: xyz | scale or cov3D | rgb (in: SH dc, out: rgb) ; cov3D out aliases its own
    __shared__ __align__(16) float s_xyz[kThreads * 3];
    __shared__ __align__(16) float s_cov[kThreads * 6];   // in: scales (3/pt) or cov3D_precomp (6/pt); out: cov3D
    __shared__ __align__(16) float s_col[kThreads * 3];   // in: SH dc (M==1) ; out: rgb
    __shared__ __align__(16) uint8_t s_clamp[kThreads * 3];
    extern __shared__ uint32_t s_cnt[];   // [a.hist_tiles] CTA-local tile histogram (dense frames, see below)

    const int base = blockIdx.x * kThreads;
    const int n_here = min(kThreads, a.P - base);
    const int t = threadIdx.x;
    const int idx = base + t;
    const bool have_cov = a.cov3D_precomp != nullptr;
    const bool have_col = a.colors_precomp != nullptr;
    const bool sh_staged = !have_col && a.M == 1;
    if (a.tile_counts && blockIdx.x == 0 && t == 0) *a.counts_valid = 1u;

    block_load(s_xyz, a.means3D + (size_t)base * 3, n_here * 3);
    if (have_cov) block_load(s_cov, a.cov3D_precomp + (size_t)base * 6, n_here * 6);
    else          block_load(s_cov, a.scales + (size_t)base * 3, n_here * 3);
    if (sh_staged) block_load(s_col, a.shs + (size_t)base * 3, n_here * 3);
    __syncthreads();

    int my_radius_i = 0;
    uint32_t my_tiles = 0;
    uint2 my_rmin = make_uint2(0u, 0u), my_rmax = make_uint2(0u, 0u);
    float cov6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float col3[3] = {0.f, 0.f, 0.f};
    bool col_valid = false, cov_valid = false;
    uint8_t cl0 = 0, cl1 = 0, cl2 = 0;

    if (t < n_here) {
        const float3 p = make_float3(s_xyz[3 * t], s_xyz[3 * t + 1], s_xyz[3 * t + 2]);
        const float4 hom = xform4x4(p, a.proj);
        const float pw = 1.0f / (hom.w + 0.0000001f);
        const float3 proj = make_float3(hom.x * pw, hom.y * pw, hom.z * pw);
        const float3 pv = xform4x3(p, a.view);

        bool alive = !(pv.z <= 0.2f);  // same NaN behaviour as the reference test (auxiliary.h:154)
        if (!alive && a.prefiltered) {
            printf("Point is filtered although prefiltered is set. This shouldn't happen!");
            __trap();
        }
        if (alive) {
            if (have_cov) {
#pragma unroll
                for (int k = 0; k < 6; ++k) cov6[k] = s_cov[6 * t + k];
            } else {
                // Sigma = (S R)^T (S R); rotation is NOT normalised here (forward.cu:127).
                const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
                const float r = q.x, x = q.y, y = q.z, z = q.w;
                M3 S = {{{1.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, {0.f, 0.f, 1.f}}};
                S.m[0][0] = a.scale_modifier * s_cov[3 * t];
                S.m[1][1] = a.scale_modifier * s_cov[3 * t + 1];
                S.m[2][2] = a.scale_modifier * s_cov[3 * t + 2];
                // The nine rotation entries share products (x*y, r*z, ...).  For shared products
                // the compiler leaves mul/add pairs unfused in PTX and ptxas then picks which
                // product to fold into an FFMA; that choice is context dependent, so it is
                // pinned here to what the reference build does (read from its SASS): the
                // r-product is the rounded addend for xy/yz, the xz product for the xz pair.
                const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
                const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
                const float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
                const float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
                const float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
                const float yy_p_zz = __fadd_rn(yy, zz);
                const float xx_p_zz = __fmaf_rn(x, x, zz), xx_p_yy = __fmaf_rn(x, x, yy);
                const M3 R = {{{1.f - 2.f * yy_p_zz, 2.f * xy_m_rz, 2.f * xz_p_ry},
                               {2.f * xy_p_rz, 1.f - 2.f * xx_p_zz, 2.f * yz_m_rx},
                               {2.f * xz_m_ry, 2.f * yz_p_rx, 1.f - 2.f * xx_p_yy}}};
                const M3 Mm = mul(S, R);
                const M3 Sig = mul(transpose(Mm), Mm);
                cov6[0] = Sig.m[0][0]; cov6[1] = Sig.m[0][1]; cov6[2] = Sig.m[0][2];
                cov6[3] = Sig.m[1][1]; cov6[4] = Sig.m[1][2]; cov6[5] = Sig.m[2][2];
                cov_valid = true;
            }

            // EWA projection of the 3D covariance (forward.cu:74-113)
            float3 tv = pv;  // same transform, same bits
            const float limx = 1.3f * a.tan_fovx;
            const float limy = 1.3f * a.tan_fovy;
            const float txtz = tv.x / tv.z;
            const float tytz = tv.y / tv.z;
            tv.x = min(limx, max(-limx, txtz)) * tv.z;
            tv.y = min(limy, max(-limy, tytz)) * tv.z;
            const M3 J = {{{a.focal_x / tv.z, 0.0f, -(a.focal_x * tv.x) / (tv.z * tv.z)},
                           {0.0f, a.focal_y / tv.z, -(a.focal_y * tv.y) / (tv.z * tv.z)},
                           {0.f, 0.f, 0.f}}};
            const float* vm = a.view;
            const M3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
            const M3 T = mul(Wm, J);
            const M3 Vrk = {{{cov6[0], cov6[1], cov6[2]}, {cov6[1], cov6[3], cov6[4]}, {cov6[2], cov6[4], cov6[5]}}};
            M3 c2 = mul(mul(transpose(T), transpose(Vrk)), T);
            c2.m[0][0] += 0.3f;
            c2.m[1][1] += 0.3f;
            const float cx = c2.m[0][0], cy = c2.m[0][1], cz = c2.m[1][1];

            // same pinning as above: det = fma(cx, cz, -(cy*cy)), disc = fma(mid, mid, -det),
            // lambda = mid +- sqrt(disc) with mid rounded (reference SASS, forward.cu:216-230)
            const float det = __fmaf_rn(cx, cz, -__fmul_rn(cy, cy));
            if (det != 0.0f) {
                const float det_inv = 1.f / det;
                const float3 conic = make_float3(cz * det_inv, -cy * det_inv, cx * det_inv);
                const float mid = __fmul_rn(0.5f, __fadd_rn(cx, cz));
                const float disc = sqrtf(max(0.1f, __fmaf_rn(mid, mid, -det)));
                const float lambda1 = __fadd_rn(mid, disc);
                const float lambda2 = __fsub_rn(mid, disc);
                const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
                const float2 pix = make_float2(ndc_to_pix(proj.x, a.W), ndc_to_pix(proj.y, a.H));
                uint2 rmin, rmax;
                tile_rect(pix, (int)my_radius, a.grid_x, a.grid_y, rmin, rmax);
                const uint32_t area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
                // A NaN covariance (NaN / infinite scale) gives radius (int)NaN = 0 with a one-tile rectangle.  The
                // reference counts that tile (forward.cu:240-254) but duplicateWithKeys skips radius 0
                // (rasterizer_impl.cu:84), so its key slot keeps whatever the buffer held; here such a Gaussian is
                // culled.  A finite covariance has radius >= 3 (the 0.3 added to the diagonal): nothing else changes.
                if (area != 0 && (int)my_radius > 0) {
                    if (!have_col) {
                        const float* shp = sh_staged ? (s_col + 3 * t) : (a.shs + (size_t)idx * a.M * 3);
                        V3 res = kSH0 * V3{shp[0], shp[1], shp[2]};
                        if (a.D > 0) res = sh_color_hi(a.D, res, p, a.campos, shp);
                        res.x += 0.5f; res.y += 0.5f; res.z += 0.5f;
                        cl0 = res.x < 0; cl1 = res.y < 0; cl2 = res.z < 0;
                        col3[0] = cl0 ? 0.0f : res.x; col3[1] = cl1 ? 0.0f : res.y; col3[2] = cl2 ? 0.0f : res.z;
                        col_valid = true;
                    }
                    depths[idx] = pv.z;
                    means2D[idx] = pix;
                    conic_opacity[idx] = make_float4(conic.x, conic.y, conic.z, __ldg(a.opacities + idx));
                    my_radius_i = (int)my_radius;
                    my_tiles = area;
                    my_rmin = rmin; my_rmax = rmax;
                }
            }
        }
        radii[idx] = my_radius_i;
        tiles_touched[idx] = my_tiles;
    }

    // per-tile list lengths for tile_binning.cu.  A frame of small splats (the random-init decoder: 1.6 tiles per
    // Gaussian) adds one global atomic per (splat, tile).  A frame of large splats (trained-like: 13 tiles per Gaussian,
    // 6.4 M instances on 1024 counters) would serialise on those counters: there the CTA counts into a shared-memory
    // histogram first and flushes one atomic per non-empty tile (~3x fewer, each carrying a count).  The choice is per
    // CTA, from its own splats.
    if (a.tile_counts) {
        // A rectangle of kRectDiffMin tiles or more is not counted tile by tile: it adds +1 / -1 at its four corners to a
        // 2-D difference array (word 3 of the tile records) and tile_scan_kernel turns that array into per-tile counts with
        // one 2-D prefix sum over the tile grid — O(1) per splat instead of O(area).  The 2048^2 / 2 M frame touches 140
        // tiles per Gaussian: 2.8e8 counter updates (0.9 of the kernel's 1.0 ms) become 8e6.
        const bool big = a.diff_ok && my_tiles >= kRectDiffMin;
        if (big) {
            uint32_t* D = a.tile_counts + 3;
            const uint32_t gx = (uint32_t)a.grid_x;
            atomicAdd(D + (size_t)(my_rmin.y * gx + my_rmin.x) * 32, 1u);
            if (my_rmax.x < gx) atomicAdd(D + (size_t)(my_rmin.y * gx + my_rmax.x) * 32, 0xffffffffu);
            if (my_rmax.y < (uint32_t)a.grid_y) {
                atomicAdd(D + (size_t)(my_rmax.y * gx + my_rmin.x) * 32, 0xffffffffu);
                if (my_rmax.x < gx) atomicAdd(D + (size_t)(my_rmax.y * gx + my_rmax.x) * 32, 1u);
            }
            *a.diff_used = 1u;          // same value from everyone
            my_rmax.y = my_rmin.y;      // nothing left for the loops below
            my_tiles = 0;
        }
        const bool dense = a.hist_tiles > 0 && __syncthreads_count(my_tiles > 4u) >= kThreads / 2;
        if (dense) {
            for (int i = t; i < a.hist_tiles; i += kThreads) s_cnt[i] = 0u;
            __syncthreads();
            for (uint32_t ty = my_rmin.y; ty < my_rmax.y; ++ty)
                for (uint32_t tx = my_rmin.x; tx < my_rmax.x; ++tx) atomicAdd(&s_cnt[ty * a.grid_x + tx], 1u);
            __syncthreads();
            for (int i = t; i < a.hist_tiles; i += kThreads) {
                const uint32_t c = s_cnt[i];
                if (c) atomicAdd(a.tile_counts + (size_t)i * 32, c);
            }
        } else {
            for (uint32_t ty = my_rmin.y; ty < my_rmax.y; ++ty)
                for (uint32_t tx = my_rmin.x; tx < my_rmax.x; ++tx) atomicAdd(a.tile_counts + (size_t)(ty * a.grid_x + tx) * 32, 1u);
        }
    }

    // staged, coalesced write-back of the stride-3/6 outputs. Lanes that did not
    // produce a value write zeros (the reference leaves stale bytes there; nothing reads them).
    __syncthreads();
    if (t < n_here) {
        if (!have_cov) {
#pragma unroll
            for (int k = 0; k < 6; ++k) s_cov[6 * t + k] = cov_valid ? cov6[k] : 0.f;
        }
        if (!have_col) {
            s_col[3 * t] = col_valid ? col3[0] : 0.f;
            s_col[3 * t + 1] = col_valid ? col3[1] : 0.f;
            s_col[3 * t + 2] = col_valid ? col3[2] : 0.f;
            s_clamp[3 * t] = cl0; s_clamp[3 * t + 1] = cl1; s_clamp[3 * t + 2] = cl2;
        }
    }
    __syncthreads();
    if (!have_cov) block_store(cov3D_out + (size_t)base * 6, s_cov, n_here * 6);
    if (!have_col) {
        block_store(rgb_out + (size_t)base * 3, s_col, n_here * 3);
        // base*3 is a multiple of 4 bytes*... (kThreads*3 = 768) so word stores are aligned
        const int nwords = (n_here * 3) >> 2;
        uint32_t* dw = reinterpret_cast<uint32_t*>(clamped_out + (size_t)base * 3);
        const uint32_t* sw = reinterpret_cast<const uint32_t*>(s_clamp);
        for (int i = t; i < nwords; i += kThreads) dw[i] = sw[i];
        for (int i = (nwords << 2) + t; i < n_here * 3; i += kThreads)
            clamped_out[(size_t)base * 3 + i] = s_clamp[i];
    }
}

// ---- single-pass inclusive scan (decoupled look-back) ---------------------
// status word: [63:62] flag (0 empty, 1 tile aggregate, 2 inclusive prefix), [31:0] value
constexpr uint64_t kFlagAgg = 1ull << 62;
constexpr uint64_t kFlagPre = 2ull << 62;

__global__ void __launch_bounds__(kScanThreads)
scan_kernel(int P, const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
            uint32_t* __restrict__ counters, uint64_t* __restrict__ status)
{
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(&counters[2], 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int base = tile * kScanTile;
    if (base >= P) return;

    // blocked arrangement: thread t owns items [t*8, t*8+8)
    uint32_t v[kScanItems];
    const int i0 = base + threadIdx.x * kScanItems;
    if (i0 + kScanItems <= P) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(in + i0));
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(in + i0) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) v[k] = (i0 + k < P) ? __ldg(in + i0 + k) : 0u;
    }
    uint32_t tsum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) { tsum += v[k]; v[k] = tsum; }

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, block_sum = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
        const uint32_t ws = s_warp[w];
        if (w < warp) warp_off += ws;
        block_sum += ws;
    }
    const uint32_t thread_excl = warp_off + incl - tsum;

    if (warp == 0) {
        // decoupled look-back, 32 predecessors per step: lane i inspects tile (look - i); everything up to the nearest
        // tile that already knows its inclusive prefix must at least have published its aggregate
        uint32_t excl = 0;
        if (lane == 0)
            atomicExch(reinterpret_cast<unsigned long long*>(&status[tile]), (tile == 0 ? (2ull << 62) : (1ull << 62)) | block_sum);
        if (tile != 0) {
            int look = (int)tile - 1;
            while (true) {
                const int idx = look - lane;
                const uint64_t w = idx >= 0 ? *reinterpret_cast<volatile uint64_t*>(&status[idx]) : (2ull << 62);   // before tile 0: prefix 0
                const uint32_t f = (uint32_t)(w >> 62);
                const uint32_t pre = __ballot_sync(0xffffffffu, f == 2), empty = __ballot_sync(0xffffffffu, f == 0);
                const int first_pre = pre ? __ffs(pre) - 1 : 32;
                const uint32_t need = first_pre >= 31 ? 0xffffffffu : ((2u << first_pre) - 1u);
                if (empty & need) continue;                       // a tile in the window has published nothing yet
                uint32_t part = (lane <= first_pre) ? (uint32_t)w : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                excl += part;
                if (first_pre < 32) break;
                look -= 32;
            }
            if (lane == 0)
                atomicExch(reinterpret_cast<unsigned long long*>(&status[tile]), kFlagPre | (uint64_t)(excl + block_sum));
        }
        if (lane == 0) {
            s_prefix = excl;
            if (base + kScanTile >= P) counters[0] = excl + block_sum;  // last tile: R
        }
    }
    __syncthreads();
    const uint32_t add = s_prefix + thread_excl;
    if (i0 + kScanItems <= P) {
        uint4 a = make_uint4(v[0] + add, v[1] + add, v[2] + add, v[3] + add);
        uint4 b = make_uint4(v[4] + add, v[5] + add, v[6] + add, v[7] + add);
        reinterpret_cast<uint4*>(out + i0)[0] = a;
        reinterpret_cast<uint4*>(out + i0)[1] = b;
    } else {
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
            if (i0 + k < P) out[i0 + k] = v[k] + add;
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ pts, const float* __restrict__ view,
                                    const float* __restrict__ proj, uint8_t* __restrict__ present)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p = make_float3(pts[3 * idx], pts[3 * idx + 1], pts[3 * idx + 2]);
    const float3 pv = xform4x3(p, view);
    (void)proj;  // the reference projects too but only tests view-space depth (auxiliary.h:154)
    present[idx] = !(pv.z <= 0.2f);
}

}  // namespace

int launch_preprocess(cudaStream_t s, const PreprocessArgs& a, const GeomState& g, int* radii)
{
    const int blocks = (a.P + kThreads - 1) / kThreads;
    PreprocessArgs b = a;
    const int tiles = a.grid_x * a.grid_y;
    b.diff_ok = (a.tile_counts && tiles <= kTileDiffMaxTiles) ? 1 : 0;
    // CTA-local histogram of the small rectangles: 16 KB at most (with the large ones on the difference array the
    // 64 KB histogram a 2048^2 image used to get only cost occupancy)
    b.hist_tiles = (a.tile_counts && tiles <= (b.diff_ok ? 4096 : 16384)) ? tiles : 0;
    if (b.hist_tiles > 8192)
        GGD_CUDA_CHECK(cudaFuncSetAttribute(preprocess_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    preprocess_kernel<<<blocks, kThreads, sizeof(uint32_t) * (size_t)b.hist_tiles, s>>>(b, g.depths, radii, g.means2D, g.cov3D,
                                                  g.conic_opacity, g.rgb, g.tiles_touched, g.clamped);
    GGD_LAUNCH_CHECK("preprocess_kernel");
    return GGD_OK;
}

int launch_scan(cudaStream_t s, int P, const GeomState& g)
{
    const int tiles = (P + kScanTile - 1) / kScanTile;
    scan_kernel<<<tiles, kScanThreads, 0, s>>>(P, g.tiles_touched, g.point_offsets, g.counters,
                                               g.scan_status);
    GGD_LAUNCH_CHECK("scan_kernel");
    return GGD_OK;
}

int launch_mark_visible(cudaStream_t s, int P, const float* means3D, const float* view,
                        const float* proj, uint8_t* present)
{
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, proj, present);
    GGD_LAUNCH_CHECK("mark_visible_kernel");
    return GGD_OK;
}

}  // namespace ggd
