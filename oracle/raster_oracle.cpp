// TEST INFRASTRUCTURE — CPU restatement of the reference rasteriser. NOT the product.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load this library.  It restates, in plain C++17 + OpenMP,
// the algorithm of (DGR = gaussian_splatting/submodules/diff-gaussian-rasterization):
//   DGR/cuda_rasterizer/auxiliary.h:41-164   ndc2Pix (double), getRect, transforms, in_frustum
//   DGR/cuda_rasterizer/forward.cu:20-71     SH -> RGB
//   DGR/cuda_rasterizer/forward.cu:74-113    EWA 2D covariance
//   DGR/cuda_rasterizer/forward.cu:118-152   3D covariance from scale/rotation
//   DGR/cuda_rasterizer/forward.cu:155-256   per-Gaussian preprocess
//   DGR/cuda_rasterizer/rasterizer_impl.cu:35-50   getHigherMsb
//   DGR/cuda_rasterizer/rasterizer_impl.cu:70-111  key layout + emission order
//   DGR/cuda_rasterizer/rasterizer_impl.cu:116-138 tile ranges
//   DGR/cuda_rasterizer/forward.cu:261-374   blending
//   DGR/cuda_rasterizer/backward.cu:20-557   all gradients
// cub::DeviceScan::InclusiveSum / DeviceRadixSort::SortPairs (CUDA toolkit, not
// vendored) are restated by their contract: exact uint32 prefix sum; stable
// sort on the low (32+bit) key bits.
//
// PARITY STATUS: the reference ships no tests or golden vectors for this path
// (SURVEY.md §4), so this restatement is pinned against OUTPUTS OF THE
// REFERENCE ITSELF: tests/golden/raster_*.npz were produced on a B200 by
// oracle/_ref/libggd_ref_cuda.so (the unmodified reference compiled from
// /root/reference) with tests/golden/make_raster_golden.py.  Floats here are
// tolerance-grade (host libm expf, gcc FMA contraction); the integer outputs
// agree except where a float lands within an ulp of a rounding boundary —
// bit-exactness of the product is certified against the reference CUDA build.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <parallel/algorithm>
#include <omp.h>

namespace {

constexpr int BX = 16, BY = 16, BS = BX * BY;

struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };
struct m3 { float c[3][3]; };  // column-major c[col][row]

inline m3 mul(const m3& a, const m3& b)
{
    m3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
    return r;
}
inline m3 tr(const m3& a)
{
    m3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[i][j];
    return r;
}
inline f3 xf4x3(f3 p, const float* m)
{
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
inline f4 xf4x4(f3 p, const float* m)
{
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}
inline f3 xfvec4x3T(f3 p, const float* m)
{
    return {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
            m[8] * p.x + m[9] * p.y + m[10] * p.z};
}
inline float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }
inline int f2i_sat(float f)  // cvt.rzi.s32.f32 saturates; a host cast is UB out of range
{
    if (std::isnan(f)) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int)f;
}
inline void get_rect(float px, float py, int r, uint32_t gx, uint32_t gy, uint32_t* mn, uint32_t* mx)
{
    mn[0] = std::min(gx, (uint32_t)std::max(0, f2i_sat((px - r) / BX)));
    mn[1] = std::min(gy, (uint32_t)std::max(0, f2i_sat((py - r) / BY)));
    mx[0] = std::min(gx, (uint32_t)std::max(0, f2i_sat((px + r + BX - 1) / BX)));
    mx[1] = std::min(gy, (uint32_t)std::max(0, f2i_sat((py + r + BY - 1) / BY)));
}
uint32_t higher_msb(uint32_t n)
{
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
const float C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f};
const float C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                    -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
inline f3 operator*(float s, f3 v) { return {s * v.x, s * v.y, s * v.z}; }
inline f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct State {
    int P = 0, W = 0, H = 0, D = 0, M = 0;
    uint32_t gx = 0, gy = 0, R = 0;
    std::vector<float> depths, means2D, cov3D, conic_opacity, rgb, final_T, out_color;
    std::vector<int32_t> radii;
    std::vector<uint32_t> tiles_touched, point_offsets, vals_unsorted, vals, n_contrib, ranges;
    std::vector<uint8_t> clamped;
    std::vector<uint64_t> keys_unsorted, keys;
    // gradients
    std::vector<float> dmean2D, dconic, dopacity, dcolor, dmean3D, dcov3D, dsh, dscale, drot;
};

void cov3d_from(const float* sc, float mod, const float* q, float* out)
{
    m3 S = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
    S.c[0][0] = mod * sc[0]; S.c[1][1] = mod * sc[1]; S.c[2][2] = mod * sc[2];
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    m3 Rm = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
              {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
              {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    m3 Mm = mul(S, Rm);
    m3 Sg = mul(tr(Mm), Mm);
    out[0] = Sg.c[0][0]; out[1] = Sg.c[0][1]; out[2] = Sg.c[0][2];
    out[3] = Sg.c[1][1]; out[4] = Sg.c[1][2]; out[5] = Sg.c[2][2];
}

struct Cov2DCtx { m3 T, Vrk; f3 t; float txtz, tytz, limx, limy; };
inline f3 cov2d_from(f3 mean, float fx, float fy, float tfx, float tfy, const float* cov3D, const float* vm, Cov2DCtx* ctx = nullptr)
{
    f3 t = xf4x3(mean, vm);
    const float limx = 1.3f * tfx, limy = 1.3f * tfy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = std::min(limx, std::max(-limx, txtz)) * t.z;
    t.y = std::min(limy, std::max(-limy, tytz)) * t.z;
    m3 J = {{{fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z)}, {0.0f, fy / t.z, -(fy * t.y) / (t.z * t.z)}, {0, 0, 0}}};
    m3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
    m3 T = mul(Wm, J);
    m3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
    m3 cov = mul(mul(tr(T), tr(Vrk)), T);
    if (ctx) { ctx->T = T; ctx->Vrk = Vrk; ctx->t = t; ctx->txtz = txtz; ctx->tytz = tytz; ctx->limx = limx; ctx->limy = limy; }
    cov.c[0][0] += 0.3f;
    cov.c[1][1] += 0.3f;
    return {cov.c[0][0], cov.c[0][1], cov.c[1][1]};
}

f3 sh_to_rgb(int idx, int deg, int M, const float* means, const float* campos, const float* shs, uint8_t* clamped)
{
    const f3* sh = reinterpret_cast<const f3*>(shs) + (size_t)idx * M;
    f3 result = C0 * sh[0];
    if (deg > 0) {
        f3 pos = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
        f3 dir = pos - f3{campos[0], campos[1], campos[2]};
        const float len = std::sqrt(dot(dir, dir));
        dir = {dir.x / len, dir.y / len, dir.z / len};
        const float x = dir.x, y = dir.y, z = dir.z;
        result = result - C1 * y * sh[1] + C1 * z * sh[2] - C1 * x * sh[3];
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            result = result + C2[0] * xy * sh[4] + C2[1] * yz * sh[5] + C2[2] * (2.0f * zz - xx - yy) * sh[6] +
                     C2[3] * xz * sh[7] + C2[4] * (xx - yy) * sh[8];
            if (deg > 2) {
                result = result + C3[0] * y * (3.0f * xx - yy) * sh[9] + C3[1] * xy * z * sh[10] +
                         C3[2] * y * (4.0f * zz - xx - yy) * sh[11] + C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                         C3[4] * x * (4.0f * zz - xx - yy) * sh[13] + C3[5] * z * (xx - yy) * sh[14] +
                         C3[6] * x * (xx - 3.0f * yy) * sh[15];
            }
        }
    }
    result.x += 0.5f; result.y += 0.5f; result.z += 0.5f;
    clamped[3 * idx + 0] = result.x < 0; clamped[3 * idx + 1] = result.y < 0; clamped[3 * idx + 2] = result.z < 0;
    return {result.x < 0 ? 0.f : result.x, result.y < 0 ? 0.f : result.y, result.z < 0 ? 0.f : result.z};
}

}  // namespace

extern "C" {

void* oracle_create() { return new State(); }
void oracle_destroy(void* h) { delete static_cast<State*>(h); }
uint32_t oracle_higher_msb(uint32_t n) { return higher_msb(n); }

// Forward: returns R (num_rendered). All intermediate state is kept in the handle.
int oracle_forward(void* h, int P, int D, int M, const float* bg, int W, int H, const float* means3D,
                   const float* shs, const float* colors_precomp, const float* opacities,
                   const float* scales, float scale_modifier, const float* rotations,
                   const float* cov3D_precomp, const float* view, const float* proj,
                   const float* campos, float tan_fovx, float tan_fovy, int stop_after_preprocess)
{
    State& s = *static_cast<State*>(h);
    s.P = P; s.W = W; s.H = H; s.D = D; s.M = M;
    s.gx = (W + BX - 1) / BX; s.gy = (H + BY - 1) / BY;
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
    s.depths.assign(P, 0.f); s.means2D.assign(2 * (size_t)P, 0.f); s.cov3D.assign(6 * (size_t)P, 0.f);
    s.conic_opacity.assign(4 * (size_t)P, 0.f); s.rgb.assign(3 * (size_t)P, 0.f); s.radii.assign(P, 0);
    s.tiles_touched.assign(P, 0); s.point_offsets.assign(P, 0); s.clamped.assign(3 * (size_t)P, 0);

#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        f3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
        f4 hom = xf4x4(p, proj);
        float pw = 1.0f / (hom.w + 0.0000001f);
        f3 pp = {hom.x * pw, hom.y * pw, hom.z * pw};
        f3 pv = xf4x3(p, view);
        if (pv.z <= 0.2f) continue;
        const float* c3;
        if (cov3D_precomp) c3 = cov3D_precomp + 6 * (size_t)idx;
        else { cov3d_from(scales + 3 * (size_t)idx, scale_modifier, rotations + 4 * (size_t)idx, &s.cov3D[6 * (size_t)idx]); c3 = &s.cov3D[6 * (size_t)idx]; }
        f3 cov = cov2d_from(p, focal_x, focal_y, tan_fovx, tan_fovy, c3, view);
        float det = (cov.x * cov.z - cov.y * cov.y);
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        f3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
        float mid = 0.5f * (cov.x + cov.z);
        float l1 = mid + std::sqrt(std::max(0.1f, mid * mid - det));
        float l2 = mid - std::sqrt(std::max(0.1f, mid * mid - det));
        float my_radius = std::ceil(3.f * std::sqrt(std::max(l1, l2)));
        float pix[2] = {ndc2pix(pp.x, W), ndc2pix(pp.y, H)};
        uint32_t mn[2], mx[2];
        get_rect(pix[0], pix[1], f2i_sat(my_radius), s.gx, s.gy, mn, mx);
        if ((mx[0] - mn[0]) * (mx[1] - mn[1]) == 0) continue;
        if (!colors_precomp) {
            f3 c = sh_to_rgb(idx, D, M, means3D, campos, shs, s.clamped.data());
            s.rgb[3 * (size_t)idx] = c.x; s.rgb[3 * (size_t)idx + 1] = c.y; s.rgb[3 * (size_t)idx + 2] = c.z;
        }
        s.depths[idx] = pv.z;
        s.radii[idx] = f2i_sat(my_radius);
        s.means2D[2 * (size_t)idx] = pix[0]; s.means2D[2 * (size_t)idx + 1] = pix[1];
        float* co = &s.conic_opacity[4 * (size_t)idx];
        co[0] = conic.x; co[1] = conic.y; co[2] = conic.z; co[3] = opacities[idx];
        s.tiles_touched[idx] = (mx[1] - mn[1]) * (mx[0] - mn[0]);
    }
    uint32_t run = 0;
    for (int i = 0; i < P; ++i) { run += s.tiles_touched[i]; s.point_offsets[i] = run; }
    s.R = run;
    if (stop_after_preprocess) return (int)s.R;

    const uint32_t R = s.R;
    s.keys_unsorted.assign(R, 0); s.vals_unsorted.assign(R, 0);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int idx = 0; idx < P; ++idx) {
        if (s.radii[idx] <= 0) continue;
        uint32_t off = idx == 0 ? 0 : s.point_offsets[idx - 1];
        uint32_t mn[2], mx[2];
        get_rect(s.means2D[2 * (size_t)idx], s.means2D[2 * (size_t)idx + 1], s.radii[idx], s.gx, s.gy, mn, mx);
        uint32_t dbits;
        std::memcpy(&dbits, &s.depths[idx], 4);
        for (uint32_t y = mn[1]; y < mx[1]; ++y)
            for (uint32_t x = mn[0]; x < mx[0]; ++x) {
                uint64_t key = y * s.gx + x;
                key <<= 32;
                key |= dbits;
                s.keys_unsorted[off] = key;
                s.vals_unsorted[off] = (uint32_t)idx;
                ++off;
            }
    }
    const uint32_t bit = higher_msb(s.gx * s.gy);
    const uint64_t mask = (32 + bit >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1);
    std::vector<std::pair<uint64_t, uint32_t>> kv(R);
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)R; ++i) kv[i] = {s.keys_unsorted[i], s.vals_unsorted[i]};
    __gnu_parallel::stable_sort(kv.begin(), kv.end(), [mask](const auto& a, const auto& b) { return (a.first & mask) < (b.first & mask); });
    s.keys.resize(R); s.vals.resize(R);
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)R; ++i) { s.keys[i] = kv[i].first; s.vals[i] = kv[i].second; }

    const uint32_t tiles = s.gx * s.gy;
    s.ranges.assign(2 * (size_t)tiles, 0);
    for (uint32_t i = 0; i < R; ++i) {
        uint32_t cur = (uint32_t)(s.keys[i] >> 32);
        if (i == 0) s.ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(s.keys[i - 1] >> 32);
            if (cur != prev) { s.ranges[2 * prev + 1] = i; s.ranges[2 * cur] = i; }
        }
        if (i == R - 1) s.ranges[2 * cur + 1] = R;
    }

    const float* feat = colors_precomp ? colors_precomp : s.rgb.data();
    s.final_T.assign((size_t)W * H, 0.f); s.n_contrib.assign((size_t)W * H, 0); s.out_color.assign(3 * (size_t)W * H, 0.f);
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < (int)tiles; ++tile) {
        const uint32_t tx = tile % s.gx, ty = tile / s.gx;
        const uint32_t r0 = s.ranges[2 * tile], r1 = s.ranges[2 * tile + 1];
        for (int ly = 0; ly < BY; ++ly)
            for (int lx = 0; lx < BX; ++lx) {
                const uint32_t px = tx * BX + lx, py = ty * BY + ly;
                if (px >= (uint32_t)W || py >= (uint32_t)H) continue;
                const float fx = (float)px, fy = (float)py;
                float T = 1.0f, C[3] = {0, 0, 0};
                uint32_t contributor = 0, last = 0;
                for (uint32_t k = r0; k < r1; ++k) {
                    contributor++;
                    const uint32_t id = s.vals[k];
                    const float dx = s.means2D[2 * (size_t)id] - fx, dy = s.means2D[2 * (size_t)id + 1] - fy;
                    const float* co = &s.conic_opacity[4 * (size_t)id];
                    const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                    if (power > 0.0f) continue;
                    const float alpha = std::min(0.99f, co[3] * std::exp(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break;  // `done`: nothing after this changes the pixel
                    for (int ch = 0; ch < 3; ++ch) C[ch] += feat[3 * (size_t)id + ch] * alpha * T;
                    T = test_T;
                    last = contributor;
                }
                const size_t pid = (size_t)W * py + px;
                s.final_T[pid] = T;
                s.n_contrib[pid] = last;
                for (int ch = 0; ch < 3; ++ch) s.out_color[(size_t)ch * H * W + pid] = C[ch] + T * bg[ch];
            }
    }
    return (int)R;
}

// Backward (DGR/cuda_rasterizer/backward.cu): call after oracle_forward on the same handle.
int oracle_backward(void* h, const float* bg, const float* means3D, const float* shs,
                    const float* colors_precomp, const float* scales, float scale_modifier,
                    const float* rotations, const float* cov3D_precomp, const float* view,
                    const float* proj, const float* campos, float tan_fovx, float tan_fovy,
                    const float* dL_dpix)
{
    State& s = *static_cast<State*>(h);
    const int P = s.P, W = s.W, H = s.H, D = s.D, M = s.M;
    const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
    s.dmean2D.assign(3 * (size_t)P, 0.f); s.dconic.assign(4 * (size_t)P, 0.f); s.dopacity.assign(P, 0.f);
    s.dcolor.assign(3 * (size_t)P, 0.f); s.dmean3D.assign(3 * (size_t)P, 0.f); s.dcov3D.assign(6 * (size_t)P, 0.f);
    s.dsh.assign(3 * (size_t)P * std::max(M, 0), 0.f); s.dscale.assign(3 * (size_t)P, 0.f); s.drot.assign(4 * (size_t)P, 0.f);
    const float* colors = colors_precomp ? colors_precomp : s.rgb.data();
    const uint32_t tiles = s.gx * s.gy;
    // per-pixel replay, back to front (backward.cu:399-557); accumulate in double to make the
    // oracle's sums order-independent to ~1e-16 relative.
    std::vector<double> a_m2d(2 * (size_t)P, 0.0), a_con(3 * (size_t)P, 0.0), a_op(P, 0.0), a_col(3 * (size_t)P, 0.0);
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < (int)tiles; ++tile) {
        const uint32_t tx = tile % s.gx, ty = tile / s.gx;
        const uint32_t r0 = s.ranges[2 * tile], r1 = s.ranges[2 * tile + 1];
        for (int ly = 0; ly < BY; ++ly)
            for (int lx = 0; lx < BX; ++lx) {
                const uint32_t px = tx * BX + lx, py = ty * BY + ly;
                if (px >= (uint32_t)W || py >= (uint32_t)H) continue;
                const size_t pid = (size_t)W * py + px;
                const float fx = (float)px, fy = (float)py;
                const float T_final = s.final_T[pid];
                float T = T_final;
                uint32_t contributor = r1 - r0;
                const uint32_t last_contributor = s.n_contrib[pid];
                float accum_rec[3] = {0, 0, 0}, dL_dpixel[3], last_color[3] = {0, 0, 0};
                for (int ch = 0; ch < 3; ++ch) dL_dpixel[ch] = dL_dpix[(size_t)ch * H * W + pid];
                float last_alpha = 0;
                for (uint32_t k = r1; k-- > r0;) {
                    contributor--;
                    if (contributor >= last_contributor) continue;
                    const uint32_t id = s.vals[k];
                    const float dx = s.means2D[2 * (size_t)id] - fx, dy = s.means2D[2 * (size_t)id + 1] - fy;
                    const float* co = &s.conic_opacity[4 * (size_t)id];
                    const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                    if (power > 0.0f) continue;
                    const float G = std::exp(power);
                    const float alpha = std::min(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    for (int ch = 0; ch < 3; ++ch) {
                        const float c = colors[3 * (size_t)id + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        const float dL_dchannel = dL_dpixel[ch];
                        dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
#pragma omp atomic
                        a_col[3 * (size_t)id + ch] += (double)(dchannel_dcolor * dL_dchannel);
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    float bg_dot = 0;
                    for (int i = 0; i < 3; ++i) bg_dot += bg[i] * dL_dpixel[i];
                    dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                    const float dL_dG = co[3] * dL_dalpha;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    const float dG_ddely = -gdy * co[2] - gdx * co[1];
#pragma omp atomic
                    a_m2d[2 * (size_t)id] += (double)(dL_dG * dG_ddelx * ddelx_dx);
#pragma omp atomic
                    a_m2d[2 * (size_t)id + 1] += (double)(dL_dG * dG_ddely * ddely_dy);
#pragma omp atomic
                    a_con[3 * (size_t)id] += (double)(-0.5f * gdx * dx * dL_dG);
#pragma omp atomic
                    a_con[3 * (size_t)id + 1] += (double)(-0.5f * gdx * dy * dL_dG);
#pragma omp atomic
                    a_con[3 * (size_t)id + 2] += (double)(-0.5f * gdy * dy * dL_dG);
#pragma omp atomic
                    a_op[id] += (double)(G * dL_dalpha);
                }
            }
    }
    for (int i = 0; i < P; ++i) {
        s.dmean2D[3 * (size_t)i] = (float)a_m2d[2 * (size_t)i]; s.dmean2D[3 * (size_t)i + 1] = (float)a_m2d[2 * (size_t)i + 1];
        s.dconic[4 * (size_t)i] = (float)a_con[3 * (size_t)i]; s.dconic[4 * (size_t)i + 1] = (float)a_con[3 * (size_t)i + 1];
        s.dconic[4 * (size_t)i + 3] = (float)a_con[3 * (size_t)i + 2];
        s.dopacity[i] = (float)a_op[i];
        for (int ch = 0; ch < 3; ++ch) s.dcolor[3 * (size_t)i + ch] = (float)a_col[3 * (size_t)i + ch];
    }

#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; ++idx) {
        if (!(s.radii[idx] > 0)) continue;
        // ---- conic -> cov2D -> cov3D / mean (backward.cu:144-274)
        const float* cov3D = cov3D_precomp ? cov3D_precomp + 6 * (size_t)idx : &s.cov3D[6 * (size_t)idx];
        f3 mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
        f3 dL_dconic = {s.dconic[4 * (size_t)idx], s.dconic[4 * (size_t)idx + 1], s.dconic[4 * (size_t)idx + 3]};
        Cov2DCtx cx;
        f3 c2 = cov2d_from(mean, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, view, &cx);
        const m3& T = cx.T; const m3& Vrk = cx.Vrk; const f3 t = cx.t;
        const float x_grad_mul = (cx.txtz < -cx.limx || cx.txtz > cx.limx) ? 0 : 1;
        const float y_grad_mul = (cx.tytz < -cx.limy || cx.tytz > cx.limy) ? 0 : 1;
        const float a = c2.x, b = c2.y, c = c2.z;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcov = &s.dcov3D[6 * (size_t)idx];
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
            dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
            dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
            dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
            dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
            dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
            dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
            dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
            dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
        } else {
            for (int i = 0; i < 6; ++i) dcov[i] = 0;
        }
        const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                              (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
        const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                              (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
        const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                              (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
        const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                              (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
        const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                              (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
        const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                              (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;
        const float* vm = view;
        m3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
        const float dL_dJ00 = Wm.c[0][0] * dL_dT00 + Wm.c[0][1] * dL_dT01 + Wm.c[0][2] * dL_dT02;
        const float dL_dJ02 = Wm.c[2][0] * dL_dT00 + Wm.c[2][1] * dL_dT01 + Wm.c[2][2] * dL_dT02;
        const float dL_dJ11 = Wm.c[1][0] * dL_dT10 + Wm.c[1][1] * dL_dT11 + Wm.c[1][2] * dL_dT12;
        const float dL_dJ12 = Wm.c[2][0] * dL_dT10 + Wm.c[2][1] * dL_dT11 + Wm.c[2][2] * dL_dT12;
        const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -focal_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -focal_y * tz2 * dL_dJ12;
        const float dL_dtz = -focal_x * tz2 * dL_dJ00 - focal_y * tz2 * dL_dJ11 + (2 * focal_x * t.x) * tz3 * dL_dJ02 + (2 * focal_y * t.y) * tz3 * dL_dJ12;
        f3 dmean = xfvec4x3T({dL_dtx, dL_dty, dL_dtz}, view);

        // ---- mean2D -> mean3D (backward.cu:346-396)
        f4 mh = xf4x4(mean, proj);
        const float m_w = 1.0f / (mh.w + 0.0000001f);
        const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
        const float gx2 = s.dmean2D[3 * (size_t)idx], gy2 = s.dmean2D[3 * (size_t)idx + 1];
        dmean.x += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
        dmean.y += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
        dmean.z += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;

        // ---- SH backward (backward.cu:20-139)
        if (shs) {
            const f3* sh = reinterpret_cast<const f3*>(shs) + (size_t)idx * M;
            f3 dir_orig = mean - f3{campos ? campos[0] : 0.f, campos ? campos[1] : 0.f, campos ? campos[2] : 0.f};
            const float len = std::sqrt(dot(dir_orig, dir_orig));
            f3 dir = {dir_orig.x / len, dir_orig.y / len, dir_orig.z / len};
            f3 dRGB = {s.dcolor[3 * (size_t)idx], s.dcolor[3 * (size_t)idx + 1], s.dcolor[3 * (size_t)idx + 2]};
            dRGB.x *= s.clamped[3 * (size_t)idx] ? 0 : 1; dRGB.y *= s.clamped[3 * (size_t)idx + 1] ? 0 : 1; dRGB.z *= s.clamped[3 * (size_t)idx + 2] ? 0 : 1;
            f3 dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
            const float x = dir.x, y = dir.y, z = dir.z;
            f3* dsh = reinterpret_cast<f3*>(s.dsh.data()) + (size_t)idx * M;
            dsh[0] = C0 * dRGB;
            if (D > 0) {
                dsh[1] = (-C1 * y) * dRGB; dsh[2] = (C1 * z) * dRGB; dsh[3] = (-C1 * x) * dRGB;
                dRGBdx = (-C1) * sh[3]; dRGBdy = (-C1) * sh[1]; dRGBdz = C1 * sh[2];
                if (D > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    dsh[4] = (C2[0] * xy) * dRGB; dsh[5] = (C2[1] * yz) * dRGB; dsh[6] = (C2[2] * (2.f * zz - xx - yy)) * dRGB;
                    dsh[7] = (C2[3] * xz) * dRGB; dsh[8] = (C2[4] * (xx - yy)) * dRGB;
                    dRGBdx = dRGBdx + (C2[0] * y) * sh[4] + (C2[2] * 2.f * -x) * sh[6] + (C2[3] * z) * sh[7] + (C2[4] * 2.f * x) * sh[8];
                    dRGBdy = dRGBdy + (C2[0] * x) * sh[4] + (C2[1] * z) * sh[5] + (C2[2] * 2.f * -y) * sh[6] + (C2[4] * 2.f * -y) * sh[8];
                    dRGBdz = dRGBdz + (C2[1] * y) * sh[5] + (C2[2] * 2.f * 2.f * z) * sh[6] + (C2[3] * x) * sh[7];
                    if (D > 2) {
                        dsh[9] = (C3[0] * y * (3.f * xx - yy)) * dRGB; dsh[10] = (C3[1] * xy * z) * dRGB;
                        dsh[11] = (C3[2] * y * (4.f * zz - xx - yy)) * dRGB; dsh[12] = (C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dRGB;
                        dsh[13] = (C3[4] * x * (4.f * zz - xx - yy)) * dRGB; dsh[14] = (C3[5] * z * (xx - yy)) * dRGB;
                        dsh[15] = (C3[6] * x * (xx - 3.f * yy)) * dRGB;
                        dRGBdx = dRGBdx + (C3[0] * 3.f * 2.f * xy) * sh[9] + (C3[1] * yz) * sh[10] + (C3[2] * -2.f * xy) * sh[11] +
                                 (C3[3] * -3.f * 2.f * xz) * sh[12] + (C3[4] * (-3.f * xx + 4.f * zz - yy)) * sh[13] +
                                 (C3[5] * 2.f * xz) * sh[14] + (C3[6] * 3.f * (xx - yy)) * sh[15];
                        dRGBdy = dRGBdy + (C3[0] * 3.f * (xx - yy)) * sh[9] + (C3[1] * xz) * sh[10] + (C3[2] * (-3.f * yy + 4.f * zz - xx)) * sh[11] +
                                 (C3[3] * -3.f * 2.f * yz) * sh[12] + (C3[4] * -2.f * xy) * sh[13] + (C3[5] * -2.f * yz) * sh[14] +
                                 (C3[6] * -3.f * 2.f * xy) * sh[15];
                        dRGBdz = dRGBdz + (C3[1] * xy) * sh[10] + (C3[2] * 4.f * 2.f * yz) * sh[11] + (C3[3] * 3.f * (2.f * zz - xx - yy)) * sh[12] +
                                 (C3[4] * 4.f * 2.f * xz) * sh[13] + (C3[5] * (xx - yy)) * sh[14];
                    }
                }
                f3 dL_ddir = {dot(dRGBdx, dRGB), dot(dRGBdy, dRGB), dot(dRGBdz, dRGB)};
                const f3 v = dir_orig, dv = dL_ddir;
                const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
                const float inv32 = 1.0f / std::sqrt(sum2 * sum2 * sum2);
                dmean.x += ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv32;
                dmean.y += (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv32;
                dmean.z += (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * inv32;
            }
        }
        s.dmean3D[3 * (size_t)idx] = dmean.x; s.dmean3D[3 * (size_t)idx + 1] = dmean.y; s.dmean3D[3 * (size_t)idx + 2] = dmean.z;

        // ---- cov3D -> scale, rotation (backward.cu:278-341; no quaternion-norm Jacobian)
        if (scales) {
            const float* q = rotations + 4 * (size_t)idx;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            m3 Rm = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                      {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                      {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
            m3 S = {{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};
            const float sx = scale_modifier * scales[3 * (size_t)idx], sy = scale_modifier * scales[3 * (size_t)idx + 1], sz = scale_modifier * scales[3 * (size_t)idx + 2];
            S.c[0][0] = sx; S.c[1][1] = sy; S.c[2][2] = sz;
            m3 Mm = mul(S, Rm);
            const float* g = dcov;
            m3 dSig = {{{g[0], 0.5f * g[1], 0.5f * g[2]}, {0.5f * g[1], g[3], 0.5f * g[4]}, {0.5f * g[2], 0.5f * g[4], g[5]}}};
            m3 M2 = Mm;
            for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) M2.c[j][i] = 2.0f * Mm.c[j][i];
            m3 dL_dM = mul(M2, dSig);
            m3 Rt = tr(Rm), dMt = tr(dL_dM);
            float* ds = &s.dscale[3 * (size_t)idx];
            ds[0] = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
            ds[1] = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
            ds[2] = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
            for (int i = 0; i < 3; ++i) { dMt.c[0][i] *= sx; dMt.c[1][i] *= sy; dMt.c[2][i] *= sz; }
            float* dq = &s.drot[4 * (size_t)idx];
            dq[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
            dq[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
            dq[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
            dq[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
        }
    }
    return 0;
}

// Named accessor: returns pointer, element count, and element size in bytes.
int oracle_get(void* h, const char* name, const void** ptr, size_t* count, int* elem)
{
    State& s = *static_cast<State*>(h);
    const std::string n(name);
#define F(NAME, VEC) if (n == NAME) { *ptr = s.VEC.data(); *count = s.VEC.size(); *elem = (int)sizeof(s.VEC[0]); return 0; }
    F("depths", depths) F("means2D", means2D) F("cov3D", cov3D) F("conic_opacity", conic_opacity) F("rgb", rgb)
    F("accum_alpha", final_T) F("out_color", out_color) F("radii", radii) F("tiles_touched", tiles_touched)
    F("point_offsets", point_offsets) F("point_list_unsorted", vals_unsorted) F("point_list", vals)
    F("n_contrib", n_contrib) F("ranges", ranges) F("clamped", clamped) F("point_list_keys_unsorted", keys_unsorted)
    F("point_list_keys", keys) F("dL_dmean2D", dmean2D) F("dL_dconic", dconic) F("dL_dopacity", dopacity)
    F("dL_dcolor", dcolor) F("dL_dmean3D", dmean3D) F("dL_dcov3D", dcov3D) F("dL_dsh", dsh) F("dL_dscale", dscale)
    F("dL_drot", drot)
#undef F
    return -1;
}

int oracle_num_threads() { return omp_get_max_threads(); }

}  // extern "C"
