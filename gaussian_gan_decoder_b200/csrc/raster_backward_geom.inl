// Per-Gaussian part of the rasteriser backward (included by raster_backward.cu).
namespace ggd {
namespace {

struct B3 { float c[3][3]; };   // column-major c[col][row]
__device__ __forceinline__ B3 bmul(const B3& a, const B3& b)
{
    B3 r;
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
    return r;
}
__device__ __forceinline__ B3 btr(const B3& a)
{
    B3 r;
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[i][j];
    return r;
}

struct GeomBwdArgs {
    int P, D, M;
    const float* means3D; const int* radii; const float* shs; const uint8_t* clamped;
    const float* scales; const float* rotations; float scale_modifier;
    const float* cov3D;            // precomputed or the forward's
    const float* view; const float* proj; const float* campos;
    float focal_x, focal_y, tan_fovx, tan_fovy;
    const float* dL_dmean2D;       // [P][3]
    const float* dL_dconic;        // [P][4]
    const float* dL_dcolor;        // [P][3]
    float* dL_dmean3D; float* dL_dcov3D; float* dL_dsh; float* dL_dscale; float* dL_drot;
};

__constant__ float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                             -1.0925484305920792f, 0.5462742152960396f};
__constant__ float kC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                             -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 operator*(float s, F3 v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ float dot3(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

__global__ void __launch_bounds__(256)
geom_backward_kernel(const GeomBwdArgs a)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P || !(a.radii[idx] > 0)) return;
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float* vm = a.view;
    const float* pj = a.proj;
    const float3 mean = make_float3(a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]);

    // ---- conic -> cov2D -> cov3D and mean (through the EWA Jacobian) ----
    const float* cov3D = a.cov3D + 6 * (size_t)idx;
    const float3 dconic = make_float3(a.dL_dconic[4 * idx], a.dL_dconic[4 * idx + 1], a.dL_dconic[4 * idx + 3]);
    float3 t = make_float3(vm[0] * mean.x + vm[4] * mean.y + vm[8] * mean.z + vm[12],
                           vm[1] * mean.x + vm[5] * mean.y + vm[9] * mean.z + vm[13],
                           vm[2] * mean.x + vm[6] * mean.y + vm[10] * mean.z + vm[14]);
    const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float h_x = a.focal_x, h_y = a.focal_y;
    const B3 J = {{{h_x / t.z, 0.f, -(h_x * t.x) / (t.z * t.z)}, {0.f, h_y / t.z, -(h_y * t.y) / (t.z * t.z)}, {0.f, 0.f, 0.f}}};
    const B3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
    const B3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
    const B3 T = bmul(Wm, J);
    const B3 c2 = bmul(bmul(btr(T), btr(Vrk)), T);
    const float ca = c2.c[0][0] + 0.3f, cb = c2.c[0][1], cc = c2.c[1][1] + 0.3f;
    const float denom = ca * cc - cb * cb;
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float* dcov = a.dL_dcov3D + 6 * (size_t)idx;
    float dc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dconic.x + 2 * cb * cc * dconic.y + (denom - ca * cc) * dconic.z);
        dL_dc = denom2inv * (-ca * ca * dconic.z + 2 * ca * cb * dconic.y + (denom - ca * cc) * dconic.x);
        dL_db = denom2inv * 2 * (cb * cc * dconic.x - (denom + 2 * cb * cb) * dconic.y + ca * cb * dconic.z);
        dc[0] = T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc;
        dc[3] = T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc;
        dc[5] = T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc;
        dc[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
        dc[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
        dc[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) dcov[i] = dc[i];

    float dT[2][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float r0 = T.c[0][0] * Vrk.c[k][0] + T.c[0][1] * Vrk.c[k][1] + T.c[0][2] * Vrk.c[k][2];
        const float r1 = T.c[1][0] * Vrk.c[k][0] + T.c[1][1] * Vrk.c[k][1] + T.c[1][2] * Vrk.c[k][2];
        dT[0][k] = 2 * r0 * dL_da + r1 * dL_db;
        dT[1][k] = 2 * r1 * dL_dc + r0 * dL_db;
    }
    const float dJ00 = Wm.c[0][0] * dT[0][0] + Wm.c[0][1] * dT[0][1] + Wm.c[0][2] * dT[0][2];
    const float dJ02 = Wm.c[2][0] * dT[0][0] + Wm.c[2][1] * dT[0][1] + Wm.c[2][2] * dT[0][2];
    const float dJ11 = Wm.c[1][0] * dT[1][0] + Wm.c[1][1] * dT[1][1] + Wm.c[1][2] * dT[1][2];
    const float dJ12 = Wm.c[2][0] * dT[1][0] + Wm.c[2][1] * dT[1][1] + Wm.c[2][2] * dT[1][2];
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const float dtx = x_grad_mul * -h_x * tz2 * dJ02;
    const float dty = y_grad_mul * -h_y * tz2 * dJ12;
    const float dtz = -h_x * tz2 * dJ00 - h_y * tz2 * dJ11 + (2 * h_x * t.x) * tz3 * dJ02 + (2 * h_y * t.y) * tz3 * dJ12;
    float3 dmean = make_float3(vm[0] * dtx + vm[1] * dty + vm[2] * dtz, vm[4] * dtx + vm[5] * dty + vm[6] * dtz,
                               vm[8] * dtx + vm[9] * dty + vm[10] * dtz);

    // ---- mean2D -> mean3D through the projection ----
    const float hw = pj[3] * mean.x + pj[7] * mean.y + pj[11] * mean.z + pj[15];
    const float m_w = 1.0f / (hw + 0.0000001f);
    const float mul1 = (pj[0] * mean.x + pj[4] * mean.y + pj[8] * mean.z + pj[12]) * m_w * m_w;
    const float mul2 = (pj[1] * mean.x + pj[5] * mean.y + pj[9] * mean.z + pj[13]) * m_w * m_w;
    const float g2x = a.dL_dmean2D[3 * idx], g2y = a.dL_dmean2D[3 * idx + 1];
    dmean.x += (pj[0] * m_w - pj[3] * mul1) * g2x + (pj[1] * m_w - pj[3] * mul2) * g2y;
    dmean.y += (pj[4] * m_w - pj[7] * mul1) * g2x + (pj[5] * m_w - pj[7] * mul2) * g2y;
    dmean.z += (pj[8] * m_w - pj[11] * mul1) * g2x + (pj[9] * m_w - pj[11] * mul2) * g2y;

    // ---- RGB -> SH (and the view-direction term of the mean gradient for degree > 0) ----
    if (a.shs) {
        const F3* sh = reinterpret_cast<const F3*>(a.shs) + (size_t)idx * a.M;
        F3* dsh = reinterpret_cast<F3*>(a.dL_dsh) + (size_t)idx * a.M;
        F3 dRGB = {a.dL_dcolor[3 * idx], a.dL_dcolor[3 * idx + 1], a.dL_dcolor[3 * idx + 2]};
        if (a.clamped[3 * idx]) dRGB.x = 0.f;
        if (a.clamped[3 * idx + 1]) dRGB.y = 0.f;
        if (a.clamped[3 * idx + 2]) dRGB.z = 0.f;
        dsh[0] = C0 * dRGB;
        if (a.D > 0) {
            const F3 dir_orig = {mean.x - a.campos[0], mean.y - a.campos[1], mean.z - a.campos[2]};
            const float len = sqrtf(dot3(dir_orig, dir_orig));
            const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
            dsh[1] = (-C1 * y) * dRGB; dsh[2] = (C1 * z) * dRGB; dsh[3] = (-C1 * x) * dRGB;
            F3 dx = (-C1) * sh[3], dy = (-C1) * sh[1], dz = C1 * sh[2];
            if (a.D > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                dsh[4] = (kC2[0] * xy) * dRGB; dsh[5] = (kC2[1] * yz) * dRGB; dsh[6] = (kC2[2] * (2.f * zz - xx - yy)) * dRGB;
                dsh[7] = (kC2[3] * xz) * dRGB; dsh[8] = (kC2[4] * (xx - yy)) * dRGB;
                dx = dx + (kC2[0] * y) * sh[4] + (kC2[2] * 2.f * -x) * sh[6] + (kC2[3] * z) * sh[7] + (kC2[4] * 2.f * x) * sh[8];
                dy = dy + (kC2[0] * x) * sh[4] + (kC2[1] * z) * sh[5] + (kC2[2] * 2.f * -y) * sh[6] + (kC2[4] * 2.f * -y) * sh[8];
                dz = dz + (kC2[1] * y) * sh[5] + (kC2[2] * 2.f * 2.f * z) * sh[6] + (kC2[3] * x) * sh[7];
                if (a.D > 2) {
                    dsh[9] = (kC3[0] * y * (3.f * xx - yy)) * dRGB; dsh[10] = (kC3[1] * xy * z) * dRGB;
                    dsh[11] = (kC3[2] * y * (4.f * zz - xx - yy)) * dRGB; dsh[12] = (kC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dRGB;
                    dsh[13] = (kC3[4] * x * (4.f * zz - xx - yy)) * dRGB; dsh[14] = (kC3[5] * z * (xx - yy)) * dRGB;
                    dsh[15] = (kC3[6] * x * (xx - 3.f * yy)) * dRGB;
                    dx = dx + (kC3[0] * 3.f * 2.f * xy) * sh[9] + (kC3[1] * yz) * sh[10] + (kC3[2] * -2.f * xy) * sh[11] +
                         (kC3[3] * -3.f * 2.f * xz) * sh[12] + (kC3[4] * (-3.f * xx + 4.f * zz - yy)) * sh[13] +
                         (kC3[5] * 2.f * xz) * sh[14] + (kC3[6] * 3.f * (xx - yy)) * sh[15];
                    dy = dy + (kC3[0] * 3.f * (xx - yy)) * sh[9] + (kC3[1] * xz) * sh[10] + (kC3[2] * (-3.f * yy + 4.f * zz - xx)) * sh[11] +
                         (kC3[3] * -3.f * 2.f * yz) * sh[12] + (kC3[4] * -2.f * xy) * sh[13] + (kC3[5] * -2.f * yz) * sh[14] +
                         (kC3[6] * -3.f * 2.f * xy) * sh[15];
                    dz = dz + (kC3[1] * xy) * sh[10] + (kC3[2] * 4.f * 2.f * yz) * sh[11] + (kC3[3] * 3.f * (2.f * zz - xx - yy)) * sh[12] +
                         (kC3[4] * 4.f * 2.f * xz) * sh[13] + (kC3[5] * (xx - yy)) * sh[14];
                }
            }
            const F3 dv = {dot3(dx, dRGB), dot3(dy, dRGB), dot3(dz, dRGB)};
            const F3 v = dir_orig;
            const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
            const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
            dmean.x += ((sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * inv32;
            dmean.y += (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * inv32;
            dmean.z += (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * inv32;
        }
    }
    a.dL_dmean3D[3 * idx] = dmean.x; a.dL_dmean3D[3 * idx + 1] = dmean.y; a.dL_dmean3D[3 * idx + 2] = dmean.z;

    // ---- cov3D -> scale, rotation ----
    if (a.scales) {
        const float* q = a.rotations + 4 * (size_t)idx;
        const float r = q[0], x = q[1], y = q[2], z = q[3];
        const B3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                       {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                       {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
        const float sx = a.scale_modifier * a.scales[3 * idx], sy = a.scale_modifier * a.scales[3 * idx + 1],
                    sz = a.scale_modifier * a.scales[3 * idx + 2];
        B3 M2;   // 2 * S * R
#pragma unroll
        for (int j = 0; j < 3; ++j) { M2.c[j][0] = 2.f * sx * R.c[j][0]; M2.c[j][1] = 2.f * sy * R.c[j][1]; M2.c[j][2] = 2.f * sz * R.c[j][2]; }
        const B3 dSig = {{{dc[0], 0.5f * dc[1], 0.5f * dc[2]}, {0.5f * dc[1], dc[3], 0.5f * dc[4]}, {0.5f * dc[2], 0.5f * dc[4], dc[5]}}};
        B3 dMt = btr(bmul(M2, dSig));
        const B3 Rt = btr(R);
        float* ds = a.dL_dscale + 3 * (size_t)idx;
        ds[0] = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
        ds[1] = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
        ds[2] = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
#pragma unroll
        for (int i = 0; i < 3; ++i) { dMt.c[0][i] *= sx; dMt.c[1][i] *= sy; dMt.c[2][i] *= sz; }
        float* dq = a.dL_drot + 4 * (size_t)idx;
        dq[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
        dq[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
        dq[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
        dq[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
    }
}

}  // namespace
}  // namespace ggd

using namespace ggd;

extern "C" int ggd_raster_backward(void* stream, int P, int D, int M, const float* background, int W, int H,
                                   const float* means3D, const float* shs, const float* colors_precomp,
                                   const float* scales, float scale_modifier, const float* rotations,
                                   const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                   const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                                   const void* geom_ws, const void* bin_ws, uint64_t bin_capacity, const void* img_ws,
                                   const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                                   float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                                   float* dL_dscale, float* dL_drot)
{
    if (P <= 0 || W <= 0 || H <= 0) { set_error("ggd_raster_backward: bad sizes"); return GGD_E_INVALID; }
    if (!background || !means3D || !viewmatrix || !projmatrix || !geom_ws || !bin_ws || !img_ws || !dL_dpix ||
        !dL_dmean2D || !dL_dconic || !dL_dopacity || !dL_dcolor || !dL_dmean3D || !dL_dcov3D) {
        set_error("ggd_raster_backward: null required pointer"); return GGD_E_INVALID;
    }
    if ((shs == nullptr) == (colors_precomp == nullptr)) { set_error("provide exactly one of shs / colors_precomp"); return GGD_E_INVALID; }
    if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr)) { set_error("provide exactly one of (scales, rotations) / cov3D_precomp"); return GGD_E_INVALID; }
    if (shs && !dL_dsh) { set_error("dL_dsh required with shs"); return GGD_E_INVALID; }
    if (scales && (!dL_dscale || !dL_drot)) { set_error("dL_dscale / dL_drot required with scales"); return GGD_E_INVALID; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GeomState g = carve_geom(const_cast<void*>(geom_ws), (size_t)P);
    BinState b = carve_bin(const_cast<void*>(bin_ws), bin_capacity);
    ImgState im = carve_img(const_cast<void*>(img_ws), W, H);
    if (!radii) radii = g.radii;
    const int grid_x = (W + kTileX - 1) / kTileX, grid_y = (H + kTileY - 1) / kTileY;
    const float* colors = colors_precomp ? colors_precomp : g.rgb;
    // the state buffers are inputs here except for two words of the geometry state's counter block, which serve as the
    // (self-resetting) tile queue of the persistent kernel
    const int sms = device_sm_count();
    const int tiles = grid_x * grid_y;
    blend_backward_kernel<<<tiles < 4 * sms ? tiles : 4 * sms, kTilePixels, 0, s>>>(
        im.ranges, b.vals, W, H, background, g.means2D, g.conic_opacity, colors, im.accum_alpha, im.n_contrib,
        dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, g.counters + 10, (uint32_t)tiles);
    GGD_LAUNCH_CHECK("blend_backward_kernel");
    GeomBwdArgs a;
    a.P = P; a.D = D; a.M = shs ? M : 0;
    a.means3D = means3D; a.radii = radii; a.shs = shs; a.clamped = g.clamped;
    a.scales = scales; a.rotations = rotations; a.scale_modifier = scale_modifier;
    a.cov3D = cov3D_precomp ? cov3D_precomp : g.cov3D;
    a.view = viewmatrix; a.proj = projmatrix; a.campos = campos;
    a.focal_y = H / (2.0f * tan_fovy); a.focal_x = W / (2.0f * tan_fovx);
    a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.dL_dmean2D = dL_dmean2D; a.dL_dconic = dL_dconic; a.dL_dcolor = dL_dcolor;
    a.dL_dmean3D = dL_dmean3D; a.dL_dcov3D = dL_dcov3D; a.dL_dsh = dL_dsh; a.dL_dscale = dL_dscale; a.dL_drot = dL_drot;
    geom_backward_kernel<<<(P + 255) / 256, 256, 0, s>>>(a);
    GGD_LAUNCH_CHECK("geom_backward_kernel");
    return GGD_OK;
}
