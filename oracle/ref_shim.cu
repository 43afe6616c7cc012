// TEST INFRASTRUCTURE — not part of the product.
//
// C-ABI shim around the UNMODIFIED reference rasteriser so tests (and the
// reported "reference CUDA on the same B200" baseline) can execute it on the
// GPU box, where /root/reference does not exist.  This file is OURS; it only
// #includes the reference headers where they lie and links the reference's
// three .cu files, compiled in place by oracle/Makefile into oracle/_ref/.
//
// Reference entry points wrapped:
//   CudaRasterizer::Rasterizer::forward   DGR/cuda_rasterizer/rasterizer_impl.cu:198-337
//   CudaRasterizer::Rasterizer::backward  DGR/cuda_rasterizer/rasterizer_impl.cu:340-435
//   CudaRasterizer::Rasterizer::markVisible  rasterizer_impl.cu:141-152
//   {Geometry,Image,Binning}State::fromChunk rasterizer_impl.cu:155-194 (buffer layouts)
// (DGR = gaussian_splatting/submodules/diff-gaussian-rasterization)
#include <cstdint>
#include <cstdio>
#include <functional>
#include <cuda_runtime.h>
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"

namespace {
struct Grow {
    char* ptr = nullptr;
    size_t cap = 0;
    char* operator()(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            cudaMalloc(&ptr, n);
            cap = n;
        }
        return ptr;
    }
};
Grow g_geom, g_bin, g_img;
}  // namespace

extern "C" {

// Runs the reference forward. The three state buffers stay owned by the shim
// (grown on demand, reused across calls); their base pointers are returned.
int ggd_ref_forward(int P, int D, int M, const float* bg, int W, int H,
                    const float* means3D, const float* shs, const float* colors_precomp,
                    const float* opacities, const float* scales, float scale_modifier,
                    const float* rotations, const float* cov3D_precomp,
                    const float* view, const float* proj, const float* campos,
                    float tanfovx, float tanfovy, int prefiltered,
                    float* out_color, int* radii, int debug,
                    char** geom, char** binning, char** img)
{
    std::function<char*(size_t)> fg = [](size_t n) { return g_geom(n); };
    std::function<char*(size_t)> fb = [](size_t n) { return g_bin(n); };
    std::function<char*(size_t)> fi = [](size_t n) { return g_img(n); };
    int R = -1;
    try {
        R = CudaRasterizer::Rasterizer::forward(fg, fb, fi, P, D, M, bg, W, H, means3D, shs,
            colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp,
            view, proj, campos, tanfovx, tanfovy, prefiltered != 0, out_color, radii, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "ggd_ref_forward: %s\n", e.what());
        return -1;
    }
    if (geom) *geom = g_geom.ptr;
    if (binning) *binning = g_bin.ptr;
    if (img) *img = g_img.ptr;
    return R;
}

int ggd_ref_backward(int P, int D, int M, int R, const float* bg, int W, int H,
                     const float* means3D, const float* shs, const float* colors_precomp,
                     const float* scales, float scale_modifier, const float* rotations,
                     const float* cov3D_precomp, const float* view, const float* proj,
                     const float* campos, float tanfovx, float tanfovy, const int* radii,
                     const float* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                     float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D,
                     float* dL_dsh, float* dL_dscale, float* dL_drot, int debug)
{
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp,
            scales, scale_modifier, rotations, cov3D_precomp, view, proj, campos, tanfovx, tanfovy,
            radii, g_geom.ptr, g_bin.ptr, g_img.ptr, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity,
            dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "ggd_ref_backward: %s\n", e.what());
        return -1;
    }
    return 0;
}

// Copy `n` bytes at byte offset `off` of shim-owned state buffer `which`
// (0 geom, 1 binning, 2 img) to `dst` (device or host pointer).
int ggd_ref_read(int which, size_t off, void* dst, size_t n)
{
    const char* base = which == 0 ? g_geom.ptr : which == 1 ? g_bin.ptr : g_img.ptr;
    if (!base) return -1;
    return (int)cudaMemcpy(dst, base + off, n, cudaMemcpyDefault);
}

void ggd_ref_mark_visible(int P, float* means3D, float* view, float* proj, bool* present)
{
    CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
}

// Byte offsets of every field inside the three opaque state buffers, obtained
// from the reference's own fromChunk on a null base (so the CUB temp sizes are
// whatever this toolkit's CUB reports).
// geom: depths, clamped, internal_radii, means2D, cov3D, conic_opacity, rgb,
//       tiles_touched, scanning_space, point_offsets   (10 entries)
void ggd_ref_geom_offsets(size_t P, size_t* off)
{
    char* c = nullptr;
    auto s = CudaRasterizer::GeometryState::fromChunk(c, P);
    off[0] = (size_t)s.depths; off[1] = (size_t)s.clamped; off[2] = (size_t)s.internal_radii;
    off[3] = (size_t)s.means2D; off[4] = (size_t)s.cov3D; off[5] = (size_t)s.conic_opacity;
    off[6] = (size_t)s.rgb; off[7] = (size_t)s.tiles_touched; off[8] = (size_t)s.scanning_space;
    off[9] = (size_t)s.point_offsets;
}
// binning: point_list, point_list_unsorted, point_list_keys, point_list_keys_unsorted (4)
void ggd_ref_binning_offsets(size_t R, size_t* off)
{
    char* c = nullptr;
    auto s = CudaRasterizer::BinningState::fromChunk(c, R);
    off[0] = (size_t)s.point_list; off[1] = (size_t)s.point_list_unsorted;
    off[2] = (size_t)s.point_list_keys; off[3] = (size_t)s.point_list_keys_unsorted;
}
// img: accum_alpha, n_contrib, ranges (3)
void ggd_ref_img_offsets(size_t N, size_t* off)
{
    char* c = nullptr;
    auto s = CudaRasterizer::ImageState::fromChunk(c, N);
    off[0] = (size_t)s.accum_alpha; off[1] = (size_t)s.n_contrib; off[2] = (size_t)s.ranges;
}

}  // extern "C"
