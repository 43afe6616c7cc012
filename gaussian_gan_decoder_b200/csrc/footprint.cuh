// Conservative footprint test shared by the forward blend (blend.cu) and the backward blend (raster_backward.cu).
#pragma once
#include <cstdint>

namespace ggd {

// Second test of warp_block_mask (see there), for splats larger than a block.  Not inlined: inside the blend kernel its
// arithmetic raised the register allocation of the whole kernel by four and cost the per-hit loop its uniform-register
// addressing (3 more instructions per hit, +4.5 % on the frame of small splats that never gets here).
static __device__ __noinline__ uint32_t refine_block_mask(uint32_t box, float cx, float cy, float a, float b, float c, float s)
{
    const float nb_over_c = -b / c, nb_over_a = -b / a, s_far = s * 1.01f;
    uint32_t m = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const float bx0 = (float)((w & 1) * 8), by0 = (float)((w >> 1) * 4);
        if (!((box >> w) & 1u)) continue;
        const float dx0 = cx - bx0 - 7.f, dx1 = cx - bx0, dy0 = cy - by0 - 3.f, dy1 = cy - by0;   // d = centre - pixel
        const float ex = fminf(fmaxf(0.f, dx0), dx1), ey = fminf(fmaxf(0.f, dy0), dy1);
        float qmin = 0.f;
        if (ex != 0.f || ey != 0.f) {
            qmin = 3.0e38f;
            if (ex != 0.f) {
                const float dy = fminf(fmaxf(nb_over_c * ex, dy0), dy1);
                qmin = fminf(qmin, a * ex * ex + dy * (2.f * b * ex + c * dy));
            }
            if (ey != 0.f) {
                const float dx = fminf(fmaxf(nb_over_a * ey, dx0), dx1);
                qmin = fminf(qmin, c * ey * ey + dx * (2.f * b * ey + a * dx));
            }
        }
        if (!(qmin > s_far)) m |= 1u << w;     // NaN keeps the block
    }
    return m;
}

// Conservative footprint test.  An instance can only change a pixel if
// alpha = min(0.99, o * exp(power)) >= 1/255, i.e. d^T Q d <= 2 ln(255 o) with Q the conic.
// The staging thread turns that ellipse into a pixel-space box, inflated by 1 % (the fp32
// evaluation of `power` in the blend loop is accurate to ~4e-4 relative once the conic's
// condition number is bounded, which is checked: ill-conditioned conics are never culled),
// and records which of the eight 8x4-pixel warp blocks the box touches.  A warp then walks
// only the instances whose bit is set.  Skipped (pixel, instance) pairs are exactly pairs the
// reference rejects with `alpha < 1/255`, so T, C, n_contrib keep their reference bits; the
// contributor index is recovered from the instance's position in the tile list.
__device__ __forceinline__ uint32_t warp_block_mask(float x, float y, float a, float b, float c, float o,
                                                    float tile_x0, float tile_y0)
{
    if (!(o >= 1.0f / 255.0f)) return (o == o) ? 0u : 0xffu;   // alpha <= o < 1/255: never accepted (NaN: keep)
    const float det = a * c - b * b, tr = a + c;
    if (!(det > 0.f) || !(tr * tr <= 1000.f * det) || !(o <= 1e30f) || !(a > 0.f) || !(c > 0.f)) return 0xffu;
    const float s = 2.02f * logf(255.0f * o) + 1e-4f;
    const float inv = 1.0f / det;
    const float rx = sqrtf(s * c * inv) * 1.005f + 1e-3f, ry = sqrtf(s * a * inv) * 1.005f + 1e-3f;
    const float lx0 = x - rx - tile_x0, lx1 = x + rx - tile_x0, ly0 = y - ry - tile_y0, ly1 = y + ry - tile_y0;
    if (!(lx0 == lx0) || !(lx1 == lx1) || !(ly0 == ly0) || !(ly1 == ly1)) return 0xffu;
    // A block the box touches is looked at once more: the smallest value the quadratic form d^T Q d takes over the
    // block's rectangle of pixel centres (a convex function: 0 if the centre lies inside, else the minimum over the one or
    // two edges facing the centre, each a clamped 1-D parabola).  Beyond s * 1.01 no pixel of the block can reach
    // alpha >= 1/255.  The box alone passes 38 % of a trained-like frame's (block, instance) visits for nothing — the
    // corners of the boxes of elongated, rotated splats; 6 % for the small round splats of a random-init decoder.
    // (A splat whose box is smaller than a block gains nothing from it — the random-init frame — and keeps the box test.)
    uint32_t box = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const float bx0 = (float)((w & 1) * 8), by0 = (float)((w >> 1) * 4);
        if (lx1 >= bx0 && lx0 <= bx0 + 7.f && ly1 >= by0 && ly0 <= by0 + 3.f) box |= 1u << w;
    }
    if (!(rx + ry > 6.0f)) return box;
    return refine_block_mask(box, x - tile_x0, y - tile_y0, a, b, c, s);
}

}  // namespace ggd
