// Stage 2b of the rasteriser: per-tile front-to-back alpha blending.
//
// Replaces FORWARD::render → renderCUDA (DGR/cuda_rasterizer/forward.cu:261-400).
// A CTA works on one 16x16 tile at a time, one thread per pixel; six persistent CTAs per SM pull
// tiles from a queue (the reference launches one CTA per tile).  The per-pixel arithmetic —
// the falloff exponent, alpha = min(0.99, o·exp(power)), the 1/255 and 1e-4
// tests, C += c·alpha·T, and the contributor bookkeeping — keeps the
// reference's operation order exactly, because n_contrib (an integer output)
// and the final T depend on every rounding along the way.
//
// What differs: a batch stages, per instance, the complete blend record
// {xy, conic, opacity, rgb} in shared memory as a 48-byte record, so the
// inner loop has no global loads at all (the reference gathers the colour from
// global memory for every contributing (pixel, Gaussian) pair, forward.cu:355)
// and every shared-memory read is a 16-byte broadcast; and each warp owns an
// 8x4 pixel block and only walks the instances whose conservative footprint
// touches that block (see warp_block_mask: the splat's box, then — for splats larger
// than a block — the minimum of its quadratic form over the block).
#include "common.cuh"
#include "footprint.cuh"
#include <algorithm>
#include <cstdlib>

namespace ggd {
namespace {

constexpr int kBatch = 2 * kTilePixels;  // instances staged per round (two per thread): half as many CTA barriers as with 256,
                                         // at which the eight warps of a tile wait for the slowest one's hit list

template <int kRound>      // 0: plain; 1 / 2: the two rounds of a lazily sorted frame (see below) — a template argument so that
                           // the plain kernel keeps its register allocation
__global__ void __launch_bounds__(kTilePixels)         // 40 registers: six CTAs per SM (asking ptxas for that with a
                                                       // min-blocks bound cost the hit loop its uniform-register addressing)
blend_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
             const float2* __restrict__ means2D, const float* __restrict__ features,
             const float4* __restrict__ conic_opacity, float* __restrict__ final_T,
             uint32_t* __restrict__ n_contrib, const float* __restrict__ bg,
             float* __restrict__ out_color, uint32_t* __restrict__ work /*[2]: next tile, finished CTAs*/, uint32_t num_tiles,
             uint32_t* __restrict__ recs /*tile records*/, const uint32_t* __restrict__ counters)
{
    constexpr int round = kRound;
    // round 0: every tile, whole lists.  Lean pipelines sort long tile lists lazily (tile_sort_large_kernel): round 1 walks a
    // long list only as far as it is sorted (word 7 of the tile's record) and marks the tile (word 8) when it gets there with
    // live pixels; round 2 blends the marked tiles again, from the start, once their lists are sorted completely.  A tile that
    // saturates inside the sorted part — almost all of them — has read exactly the instances it would read of the full order.
    const bool lazy = round != 0 && __ldg(counters + 4) != 0u;      // (a frame that fell back to the radix sort has full lists)
    if (round == 2 && !lazy) return;
    __shared__ uint32_t s_tile_id;
    __shared__ __align__(16) float4 s_rec[3 * kBatch];  // per instance: {x, y, conic.x, conic.y}, {conic.z, opacity, r, g}, {b, -, -, -}
    __shared__ uint32_t s_mask[kBatch];
    // 32-bit shared-window addresses, formed once: inside the list walk the compiler would otherwise rebuild the
    // generic -> shared conversion for every record it touches (8 of the loop's 68 instructions)
    // (the mov keeps ptxas from re-materialising the conversion inside the loop)
    uint32_t rec_addr;
    asm volatile("mov.u32 %0, %1;" : "=r"(rec_addr) : "r"((uint32_t)__cvta_generic_to_shared(s_rec)));

    const uint32_t tiles_x = (W + kTileX - 1) / kTileX;
    const int tr = threadIdx.x, warp = tr >> 5, lane = tr & 31;
    // The grid is a few CTAs per SM that pull tiles from a counter until none is left: tile lists differ in length
    // by more than 10x inside one image, and with one CTA per tile (all resident at once on a 512^2 image) the SMs
    // that drew long lists finish long after the others.
    for (;;) {
    __syncthreads();                      // previous tile's shared records are no longer read
    if (tr == 0) s_tile_id = atomicAdd(work, 1u);
    __syncthreads();
    const uint32_t tile_id = s_tile_id;
    if (tile_id >= num_tiles) break;
    if (round == 2 && recs[(size_t)tile_id * kTileCountStride + 8] == 0u) continue;
    const uint32_t tile_bx = tile_id % tiles_x, tile_by = tile_id / tiles_x;
    // warp w owns the 8x4 pixel block at (8*(w&1), 4*(w>>1)) of the tile
    const uint32_t px = tile_bx * kTileX + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_by * kTileY + (warp >> 1) * 4 + (lane >> 3);
    const uint32_t pix_id = W * py + px;
    const float fx = (float)px, fy = (float)py;
    const float tile_x0 = (float)(tile_bx * kTileX), tile_y0 = (float)(tile_by * kTileY);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    // Tg is the transmittance the list walk multiplies with: equal to T while the pixel is live, 0 once it has
    // terminated (or lies outside the image).  A dead pixel then fails the 1e-4 test by itself, so the walk needs
    // no separate "done" branch; T keeps the value at termination for the output.
    float Tg = inside ? 1.0f : 0.0f;

    uint2 range = ranges[tile_id];
    const uint32_t list_len = range.y - range.x;
    if (lazy && round == 1 && list_len > kTileSortMax) range.y = range.x + min(list_len, recs[(size_t)tile_id * kTileCountStride + 7]);
    const int rounds = (int)((range.y - range.x + kBatch - 1) / kBatch);
    int to_do = (int)(range.y - range.x);

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f;

    for (int i = 0; i < rounds; ++i, to_do -= kBatch) {
        if (__syncthreads_count(Tg == 0.0f) == kTilePixels) break;
#pragma unroll
        for (int q = 0; q < kBatch / kTilePixels; ++q) {
            const int slot = q * kTilePixels + tr;
            const uint32_t progress = (uint32_t)i * kBatch + (uint32_t)slot;
            if (range.x + progress < range.y) {
                const uint32_t id = __ldg(point_list + range.x + progress);
                const float2 xy = __ldg(means2D + id);
                const float4 co = __ldg(conic_opacity + id);
                const float r = __ldg(features + 3 * (size_t)id);
                const float g = __ldg(features + 3 * (size_t)id + 1);
                const float b = __ldg(features + 3 * (size_t)id + 2);
                s_rec[3 * slot] = make_float4(xy.x, xy.y, co.x, co.y);
                s_rec[3 * slot + 1] = make_float4(co.z, co.w, r, g);
                s_rec[3 * slot + 2].x = b;
                s_mask[slot] = warp_block_mask(xy.x, xy.y, co.x, co.y, co.z, co.w, tile_x0, tile_y0);
            }
        }
        __syncthreads();

        const int lim = min(kBatch, to_do);
        for (int j0 = 0; j0 < lim; j0 += 32) {
            if (__all_sync(0xffffffffu, Tg == 0.0f)) break;
            const bool hit = (j0 + lane < lim) && ((s_mask[j0 + lane] >> warp) & 1u);
            uint32_t bits = __ballot_sync(0xffffffffu, hit);
            while (bits) {
                const int j = j0 + __ffs(bits) - 1;
                bits &= bits - 1;
                const uint32_t ra = rec_addr + 48u * j;
                float4 A, B;
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(A.x), "=f"(A.y), "=f"(A.z), "=f"(A.w) : "r"(ra));
                asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+16];" : "=f"(B.x), "=f"(B.y), "=f"(B.z), "=f"(B.w) : "r"(ra));
                // Every product/sum below is pinned (_rn intrinsics are never re-fused by ptxas) to
                // the FMA placement of the reference build: q = fma(dx, cx*dx, cz*dy*dy),
                // power = fma(q, -0.5, -(cy*dx*dy)), C = fma(T, alpha*c, C).
                const float dx = __fsub_rn(A.x, fx), dy = __fsub_rn(A.y, fy);
                const float q = __fmaf_rn(dx, __fmul_rn(dx, A.z), __fmul_rn(dy, __fmul_rn(dy, B.x)));
                const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, A.w)));
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, __fmul_rn(B.y, expf(power)));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = __fmul_rn(Tg, __fsub_rn(1.0f, alpha));
                if (test_T < 0.0001f) {
                    Tg = 0.0f;       // terminated (a dead pixel lands here again and again, harmlessly)
                    continue;
                }
                float cb;
                asm volatile("ld.shared.f32 %0, [%1+32];" : "=f"(cb) : "r"(ra));
                C0 = __fmaf_rn(Tg, __fmul_rn(alpha, B.z), C0);
                C1 = __fmaf_rn(Tg, __fmul_rn(alpha, B.w), C1);
                C2 = __fmaf_rn(Tg, __fmul_rn(alpha, cb), C2);
                T = Tg = test_T;
                last_contributor = (uint32_t)(i * kBatch + j + 1);
            }
        }
    }

    if (lazy && round == 1) {      // not saturated at the end of the sorted part: the tile is done again in round 2
        const int live = __syncthreads_count(Tg != 0.0f);
        if (tr == 0) recs[(size_t)tile_id * kTileCountStride + 8] = (live != 0 && range.y - range.x < list_len) ? 1u : 0u;
    }
    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        const size_t plane = (size_t)H * W;
        out_color[pix_id] = __fmaf_rn(T, bg[0], C0);
        out_color[plane + pix_id] = __fmaf_rn(T, bg[1], C1);
        out_color[2 * plane + pix_id] = __fmaf_rn(T, bg[2], C2);
    }
    }   // tile loop
    // the last CTA out leaves the counters at zero for the next frame
    if (tr == 0 && atomicAdd(work + 1, 1u) == gridDim.x - 1) { work[0] = 0; work[1] = 0; }
}

}  // namespace

int launch_blend(cudaStream_t s, int W, int H, int grid_x, int grid_y, const GeomState& g,
                 const BinState& b, const ImgState& im, const float* features,
                 const float* background, float* out_color, int round)
{
    const int sms = device_sm_count();
    const char* e = getenv("GGD_BLEND_CTAS");
    const int per_sm = e ? atoi(e) : 6;
    const int tiles = grid_x * grid_y;
    const int grid = std::min(tiles, sms * per_sm);
    auto kern = round == 0 ? blend_kernel<0> : round == 1 ? blend_kernel<1> : blend_kernel<2>;
    kern<<<grid, kTilePixels, 0, s>>>(im.ranges, b.vals, W, H, g.means2D, features, g.conic_opacity, im.accum_alpha, im.n_contrib,
                                      background, out_color, g.counters + 8, (uint32_t)tiles, g.tile_counts, g.counters);
    GGD_LAUNCH_CHECK("blend_kernel");
    return GGD_OK;
}

}  // namespace ggd
