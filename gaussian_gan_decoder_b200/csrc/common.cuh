// Shared host/device declarations of the sm_100a hot path.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../include/ggd.h"

namespace ggd {

// ---- error plumbing -------------------------------------------------------
void set_error(const char* fmt, ...);
#define GGD_CUDA_CHECK(expr)                                                          \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            ::ggd::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),  \
                             __FILE__, __LINE__);                                     \
            return GGD_E_CUDA;                                                        \
        }                                                                             \
    } while (0)
void count_launch();   // one call per kernel launch: the counter behind ggd_launch_count()
#define GGD_LAUNCH_CHECK(name)                                                        \
    do {                                                                              \
        ::ggd::count_launch();                                                        \
        cudaError_t _e = cudaGetLastError();                                          \
        if (_e != cudaSuccess) {                                                      \
            ::ggd::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e)); \
            return GGD_E_CUDA;                                                        \
        }                                                                             \
    } while (0)

// ---- optional per-stage device timing (ggd_stage_timing / ggd_stage_times) ----
enum Stage { kStPlanes = 0, kStDecode, kStPreprocess, kStScan, kStDuplicate, kStSort, kStRanges, kStBlend, kStCount };
void stage_mark(cudaStream_t s, int stage, bool begin);   // no-op unless timing is enabled
struct StageScope {
    cudaStream_t s; int st;
    StageScope(cudaStream_t s_, int st_) : s(s_), st(st_) { stage_mark(s, st, true); }
    ~StageScope() { stage_mark(s, st, false); }
};

constexpr int kTileX = GGD_TILE_X;
constexpr int kTileY = GGD_TILE_Y;
constexpr int kTilePixels = kTileX * kTileY;

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// SM count of the CURRENT device (cached per device ordinal, capi.cu)
int device_sm_count();

// ---- radix sort geometry (binning.cu) -------------------------------------
constexpr int kRadixBits = 8;
constexpr int kRadixBins = 1 << kRadixBits;
constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;                       // keys per thread
constexpr int kSortTile = kSortThreads * kSortItems;  // keys per CTA tile
constexpr int kMaxSortPasses = 8;                     // 64 key bits / 8

// Scratch layout of one sort: [pass histograms][tile tickets][look-back words]
struct SortScratch {
    uint32_t* hist;      // [kMaxSortPasses][kRadixBins] digit counts, then exclusive offsets
    uint32_t* tickets;   // [kMaxSortPasses + 1] dynamic tile ids (last one: histogram kernel); word 16: identity-pass mask
    uint32_t* lookback;  // [kMaxSortPasses][tiles][kRadixBins] decoupled look-back words
    size_t bytes;        // total
    size_t zero_bytes;   // prefix that must be zero before a sort starts
};
SortScratch carve_sort_scratch(void* base, uint64_t capacity);

int sort_pairs(cudaStream_t s, const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in,
               uint32_t* vals_out, const uint32_t* n_dev, uint64_t capacity, int begin_bit,
               int end_bit, uint64_t* keys_tmp, uint32_t* vals_tmp, const SortScratch& sc,
               bool hist_precomputed);

// ---- rasteriser state buffers ---------------------------------------------
// Every field starts on a 256-byte boundary from the (256-byte aligned) base.
struct GeomState {
    float* depths;            // [P]
    int* radii;               // [P] internal radii (used when the caller passes none)
    float2* means2D;          // [P]
    float* cov3D;             // [P][6]
    float4* conic_opacity;    // [P]
    float* rgb;               // [P][3]
    uint32_t* tiles_touched;  // [P]
    uint32_t* point_offsets;  // [P] inclusive scan of tiles_touched
    uint8_t* clamped;         // [P][3]
    uint32_t* counters;       // [0] num_rendered, [1] overflow flag, [2] scan ticket, [3] clamped count,
                              // [4] 1 = this frame is binned per tile (tile_binning.cu), [5] longest tile list,
                              // [6] element count the radix kernels see, [7] number of tile lists > kTileSortMax,
                              // [8], [10] tile queues of the blend kernels, [12] 1 = preprocess counted tile_counts
                              // for this frame (so stage 2 never trusts counts that were not written), [13] 1 = part of
                              // the counts sits in the difference array (word 3 of the tile records)
    uint64_t* scan_status;    // [scan tiles] decoupled look-back words of the scan
    uint32_t* tile_counts;    // [tiles][kTileCountStride] word 0: instances per tile, counted by preprocess; word 2: the
                              // tile's write cursor; word 1 of entry i: id of the i-th long tile list; word 3: 2-D
                              // difference array of the large rectangles (+1 / -1 at their corners).  One 128-byte
                              // line per tile so the atomics spread over the L2 slices (last field: size depends on W, H)
    size_t bytes;
};
struct BinState {
    uint64_t* keys_unsorted;  // [cap]
    uint64_t* keys;           // [cap] sorted
    uint64_t* keys_tmp;       // [cap]
    uint32_t* vals_unsorted;  // [cap]
    uint32_t* vals;           // [cap] sorted  (= point_list)
    uint32_t* vals_tmp;       // [cap]
    void* sort_scratch;
    size_t sort_scratch_bytes;
    size_t bytes;
};
struct ImgState {
    float* accum_alpha;   // [W*H] final transmittance T
    uint32_t* n_contrib;  // [W*H]
    uint2* ranges;        // [tiles]
    size_t bytes;
};
GeomState carve_geom(void* base, size_t P, size_t tiles = 0);
BinState carve_bin(void* base, uint64_t capacity);
ImgState carve_img(void* base, int W, int H);

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

// ---- kernel launchers (one translation unit each) -------------------------
struct PreprocessArgs {
    int P, D, M, W, H;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* opacities;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    const float* view;
    const float* proj;
    const float* campos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int prefiltered;
    int grid_x, grid_y;
    uint32_t* tile_counts;   // optional [grid_x * grid_y][32]: word 0 += 1 for every tile a splat touches
    uint32_t* counts_valid;  // set to 1 by the kernel when tile_counts were written (GeomState::counters[12])
    int hist_tiles;          // set by launch_preprocess: size of the CTA-local tile histogram (0 = none)
    uint32_t* diff_used;     // GeomState::counters[13]: set when a large rectangle was counted through the difference array
    int diff_ok;             // set by launch_preprocess: tile_scan_kernel can undo a difference array of this grid
};
int launch_preprocess(cudaStream_t s, const PreprocessArgs& a, const GeomState& g, int* radii);
int launch_scan(cudaStream_t s, int P, const GeomState& g);
// `skip` (optional, device): the kernels return at once when *skip != 0 (frame binned per tile instead)
int launch_duplicate(cudaStream_t s, int P, int grid_x, int grid_y, const GeomState& g,
                     const int* radii, const BinState& b, uint64_t capacity,
                     const SortScratch& sc, int num_passes, const uint32_t* skip = nullptr);
int launch_tile_ranges(cudaStream_t s, const GeomState& g, const BinState& b, uint64_t capacity,
                       const ImgState& im, int tiles, const uint32_t* skip = nullptr);
// per-tile binning (tile_binning.cu)
constexpr uint32_t kTileCountStride = 32;  // words between two tiles' counters
constexpr uint32_t kTileSortMax = 4096;   // longest tile list the register-resident shared-memory sort takes
constexpr uint32_t kTileSortLargeMax = 24576;   // words the one-CTA-per-SM variant sorts at once (192 KB of words + 32 KB of buckets)
constexpr uint32_t kTileSortHugeMax = 4 * kTileSortLargeMax;   // longest list of the per-tile path: sorted in depth chunks of <= kTileSortLargeMax
int launch_tile_scan(cudaStream_t s, const GeomState& g, const ImgState& im, int grid_x, int grid_y, uint64_t capacity, int P, bool strict);
constexpr int kTileDiffMaxTiles = 16384;   // largest grid whose difference array tile_scan_kernel sums in shared memory (2048^2)
int launch_tile_scatter(cudaStream_t s, int P, int grid_x, int grid_y, const GeomState& g, const int* radii,
                        const BinState& b, const ImgState& im, uint64_t capacity, bool ref_order_arrays);
int launch_tile_sort(cudaStream_t s, const GeomState& g, const BinState& b, const ImgState& im, int tiles, int lazy = 0);
int launch_blend(cudaStream_t s, int W, int H, int grid_x, int grid_y, const GeomState& g,
                 const BinState& b, const ImgState& im, const float* features,
                 const float* background, float* out_color, int round = 0);
int launch_mark_visible(cudaStream_t s, int P, const float* means3D, const float* view,
                        const float* proj, uint8_t* present);

#ifdef __CUDACC__
// Tile rectangle of a splat (reference: getRect, DGR/cuda_rasterizer/auxiliary.h:46-56).
// Used by preprocess (tiles_touched) and by duplication (key emission); both must
// see the very same rectangle, so there is exactly one definition.
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, int grid_x, int grid_y,
                                          uint2& rmin, uint2& rmax)
{
    const unsigned gx = (unsigned)grid_x, gy = (unsigned)grid_y;
    rmin.x = min(gx, (unsigned)max((int)0, (int)((p.x - max_radius) / kTileX)));
    rmin.y = min(gy, (unsigned)max((int)0, (int)((p.y - max_radius) / kTileY)));
    rmax.x = min(gx, (unsigned)max((int)0, (int)((p.x + max_radius + kTileX - 1) / kTileX)));
    rmax.y = min(gy, (unsigned)max((int)0, (int)((p.y + max_radius + kTileY - 1) / kTileY)));
}
#endif

// Number of key bits the tile id occupies: the bit index one above the MSB of
// `n` (reference: getHigherMsb, DGR/cuda_rasterizer/rasterizer_impl.cu:35-50).
inline int tile_key_bits(uint32_t n)
{
    int b = 0;
    while (b < 32 && (n >> b)) ++b;
    return b;
}

}  // namespace ggd
