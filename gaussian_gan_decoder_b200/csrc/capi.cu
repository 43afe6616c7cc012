// extern "C" entry points declared in include/ggd.h: argument validation,
// state-buffer carving and stage orchestration.  No kernels live here.
#include "common.cuh"
#include <atomic>
#include <cstdarg>
#include <cstring>
#include <algorithm>

namespace ggd {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- stage timing ---------------------------------------------------------
// A benchmarking aid, off by default.  One event set per device; while it is on, drive each device from
// one host thread and one stream at a time (the events of two interleaved frames would overwrite each other).
static std::atomic<bool> g_timing{false};
struct StageEvents { cudaEvent_t ev[kStCount][2]; bool made = false, used[kStCount]; };
static StageEvents g_stage[64];
static StageEvents* stage_events()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    return &g_stage[dev];
}

void stage_mark(cudaStream_t s, int stage, bool begin)
{
    if (!g_timing.load(std::memory_order_relaxed)) return;
    StageEvents* e = stage_events();
    if (!e) return;
    if (!e->made) {
        for (int i = 0; i < kStCount; ++i) { cudaEventCreate(&e->ev[i][0]); cudaEventCreate(&e->ev[i][1]); e->used[i] = false; }
        e->made = true;
    }
    cudaEventRecord(e->ev[stage][begin ? 0 : 1], s);
    if (!begin) e->used[stage] = true;
}

namespace {
template <typename T>
T* take(char*& p, size_t count)
{
    T* r = reinterpret_cast<T*>(p);
    p += align_up(count * sizeof(T), 256);
    return r;
}
}  // namespace

GeomState carve_geom(void* base, size_t P, size_t tiles)
{
    GeomState g;
    char* p = static_cast<char*>(base);
    g.depths = take<float>(p, P);
    g.radii = take<int>(p, P);
    g.means2D = take<float2>(p, P);
    g.cov3D = take<float>(p, 6 * P);
    g.conic_opacity = take<float4>(p, P);
    g.rgb = take<float>(p, 3 * P);
    g.tiles_touched = take<uint32_t>(p, P);
    g.point_offsets = take<uint32_t>(p, P);
    g.clamped = take<uint8_t>(p, 3 * P);
    g.counters = take<uint32_t>(p, 64);  // counters and scan_status are adjacent: one memset
    g.scan_status = take<uint64_t>(p, (P + kScanTile - 1) / kScanTile + 1);
    g.tile_counts = take<uint32_t>(p, std::max<size_t>(tiles, 1) * kTileCountStride);   // zeroed by the same memset
    g.bytes = (size_t)(p - static_cast<char*>(base));
    return g;
}

BinState carve_bin(void* base, uint64_t cap)
{
    BinState b;
    char* p = static_cast<char*>(base);
    const size_t n = std::max<uint64_t>(cap, 1);
    b.keys_unsorted = take<uint64_t>(p, n);
    b.keys = take<uint64_t>(p, n);
    b.keys_tmp = take<uint64_t>(p, n);
    b.vals_unsorted = take<uint32_t>(p, n);
    b.vals = take<uint32_t>(p, n);
    b.vals_tmp = take<uint32_t>(p, n);
    b.sort_scratch = p;
    b.sort_scratch_bytes = carve_sort_scratch(p, cap).bytes;
    p += b.sort_scratch_bytes;
    b.bytes = (size_t)(p - static_cast<char*>(base));
    return b;
}

ImgState carve_img(void* base, int W, int H)
{
    ImgState im;
    char* p = static_cast<char*>(base);
    const size_t N = (size_t)W * H;
    const size_t tiles = (size_t)((W + kTileX - 1) / kTileX) * ((H + kTileY - 1) / kTileY);
    im.accum_alpha = take<float>(p, N);
    im.n_contrib = take<uint32_t>(p, N);
    im.ranges = take<uint2>(p, tiles);
    im.bytes = (size_t)(p - static_cast<char*>(base));
    return im;
}

namespace {

bool aligned256(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 255) == 0; }

int check_common(int P, int W, int H)
{
    if (P <= 0) { set_error("P must be positive (got %d); callers short-circuit P == 0", P); return GGD_E_INVALID; }
    if (W <= 0 || H <= 0) { set_error("bad image size %dx%d", W, H); return GGD_E_INVALID; }
    const long long tiles = (long long)((W + kTileX - 1) / kTileX) * ((H + kTileY - 1) / kTileY);
    if (tiles >= (1ll << 31) || (W + kTileX - 1) / kTileX >= 65536 || (H + kTileY - 1) / kTileY >= 65536) {
        set_error("image %dx%d has too many tiles", W, H);
        return GGD_E_INVALID;
    }
    return GGD_OK;
}

size_t tiles_of(int W, int H) { return (size_t)((W + kTileX - 1) / kTileX) * (size_t)((H + kTileY - 1) / kTileY); }

// How stage 2a bins: 0 = global radix sort only, 1 = per tile when the frame qualifies (decided on the device,
// both kernel sets enqueued), 2 = per tile only (a frame that does not qualify raises status flag 2).
// Start-up value from GGD_BINNING ("radix" | "auto" | "tiles"), overridden by ggd_raster_binning().
// That process-wide value is only the DEFAULT a call gets when it passes binning = -1; every raster entry
// point takes the mode of its frame as an argument, so two callers never share it.
std::atomic<int> g_binning{-1};
int default_binning()
{
    int m = g_binning.load(std::memory_order_relaxed);
    if (m < 0) {
        const char* e = getenv("GGD_BINNING");
        m = (e && e[0] == 'r') ? 0 : (e && e[0] == 't') ? 2 : 1;
        g_binning.store(m, std::memory_order_relaxed);
    }
    return m;
}
// `lean` (GGD_BINNING_LEAN or-ed into the mode): the per-tile path does not write the reference-order key/value arrays
bool resolve_binning(int& mode, const char* who, bool* lean = nullptr)
{
    if (mode == -1) mode = default_binning();
    const bool l = mode >= 0 && (mode & GGD_BINNING_LEAN) != 0;
    if (mode >= 0) mode &= ~GGD_BINNING_LEAN;
    if (lean) *lean = l;
    if (mode < 0 || mode > 2) { set_error("%s: binning must be -1 (default), 0, 1 or 2, optionally | GGD_BINNING_LEAN (got %d)", who, mode); return false; }
    return true;
}

}  // namespace

int device_sm_count()
{
    static std::atomic<int> cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
    int v = dev >= 0 ? cache[dev].load(std::memory_order_relaxed) : 0;
    if (!v) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev < 0 ? 0 : dev) != cudaSuccess || v <= 0) v = 148;
        if (dev >= 0) cache[dev].store(v, std::memory_order_relaxed);
    }
    return v;
}
}  // namespace ggd

using namespace ggd;

extern "C" {

int ggd_raster_binning(int mode)
{
    const int prev = default_binning();
    if (mode >= 0 && mode <= 2) g_binning.store(mode, std::memory_order_relaxed);
    return prev;
}

const char* ggd_last_error(void) { return g_err; }
uint64_t ggd_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
int ggd_abi_version(void) { return 3; }

int ggd_raster_workspace_sizes(int P, int W, int H, uint64_t bin_capacity, size_t sizes_out[3])
{
    if (P < 0 || W <= 0 || H <= 0 || !sizes_out) { set_error("ggd_raster_workspace_sizes: bad arguments"); return GGD_E_INVALID; }
    if (bin_capacity >= (1ull << 30)) { set_error("bin_capacity %llu exceeds 2^30-1", (unsigned long long)bin_capacity); return GGD_E_INVALID; }
    sizes_out[0] = carve_geom(nullptr, (size_t)std::max(P, 1), tiles_of(W, H)).bytes;
    sizes_out[1] = carve_bin(nullptr, bin_capacity).bytes;
    sizes_out[2] = carve_img(nullptr, W, H).bytes;
    return GGD_OK;
}

int ggd_raster_preprocess(void* stream, int P, int D, int M, int W, int H,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp,
                          const float* viewmatrix, const float* projmatrix, const float* campos,
                          float tan_fovx, float tan_fovy, int prefiltered,
                          void* geom_ws, int* radii, uint32_t* num_rendered_host, int binning)
{
    if (int e = check_common(P, W, H)) return e;
    if (!resolve_binning(binning, "ggd_raster_preprocess")) return GGD_E_INVALID;
    if (!means3D || !opacities || !viewmatrix || !projmatrix || !geom_ws) { set_error("ggd_raster_preprocess: null required pointer"); return GGD_E_INVALID; }
    if ((shs == nullptr) == (colors_precomp == nullptr)) { set_error("provide exactly one of shs / colors_precomp"); return GGD_E_INVALID; }
    if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr)) { set_error("provide exactly one of (scales, rotations) / cov3D_precomp"); return GGD_E_INVALID; }
    if (shs && (M < 1 || M > 16 || D < 0 || D > 3 || (D + 1) * (D + 1) > M)) { set_error("bad SH configuration D=%d M=%d", D, M); return GGD_E_INVALID; }
    if (shs && D > 0 && !campos) { set_error("campos required for SH degree > 0"); return GGD_E_INVALID; }
    if (!aligned256(geom_ws)) { set_error("geom_ws must be 256-byte aligned"); return GGD_E_INVALID; }
    if (rotations && (reinterpret_cast<uintptr_t>(rotations) & 15)) { set_error("rotations must be 16-byte aligned"); return GGD_E_INVALID; }

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GeomState g = carve_geom(geom_ws, (size_t)P, tiles_of(W, H));
    if (!radii) radii = g.radii;

    PreprocessArgs a;
    a.tile_counts = binning ? g.tile_counts : nullptr;
    a.counts_valid = g.counters + 12;
    a.diff_used = g.counters + 13;
    a.P = P; a.D = D; a.M = shs ? M : 0; a.W = W; a.H = H;
    a.means3D = means3D; a.shs = shs; a.colors_precomp = colors_precomp; a.opacities = opacities;
    a.scales = scales; a.scale_modifier = scale_modifier; a.rotations = rotations;
    a.cov3D_precomp = cov3D_precomp; a.view = viewmatrix; a.proj = projmatrix; a.campos = campos;
    a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.focal_y = H / (2.0f * tan_fovy);
    a.focal_x = W / (2.0f * tan_fovx);
    a.prefiltered = prefiltered;
    a.grid_x = (W + kTileX - 1) / kTileX;
    a.grid_y = (H + kTileY - 1) / kTileY;

    const size_t zero = (size_t)(reinterpret_cast<char*>(geom_ws) + g.bytes - reinterpret_cast<char*>(g.counters));
    GGD_CUDA_CHECK(cudaMemsetAsync(g.counters, 0, zero, s));
    { StageScope t(s, kStPreprocess); if (int e = launch_preprocess(s, a, g, radii)) return e; }
    { StageScope t(s, kStScan); if (int e = launch_scan(s, P, g)) return e; }
    if (num_rendered_host)
        GGD_CUDA_CHECK(cudaMemcpyAsync(num_rendered_host, g.counters, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    return GGD_OK;
}

int ggd_raster_render(void* stream, int P, int W, int H, const float* background,
                      const float* colors_precomp, const int* radii,
                      void* geom_ws, void* bin_ws, uint64_t bin_capacity, void* img_ws,
                      float* out_color, int binning)
{
    if (int e = check_common(P, W, H)) return e;
    bool lean = false;
    if (!resolve_binning(binning, "ggd_raster_render", &lean)) return GGD_E_INVALID;
    if (!background || !geom_ws || !bin_ws || !img_ws || !out_color) { set_error("ggd_raster_render: null required pointer"); return GGD_E_INVALID; }
    if (!aligned256(geom_ws) || !aligned256(bin_ws) || !aligned256(img_ws)) { set_error("state buffers must be 256-byte aligned"); return GGD_E_INVALID; }
    if (bin_capacity >= (1ull << 30)) { set_error("bin_capacity %llu exceeds 2^30-1", (unsigned long long)bin_capacity); return GGD_E_INVALID; }

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GeomState g = carve_geom(geom_ws, (size_t)P, tiles_of(W, H));
    BinState b = carve_bin(bin_ws, bin_capacity);
    ImgState im = carve_img(img_ws, W, H);
    if (!radii) radii = g.radii;
    const int grid_x = (W + kTileX - 1) / kTileX, grid_y = (H + kTileY - 1) / kTileY;
    const int tiles = grid_x * grid_y;
    const bool per_tile = binning != 0 && bin_capacity > 0;
    const bool radix = !(per_tile && binning == 2);   // enqueue the radix kernels
    // GGD_BINNING_LEAN on a large image: long tile lists are sorted only as far as the blend reads them (on 512^2 frames the
    // second round costs more than the first saves: measured)
    const bool lazy_sort = lean && per_tile && tiles > 4096;
    const int end_bit = 32 + tile_key_bits((uint32_t)tiles);
    const int num_passes = (end_bit + kRadixBits - 1) / kRadixBits;

    SortScratch sc = carve_sort_scratch(b.sort_scratch, bin_capacity);
    if (radix) GGD_CUDA_CHECK(cudaMemsetAsync(b.sort_scratch, 0, sc.zero_bytes, s));
    const uint32_t* skip = per_tile ? g.counters + 4 : nullptr;
    {
        StageScope t(s, kStDuplicate);
        if (per_tile) {   // writes every range, picks the path for this frame (counters[4])
            if (int e = launch_tile_scan(s, g, im, grid_x, grid_y, bin_capacity, P, !radix)) return e;
            if (int e = launch_tile_scatter(s, P, grid_x, grid_y, g, radii, b, im, bin_capacity, !lean)) return e;
        } else {
            GGD_CUDA_CHECK(cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * (size_t)tiles, s));
        }
        if (radix) if (int e = launch_duplicate(s, P, grid_x, grid_y, g, radii, b, bin_capacity, sc, num_passes, skip)) return e;
    }
    if (bin_capacity > 0) {
        { StageScope t(s, kStSort);
          if (per_tile) if (int e = launch_tile_sort(s, g, b, im, tiles, lazy_sort ? 1 : 0)) return e;
          if (radix) if (int e = sort_pairs(s, b.keys_unsorted, b.keys, b.vals_unsorted, b.vals, g.counters + (per_tile ? 6 : 3),
                                            bin_capacity, 0, end_bit, b.keys_tmp, b.vals_tmp, sc, true)) return e; }
        { StageScope t(s, kStRanges); if (radix) if (int e = launch_tile_ranges(s, g, b, bin_capacity, im, tiles, skip)) return e; }
    }
    const float* features = colors_precomp ? colors_precomp : g.rgb;
    StageScope t(s, kStBlend);
    if (!lazy_sort) return launch_blend(s, W, H, grid_x, grid_y, g, b, im, features, background, out_color);
    // lean: long tile lists were sorted only in their nearest depth chunk.  Blend; sort the few lists that did not saturate
    // inside it completely; blend those tiles again.  (All on the device: the second round's kernels skip unmarked tiles.)
    if (int e = launch_blend(s, W, H, grid_x, grid_y, g, b, im, features, background, out_color, 1)) return e;
    if (int e = launch_tile_sort(s, g, b, im, tiles, 2)) return e;
    return launch_blend(s, W, H, grid_x, grid_y, g, b, im, features, background, out_color, 2);
}

int ggd_raster_forward(void* stream, int P, int D, int M, const float* background, int W, int H,
                       const float* means3D, const float* shs, const float* colors_precomp,
                       const float* opacities, const float* scales, float scale_modifier,
                       const float* rotations, const float* cov3D_precomp,
                       const float* viewmatrix, const float* projmatrix, const float* campos,
                       float tan_fovx, float tan_fovy, int prefiltered,
                       void* geom_ws, void* bin_ws, uint64_t bin_capacity, void* img_ws,
                       float* out_color, int* radii, uint32_t* num_rendered_host, int binning)
{
    { int check = binning; if (!resolve_binning(check, "ggd_raster_forward")) return GGD_E_INVALID; }   // both stages resolve it again
    if (int e = ggd_raster_preprocess(stream, P, D, M, W, H, means3D, shs, colors_precomp, opacities,
                                      scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
                                      projmatrix, campos, tan_fovx, tan_fovy, prefiltered, geom_ws,
                                      radii, num_rendered_host, binning)) return e;
    return ggd_raster_render(stream, P, W, H, background, colors_precomp, radii, geom_ws, bin_ws,
                             bin_capacity, img_ws, out_color, binning);
}

int ggd_raster_status(void* stream, const void* geom_ws, int P, uint32_t out[2])
{
    if (!geom_ws || !out || P <= 0) { set_error("ggd_raster_status: bad arguments"); return GGD_E_INVALID; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GeomState g = carve_geom(const_cast<void*>(geom_ws), (size_t)P);
    GGD_CUDA_CHECK(cudaMemcpyAsync(out, g.counters, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    GGD_CUDA_CHECK(cudaStreamSynchronize(s));
    return GGD_OK;
}

int ggd_raster_binning_info(void* stream, const void* geom_ws, int P, uint32_t out[2])
{
    if (!geom_ws || !out || P <= 0) { set_error("ggd_raster_binning_info: bad arguments"); return GGD_E_INVALID; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GeomState g = carve_geom(const_cast<void*>(geom_ws), (size_t)P);
    GGD_CUDA_CHECK(cudaMemcpyAsync(out, g.counters + 4, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    GGD_CUDA_CHECK(cudaStreamSynchronize(s));
    return GGD_OK;
}

int ggd_mark_visible(void* stream, int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present)
{
    if (P <= 0 || !means3D || !viewmatrix || !projmatrix || !present) { set_error("ggd_mark_visible: bad arguments"); return GGD_E_INVALID; }
    return launch_mark_visible(static_cast<cudaStream_t>(stream), P, means3D, viewmatrix, projmatrix, present);
}

int ggd_raster_field(int which, const char* name, int P, int W, int H, uint64_t bin_capacity,
                     size_t* byte_offset, size_t* count)
{
    if (!name || !byte_offset || !count) { set_error("ggd_raster_field: null argument"); return GGD_E_INVALID; }
    char* const base = nullptr;
    auto off = [&](const void* p) { return (size_t)(static_cast<const char*>(p) - base); };
    const size_t Pz = (size_t)std::max(P, 1);
    if (which == 0) {
        GeomState g = carve_geom(base, Pz);
        struct { const char* n; const void* p; size_t c; } tab[] = {
            {"depths", g.depths, Pz}, {"radii", g.radii, Pz}, {"means2D", g.means2D, 2 * Pz},
            {"cov3D", g.cov3D, 6 * Pz}, {"conic_opacity", g.conic_opacity, 4 * Pz}, {"rgb", g.rgb, 3 * Pz},
            {"tiles_touched", g.tiles_touched, Pz}, {"point_offsets", g.point_offsets, Pz},
            {"clamped", g.clamped, 3 * Pz}, {"num_rendered", g.counters, 2}};
        for (auto& e : tab) if (!strcmp(e.n, name)) { *byte_offset = off(e.p); *count = e.c; return GGD_OK; }
    } else if (which == 1) {
        BinState b = carve_bin(base, bin_capacity);
        const size_t n = (size_t)bin_capacity;
        struct { const char* n; const void* p; size_t c; } tab[] = {
            {"point_list", b.vals, n}, {"point_list_unsorted", b.vals_unsorted, n},
            {"point_list_keys", b.keys, n}, {"point_list_keys_unsorted", b.keys_unsorted, n}};
        for (auto& e : tab) if (!strcmp(e.n, name)) { *byte_offset = off(e.p); *count = e.c; return GGD_OK; }
    } else if (which == 2) {
        ImgState im = carve_img(base, W, H);
        const size_t N = (size_t)W * H;
        const size_t tiles = (size_t)((W + kTileX - 1) / kTileX) * ((H + kTileY - 1) / kTileY);
        struct { const char* n; const void* p; size_t c; } tab[] = {
            {"accum_alpha", im.accum_alpha, N}, {"n_contrib", im.n_contrib, N}, {"ranges", im.ranges, 2 * tiles}};
        for (auto& e : tab) if (!strcmp(e.n, name)) { *byte_offset = off(e.p); *count = e.c; return GGD_OK; }
    }
    set_error("ggd_raster_field: unknown field %d/%s", which, name);
    return GGD_E_INVALID;
}

int ggd_stage_timing(int enable) { g_timing.store(enable != 0); return GGD_OK; }

int ggd_stage_times(float ms_out[8])
{
    if (!ms_out) { set_error("ggd_stage_times: null pointer"); return GGD_E_INVALID; }
    StageEvents* e = stage_events();
    for (int i = 0; i < kStCount; ++i) {
        ms_out[i] = -1.f;
        if (e && e->made && e->used[i]) {
            GGD_CUDA_CHECK(cudaEventSynchronize(e->ev[i][1]));
            GGD_CUDA_CHECK(cudaEventElapsedTime(&ms_out[i], e->ev[i][0], e->ev[i][1]));
            e->used[i] = false;   // a stage that does not run in the next frame reads -1, not a stale time
        }
    }
    return GGD_OK;
}

size_t ggd_sort_scratch_bytes(uint64_t capacity) { return carve_sort_scratch(nullptr, capacity).bytes; }

int ggd_sort_pairs(void* stream, const uint64_t* keys_in, uint64_t* keys_out,
                   const uint32_t* vals_in, uint32_t* vals_out,
                   const uint32_t* n_dev, uint64_t capacity, int begin_bit, int end_bit,
                   uint64_t* keys_tmp, uint32_t* vals_tmp, void* scratch, size_t scratch_bytes)
{
    if (!keys_in || !keys_out || !vals_in || !vals_out || !n_dev || !keys_tmp || !vals_tmp || !scratch) { set_error("ggd_sort_pairs: null pointer"); return GGD_E_INVALID; }
    if (capacity >= (1ull << 30)) { set_error("ggd_sort_pairs: capacity %llu exceeds 2^30-1", (unsigned long long)capacity); return GGD_E_INVALID; }
    SortScratch sc = carve_sort_scratch(scratch, capacity);
    if (scratch_bytes < sc.bytes) { set_error("ggd_sort_pairs: scratch too small (%zu < %zu)", scratch_bytes, sc.bytes); return GGD_E_CAPACITY; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GGD_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sc.zero_bytes, s));
    return sort_pairs(s, keys_in, keys_out, vals_in, vals_out, n_dev, capacity, begin_bit, end_bit,
                      keys_tmp, vals_tmp, sc, false);
}

}  // extern "C"
