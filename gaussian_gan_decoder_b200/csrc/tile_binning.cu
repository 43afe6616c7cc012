// Stage 2a, per-tile flavour: bin the instances by tile first, then sort every tile's list by depth
// inside shared memory.
//
// Replaces the same reference code as binning.cu — duplicateWithKeys
// (DGR/cuda_rasterizer/rasterizer_impl.cu:70-111), cub::DeviceRadixSort::SortPairs (:303-308) and
// identifyTileRanges (:116-138) — and produces the same bits: the reference sorts the keys
// (tile << 32 | depth bits) with a STABLE sort whose input order is Gaussian index ascending, so within
// a tile the order is (depth bits, Gaussian index) ascending, and a Gaussian appears at most once per
// tile.  Sorting each tile's (depth bits << 32 | Gaussian index) words is therefore the same permutation.
//
// Why: the global radix sort moves 6 x 24 bytes per instance through HBM in 6 dependent passes, each a
// launch with a chained look-back; at the frame sizes of the headline workload (R ~ 10^6 instances,
// ~800 per tile) that is latency, not bandwidth.  Here an instance is written once into its tile's
// segment and sorted where it lands:
//   preprocess        counts instances per tile (atomics, fused into the projection kernel; rectangles of 32
//                     tiles or more as +1 / -1 at their corners of a 2-D difference array)
//   tile_scan_kernel  sums the difference array (2-D prefix sum), exclusive scan of the counts ->
//                     ranges[tile] = [start, end) ((0,0) for an empty tile, as the reference leaves it); sets up
//                     each tile's write cursor; decides whether the frame qualifies (longest list <=
//                     kTileSortHugeMax, R <= capacity)
//   tile_scatter      slot-centric like duplicate_kernel; claims a slot in the tile's segment with an
//                     atomic on the tile's cursor (CTAs of large splats: one atomic per tile, ranks in shared memory)
//   super_bin / super_scan / super_emit   the scatter of images with more than 4096 tiles, in two levels: Gaussians
//                     are binned by 8x8-tile super-tile, a CTA then claims a list chunk's places per tile in shared memory
//   tile_sort_kernel  one CTA per tile: distribution sort of the 64-bit words in shared memory (bitonic
//                     network for lists with heavily clustered depths)
//   tile_sort_large_kernel  the same for lists of 4097..98304 instances: persistent CTAs of 1024 threads,
//                     one per SM, walk the list of long tiles the scan wrote; lists beyond 24576 words are
//                     sorted in depth chunks
//   tile_sort_lazy_kernel   lean pipelines on large images: only the nearest depth chunk of every long list (the blend
//                     reads 7-16 % of a dense frame's sorted instances), cut at a depth bound estimated from a sample of
//                     the list and gathered in one pass; the tiles whose pixels outlive it are marked by the blend,
//                     sorted completely (tile_sort_large_kernel, lazy = 2) and blended again
// Frames that do not qualify (a tile list longer than kTileSortHugeMax) take the radix path of binning.cu;
// the decision is made on the device (counters[4]) and both sets of kernels are always enqueued, the
// ones that do not apply return at once, so no host round trip is needed.
#include "common.cuh"
#include <cstdlib>

namespace ggd {
namespace {

constexpr int kScanThreadsT = 1024;

__global__ void __launch_bounds__(kScanThreadsT)
tile_scan_kernel(uint32_t* __restrict__ counts, uint2* __restrict__ ranges, int grid_x, int grid_y, uint32_t capacity,
                 const uint32_t* __restrict__ offsets, int P, uint32_t* __restrict__ counters, int strict)
{
    extern __shared__ uint32_t s_diff[];   // [tiles] when the grid has at most kTileDiffMaxTiles tiles
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_max, s_carry, s_large;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int tiles = grid_x * grid_y;
    // Large rectangles were counted as +1 / -1 at their corners (preprocess, word 3 of the tile records): a 2-D inclusive
    // prefix sum of that array is their per-tile count.  Rows by warps (shuffle scan, 32 columns at a time), then columns
    // by threads (consecutive threads on consecutive banks).  Arithmetic is modulo 2^32, the sums are counts.
    const bool diff = counters[13] != 0u;
    if (diff) {
        for (int i = t; i < tiles; i += kScanThreadsT) s_diff[i] = counts[(size_t)i * kTileCountStride + 3];
        __syncthreads();
        for (int r = warp; r < grid_y; r += kScanThreadsT / 32) {
            uint32_t carry = 0;
            for (int x0 = 0; x0 < grid_x; x0 += 32) {
                const int x = x0 + lane;
                uint32_t v = x < grid_x ? s_diff[r * grid_x + x] : 0u;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += n;
                }
                v += carry;
                if (x < grid_x) s_diff[r * grid_x + x] = v;
                carry = __shfl_sync(0xffffffffu, v, 31);
            }
        }
        __syncthreads();
        for (int c = t; c < grid_x; c += kScanThreadsT) {
            uint32_t run = 0;
            for (int y = 0; y < grid_y; ++y) { run += s_diff[y * grid_x + c]; s_diff[y * grid_x + c] = run; }
        }
        __syncthreads();
    }
    auto count_of = [&](int i) { return counts[(size_t)i * kTileCountStride] + (diff ? s_diff[i] : 0u); };
    uint32_t mx = 0;
    for (int i = t; i < tiles; i += kScanThreadsT) mx = max(mx, count_of(i));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_warp[warp] = mx;
    __syncthreads();
    if (t == 0) {
        uint32_t m = 0;
        for (int w = 0; w < kScanThreadsT / 32; ++w) m = max(m, s_warp[w]);
        const uint32_t R = P > 0 ? __ldg(offsets + P - 1) : 0u;
        // counters[12]: preprocess counted the tile lists of THIS frame (it does so only when asked to); without it
        // the counts are the zeros of the frame's memset and must not be taken for an empty frame
        const bool per_tile = counters[12] != 0u && m <= kTileSortHugeMax && R <= capacity;
        s_max = per_tile ? 1u : 0u;
        s_carry = 0;
        s_large = 0;
        // flag 1: more instances than the buffers hold; flag 2: the caller enqueued no radix kernels but this
        // frame needs them.  Either way the image is undefined and the caller runs the stage again.
        counters[1] = ((R > capacity) ? 1u : 0u) | ((strict && !per_tile) ? 2u : 0u);
        counters[3] = min(R, capacity);
        counters[4] = per_tile ? 1u : 0u;
        counters[5] = m;
        counters[6] = per_tile ? 0u : min(R, capacity);   // element count the radix kernels see
    }
    __syncthreads();
    const bool per_tile = s_max != 0;
    if (!per_tile) {   // the radix path's range kernel expects zeros for empty tiles
        for (int i = t; i < tiles; i += kScanThreadsT) ranges[i] = make_uint2(0u, 0u);
        return;
    }
    for (int base = 0; base < tiles; base += kScanThreadsT) {
        const int i = base + t;
        const uint32_t v = i < tiles ? count_of(i) : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        __syncthreads();
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t off = s_carry;
        for (int w = 0; w < warp; ++w) off += s_warp[w];
        if (i < tiles) {
            const uint32_t start = off + incl - v;
            ranges[i] = v ? make_uint2(start, start + v) : make_uint2(0u, 0u);
            counts[(size_t)i * kTileCountStride + 2] = start;   // the tile's write cursor (word 0 keeps the count, so
                                                                // stage 2 can be run again on the same geometry state)
            if (v > kTileSortMax) counts[(size_t)atomicAdd(&s_large, 1u) * kTileCountStride + 1] = (uint32_t)i;
        }
        __syncthreads();
        if (t == kScanThreadsT - 1) s_carry = off + incl;
        __syncthreads();
    }
    if (t == 0) {
        counters[7] = s_large;
        // flag 4: the per-tile counts do not add up to R (state buffers of two different frames were mixed)
        if (s_carry != counters[3]) counters[1] |= 4u;
    }
}

constexpr int kScatterThreads = 256;
// A CTA whose 256 Gaussians emit at least kScatterDenseMin instances (large splats) claims its places in two levels:
// it counts its instances per tile in shared memory, reserves its share of every non-empty tile's segment with ONE
// global atomic per tile, and then hands out the places inside its share with a second shared-memory counter.  With one
// global atomic per instance a trained-like 512^2 frame sends 6.4 M atomics to 1024 cursors (~3300 per CTA over at most
// 1024 tiles).  Costs 8 bytes of shared memory per tile; images with more than 4096 tiles take the direct path.
constexpr uint32_t kScatterDenseMin = 1024;

// Same slot-centric walk as duplicate_kernel (binning.cu): CTA = 256 Gaussians' slot interval, thread =
// slots s, s+256, ...; additionally to the reference-order key/value arrays (kept for the debug
// accessors) every instance claims a place in its tile's segment.
template <bool kRefOrder>
__global__ void __launch_bounds__(kScatterThreads)
tile_scatter_kernel(int P, int grid_x, int grid_y, const float2* __restrict__ means2D,
                    const float* __restrict__ depths, const uint32_t* __restrict__ offsets,
                    const int* __restrict__ radii, uint64_t* __restrict__ keys_unsorted,
                    uint32_t* __restrict__ vals_unsorted, uint64_t* __restrict__ words, uint32_t* __restrict__ cursors,
                    const uint32_t* __restrict__ counters, int hist_tiles)
{
    if (__ldg(counters + 4) == 0u) return;
    extern __shared__ uint32_t s_scatter[];   // [hist_tiles] per-tile counts, then bases | [hist_tiles] per-tile ranks handed out
    __shared__ uint32_t s_end[kScatterThreads];
    __shared__ uint32_t s_depth[kScatterThreads];
    __shared__ uint32_t s_rect[kScatterThreads];
    __shared__ uint32_t s_w[kScatterThreads];
    const int t = threadIdx.x;
    const int base = blockIdx.x * kScatterThreads;
    const int idx = base + t;
    const uint32_t block_start = (base == 0) ? 0u : __ldg(offsets + base - 1);
    uint32_t my_end = 0;
    if (idx < P) {
        my_end = __ldg(offsets + idx);
        const int r = __ldg(radii + idx);
        uint32_t x0 = 0, y0 = 0, w = 0;
        if (r > 0) {
            uint2 rmin, rmax;
            tile_rect(__ldg(means2D + idx), r, grid_x, grid_y, rmin, rmax);
            x0 = rmin.x; y0 = rmin.y; w = rmax.x - rmin.x;
            s_depth[t] = __float_as_uint(__ldg(depths + idx));
        }
        s_rect[t] = x0 | (y0 << 16);
        s_w[t] = w;
    }
    s_end[t] = (idx < P) ? my_end : 0xffffffffu;
    __syncthreads();
    const int last = min(kScatterThreads, P - base) - 1;
    const uint32_t block_end = s_end[last];
    // slot s of this CTA's interval -> (Gaussian g inside the CTA, tile)
    auto locate = [&](uint32_t s, int& g, uint32_t& tile) {
        int lo = 0, hi = last;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int mid = (lo + hi) >> 1;
            if (s_end[mid] > s) hi = mid; else lo = mid + 1;
        }
        g = lo;
        const uint32_t g_start = (g == 0) ? block_start : s_end[g - 1];
        const uint32_t j = s - g_start;
        const uint32_t w = s_w[g];
        const uint32_t rc = s_rect[g];
        tile = ((rc >> 16) + j / w) * (uint32_t)grid_x + (rc & 0xffffu) + j % w;
    };
    const uint32_t n_slots = block_end - block_start;
    if (hist_tiles > 0 && n_slots >= kScatterDenseMin) {
        uint32_t* s_base = s_scatter;
        uint32_t* s_rank = s_scatter + hist_tiles;
        for (int i = t; i < hist_tiles; i += kScatterThreads) { s_base[i] = 0u; s_rank[i] = 0u; }
        __syncthreads();
        // Both passes give every thread a run of CONSECUTIVE slots: one search for the run's first slot, then the walk
        // steps from tile to tile of a rectangle and from Gaussian to Gaussian.  (With slots t, t + 256, ... every slot
        // pays locate()'s 8-step binary search and integer division — twice here: 212 instructions per instance on a
        // trained-like frame, the kernel was instruction-bound.)
        const uint32_t per = (n_slots + kScatterThreads - 1) / kScatterThreads;
        const uint32_t k_begin = min(n_slots, (uint32_t)t * per), k_end = min(n_slots, k_begin + per);
        auto walk = [&](auto&& emit) {
            if (k_begin >= k_end) return;
            uint32_t s = block_start + k_begin;
            int g; uint32_t tile_unused;
            locate(s, g, tile_unused);
            uint32_t g_start = (g == 0) ? block_start : s_end[g - 1];
            uint32_t g_end = s_end[g], w = s_w[g], rc = s_rect[g];
            uint32_t row = (s - g_start) / w, col = (s - g_start) - row * w;
            for (; s < block_start + k_end; ++s) {
                while (s >= g_end) {                   // next Gaussian with instances (s < block_end = s_end[last])
                    ++g; g_end = s_end[g]; w = s_w[g]; rc = s_rect[g]; row = 0; col = 0;
                }
                emit(g, s, ((rc >> 16) + row) * (uint32_t)grid_x + (rc & 0xffffu) + col);
                if (++col == w) { col = 0; ++row; }
            }
        };
        walk([&](int, uint32_t, uint32_t tile) { atomicAdd(&s_base[tile], 1u); });
        __syncthreads();
        // the CTA's base in every tile's segment; four claims in flight per thread (each waits ~700 clocks for its return value)
        for (int i0 = t; i0 < hist_tiles; i0 += 4 * kScatterThreads) {
            uint32_t c[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int i = i0 + u * kScatterThreads; c[u] = i < hist_tiles ? s_base[i] : 0u; }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                b[u] = c[u] ? atomicAdd(cursors + (size_t)(i0 + u * kScatterThreads) * kTileCountStride + 2, c[u]) : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (c[u]) s_base[i0 + u * kScatterThreads] = b[u];
        }
        __syncthreads();
        walk([&](int g, uint32_t s, uint32_t tile) {
            const uint32_t depth = s_depth[g], id = (uint32_t)(base + g);
            if (kRefOrder) {
                keys_unsorted[s] = ((uint64_t)tile << 32) | depth;
                vals_unsorted[s] = id;
            }
            words[s_base[tile] + atomicAdd(&s_rank[tile], 1u)] = ((uint64_t)depth << 32) | id;   // any order inside the share
        });
        return;
    }
    for (uint32_t s = block_start + t; s < block_end; s += kScatterThreads) {
        int g; uint32_t tile;
        locate(s, g, tile);
        const uint32_t depth = s_depth[g], id = (uint32_t)(base + g);
        if (kRefOrder) {   // the reference's duplicateWithKeys output (emission order); nothing downstream reads it here
            keys_unsorted[s] = ((uint64_t)tile << 32) | depth;
            vals_unsorted[s] = id;
        }
        const uint32_t pos = atomicAdd(cursors + (size_t)tile * kTileCountStride + 2, 1u);
        words[pos] = ((uint64_t)depth << 32) | id;
    }
}

// ---- two-level scatter for large images --------------------------------------------------------------------------------
// With 16 384 tiles (2048^2) and ~140 tiles per Gaussian the scatter above sends 2.8e8 atomics to the tiles' cursors and is
// bound by the L2 atomic rate (3.4 ms, 8 % of the issue slots used).  Here Gaussians are first binned by SUPER-TILE
// (kSuper x kSuper tiles, ~6 per Gaussian: 1.2e7 list entries), then a CTA takes a chunk of one super-tile's list, counts its
// instances per tile in shared memory (64 counters), reserves the chunk's share of each tile's segment with ONE global
// atomic per tile and hands out the places with a second shared-memory counter: ~4e5 global atomics per frame.
// The pair lists live in BinState::vals_tmp (scratch of the radix sort, unused on this path; pairs <= instances <= capacity);
// the super-tiles' counters are words 4-6 of the first tile records.  Any order inside a tile's segment is fine: the tile
// sort orders by (depth, Gaussian index).
constexpr int kSuper = 8;                      // tiles per super-tile edge
constexpr int kSuperMax = 1024;                // super-tiles a CTA-local histogram holds
constexpr int kSuperThreads = 256;
constexpr uint32_t kSuperChunk = 1024;         // list entries per unit of work of the emit kernel

__device__ __forceinline__ bool super_rect(const float2* __restrict__ means2D, const int* __restrict__ radii, int idx, int grid_x,
                                           int grid_y, uint2& rmin, uint2& rmax)
{
    const int r = __ldg(radii + idx);
    if (r <= 0) return false;
    tile_rect(__ldg(means2D + idx), r, grid_x, grid_y, rmin, rmax);
    return rmax.x > rmin.x && rmax.y > rmin.y;
}

// pass 1 (kCount) counts the (Gaussian, super-tile) pairs per super-tile; pass 2 writes the Gaussian ids into the lists
template <bool kCount>
__global__ void __launch_bounds__(kSuperThreads)
super_bin_kernel(int P, int grid_x, int grid_y, const float2* __restrict__ means2D, const int* __restrict__ radii,
                 uint32_t* __restrict__ recs, uint32_t* __restrict__ list, const uint32_t* __restrict__ counters)
{
    if (__ldg(counters + 4) == 0u) return;
    __shared__ uint32_t s_cnt[kSuperMax], s_rank[kSuperMax];
    const int t = threadIdx.x, idx = blockIdx.x * kSuperThreads + t;
    const int sgx = (grid_x + kSuper - 1) / kSuper, ns = sgx * ((grid_y + kSuper - 1) / kSuper);
    for (int i = t; i < ns; i += kSuperThreads) { s_cnt[i] = 0u; s_rank[i] = 0u; }
    __syncthreads();
    uint2 rmin, rmax;
    const bool live = idx < P && super_rect(means2D, radii, idx, grid_x, grid_y, rmin, rmax);
    uint32_t sx0 = 0, sx1 = 0, sy0 = 0, sy1 = 0;
    if (live) { sx0 = rmin.x / kSuper; sx1 = (rmax.x - 1) / kSuper + 1; sy0 = rmin.y / kSuper; sy1 = (rmax.y - 1) / kSuper + 1; }
    for (uint32_t sy = sy0; sy < sy1; ++sy)
        for (uint32_t sx = sx0; sx < sx1; ++sx) atomicAdd(&s_cnt[sy * sgx + sx], 1u);
    __syncthreads();
    for (int i = t; i < ns; i += kSuperThreads) {
        const uint32_t c = s_cnt[i];
        if (c) s_cnt[i] = atomicAdd(recs + (size_t)i * kTileCountStride + (kCount ? 4 : 6), c);   // count, or the CTA's base from the cursor
    }
    if (kCount) return;
    __syncthreads();
    for (uint32_t sy = sy0; sy < sy1; ++sy)
        for (uint32_t sx = sx0; sx < sx1; ++sx) {
            const uint32_t st = sy * sgx + sx;
            list[s_cnt[st] + atomicAdd(&s_rank[st], 1u)] = (uint32_t)idx;
        }
}

// exclusive scan of the super-tiles' counts -> list starts (word 5) and write cursors (word 6); the counts go back to zero
// so that stage 2 can run again on the same geometry state; counters[14] = number of pairs
__global__ void __launch_bounds__(kSuperMax)
super_scan_kernel(int ns, uint32_t* __restrict__ recs, uint32_t* __restrict__ counters)
{
    if (counters[4] == 0u) return;
    __shared__ uint32_t s_warp[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t v = t < ns ? recs[(size_t)t * kTileCountStride + 4] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    if (t < ns) {
        recs[(size_t)t * kTileCountStride + 4] = 0u;
        recs[(size_t)t * kTileCountStride + 5] = off + incl - v;
        recs[(size_t)t * kTileCountStride + 6] = off + incl - v;
    }
    if (t == kSuperMax - 1) counters[14] = off + incl;
}

// Visits the tiles [x0, x1) x [y0, y1) row by row.  (The large splats of a chunk cover the same tiles, so the lanes of a warp
// mostly visit the same tile in the same step: same-address shared-memory atomics are merged by the hardware, and the words
// of a step land next to each other in that tile's segment.  Starting every lane at a different tile was measured: slower,
// and so was claiming four places before storing the first word.)
template <typename F>
__device__ __forceinline__ void walk_rect(uint32_t x0, uint32_t x1, uint32_t y0, uint32_t y1, F&& visit)
{
    for (uint32_t ty = y0; ty < y1; ++ty)
        for (uint32_t tx = x0; tx < x1; ++tx) visit(tx, ty);
}

template <bool kRefOrder>
__global__ void __launch_bounds__(kSuperThreads, 8)
super_emit_kernel(int grid_x, int grid_y, const float2* __restrict__ means2D, const float* __restrict__ depths,
                  const uint32_t* __restrict__ offsets, const int* __restrict__ radii, uint64_t* __restrict__ keys_unsorted,
                  uint32_t* __restrict__ vals_unsorted, uint64_t* __restrict__ words, uint32_t* __restrict__ recs,
                  const uint32_t* __restrict__ list, const uint32_t* __restrict__ counters)
{
    if (__ldg(counters + 4) == 0u) return;
    __shared__ uint32_t s_start[kSuperMax + 1];
    __shared__ uint32_t s_cnt[kSuper * kSuper], s_rank[kSuper * kSuper];
    const int t = threadIdx.x;
    const int sgx = (grid_x + kSuper - 1) / kSuper, ns = sgx * ((grid_y + kSuper - 1) / kSuper);
    const uint32_t pairs = __ldg(counters + 14);
    for (int i = t; i < ns; i += kSuperThreads) s_start[i] = recs[(size_t)i * kTileCountStride + 5];
    if (t == 0) s_start[ns] = pairs;
    __syncthreads();
    const uint32_t chunks = (pairs + kSuperChunk - 1) / kSuperChunk;
    for (uint32_t ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
        uint32_t a = ch * kSuperChunk;
        const uint32_t chunk_end = min(pairs, a + kSuperChunk);
        // super-tile of the chunk's first entry: the last one whose list starts at or before it
        int st = 0;
        {
            int lo = 0, hi = ns - 1;
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_start[mid] <= a) lo = mid; else hi = mid - 1; }
            st = lo;
        }
        while (a < chunk_end) {                                       // one segment per super-tile the chunk spans
            while (s_start[st + 1] <= a) ++st;                        // (skips empty lists)
            const uint32_t b = min(chunk_end, s_start[st + 1]);
            const uint32_t tx_lo = (uint32_t)(st % sgx) * kSuper, ty_lo = (uint32_t)(st / sgx) * kSuper;
            const uint32_t tx_hi = min((uint32_t)grid_x, tx_lo + kSuper), ty_hi = min((uint32_t)grid_y, ty_lo + kSuper);
            if (t < kSuper * kSuper) { s_cnt[t] = 0u; s_rank[t] = 0u; }
            __syncthreads();
            // this thread's entries: a + t, a + t + 256, ... (at most kSuperChunk / kSuperThreads of them)
            for (uint32_t e = a + t; e < b; e += kSuperThreads) {
                const int g = (int)__ldg(list + e);
                uint2 rmin, rmax;
                super_rect(means2D, radii, g, grid_x, grid_y, rmin, rmax);
                const uint32_t x0 = max(rmin.x, tx_lo), x1 = min(rmax.x, tx_hi), y0 = max(rmin.y, ty_lo), y1 = min(rmax.y, ty_hi);
                walk_rect(x0, x1, y0, y1, [&](uint32_t tx, uint32_t ty) { atomicAdd(&s_cnt[(ty - ty_lo) * kSuper + (tx - tx_lo)], 1u); });
            }
            __syncthreads();
            if (t < kSuper * kSuper) {
                const uint32_t c = s_cnt[t];
                const uint32_t tx = tx_lo + (uint32_t)(t % kSuper), ty = ty_lo + (uint32_t)(t / kSuper);
                if (c) s_cnt[t] = atomicAdd(recs + (size_t)(ty * grid_x + tx) * kTileCountStride + 2, c);   // base in the tile's segment
            }
            __syncthreads();
            for (uint32_t e = a + t; e < b; e += kSuperThreads) {
                const int g = (int)__ldg(list + e);
                uint2 rmin, rmax;
                super_rect(means2D, radii, g, grid_x, grid_y, rmin, rmax);
                const uint32_t depth = __float_as_uint(__ldg(depths + g));
                const uint32_t x0 = max(rmin.x, tx_lo), x1 = min(rmax.x, tx_hi), y0 = max(rmin.y, ty_lo), y1 = min(rmax.y, ty_hi);
                const uint32_t w = rmax.x - rmin.x;
                const uint32_t g_start = kRefOrder ? (g == 0 ? 0u : __ldg(offsets + g - 1)) : 0u;
                walk_rect(x0, x1, y0, y1, [&](uint32_t tx, uint32_t ty) {
                    const uint32_t local = (ty - ty_lo) * kSuper + (tx - tx_lo);
                    const uint32_t tile = ty * (uint32_t)grid_x + tx;
                    if (kRefOrder) {     // the reference's emission order: row-major over the Gaussian's rectangle
                        const uint32_t slot = g_start + (ty - rmin.y) * w + (tx - rmin.x);
                        keys_unsorted[slot] = ((uint64_t)tile << 32) | depth;
                        vals_unsorted[slot] = (uint32_t)g;
                    }
                    words[s_cnt[local] + atomicAdd(&s_rank[local], 1u)] = ((uint64_t)depth << 32) | (uint32_t)g;
                });
            }
            __syncthreads();
            a = b;
        }
    }
}

constexpr int kTileSortThreads = 256;

// Bitonic sort of n_pad = 256 * E words, E consecutive words per thread ("blocked"): compare distances
// below E stay inside a thread's registers, distances below 32 E are warp shuffles, and only the three
// largest distances (6 of the network's stages, whatever n_pad is) go through shared memory.
template <int E, bool kFromShared = false>      // kFromShared: the words already wait in s[0..n)
__device__ __forceinline__ void sort_tile(uint64_t* __restrict__ s, const uint64_t* __restrict__ words, uint32_t begin,
                                          uint32_t n, uint32_t tile, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    constexpr uint32_t n_pad = kTileSortThreads * E;
    const int t = threadIdx.x;
    for (uint32_t i = t; i < n_pad; i += kTileSortThreads) {
        if (i >= n) s[i] = ~0ull;
        else if (!kFromShared) s[i] = __ldg(words + begin + i);
    }
    __syncthreads();
    uint64_t v[E];
#pragma unroll
    for (int r = 0; r < E; ++r) v[r] = s[t * E + r];
    const uint32_t e0 = (uint32_t)t * E;
#pragma unroll 1
    for (uint32_t k = 2; k <= n_pad; k <<= 1) {
#pragma unroll 1
        for (uint32_t j = k >> 1; j >= (uint32_t)E; j >>= 1) {   // partner word lives in another thread
            const uint32_t jt = j / E;                            // distance in threads
            const bool take_min = (((e0 & j) == 0) == ((e0 & k) == 0));
            if (jt >= 32) {
                __syncthreads();
#pragma unroll
                for (int r = 0; r < E; ++r) s[t * E + r] = v[r];
                __syncthreads();
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const uint64_t o = s[(t ^ jt) * E + r];
                    v[r] = take_min ? min(v[r], o) : max(v[r], o);
                }
            } else {
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const uint64_t o = __shfl_xor_sync(0xffffffffu, v[r], jt);
                    v[r] = take_min ? min(v[r], o) : max(v[r], o);
                }
            }
        }
        // distances below E: inside the thread.  Direction is per word for k <= E / 2, per thread above.
#pragma unroll
        for (int j = E >> 1; j > 0; j >>= 1) {
            if ((uint32_t)j < k) {
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    if ((r & j) == 0) {
                        const bool up = ((e0 + r) & k) == 0;
                        const uint64_t a = v[r], b = v[r | j];
                        const uint64_t lo = min(a, b), hi = max(a, b);
                        v[r] = up ? lo : hi;
                        v[r | j] = up ? hi : lo;
                    }
                }
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < E; ++r) s[t * E + r] = v[r];
    __syncthreads();
    for (uint32_t i = t; i < n; i += kTileSortThreads) {
        const uint64_t w = s[i];
        vals[begin + i] = (uint32_t)w;
        keys[begin + i] = ((uint64_t)tile << 32) | (w >> 32);
    }
}

// Distribution sort of one tile list (n <= 256 E words, thread t holds words t, t + 256, ...): a monotone map
// of the list's depth interval onto 256 E buckets (BucketMap below) leaves about one word per bucket.  Words are dropped into their bucket (shared-memory atomics
// hand out the rank inside the bucket), bucket starts come from one scan, and the thread owning a bucket
// orders its few words by insertion.  O(n) instead of the bitonic network's O(n log^2 n); a list with an
// overfull bucket (many equal or clustered depths) falls back to the network.
constexpr uint32_t kMaxBucket = 16;

// Monotone map of a list's depth bits onto n_b buckets that evens out the bucket loads: a coarse histogram
// (kCoarse bins, linear between the list's min and max) gives a piecewise-linear estimate of the depth CDF,
// and a word goes to bucket  n_b / n * (words below its coarse bin + its fraction of that bin).  Depths on a
// head cluster at the front and the back surface; a plain linear map would pile them into a few buckets.
constexpr int kCoarse = 128;
struct BucketMap {
    uint32_t dmin;
    float scale_c, ratio;            // coarse bins per depth unit, n_b / n
    const uint32_t* cstart;          // [kCoarse + 1] exclusive prefix of the coarse counts (shared memory)
    uint32_t n_b;
    __device__ __forceinline__ uint32_t coarse(uint32_t d) const
    {
        return min((uint32_t)(kCoarse - 1), (uint32_t)((float)(d - dmin) * scale_c));
    }
    __device__ __forceinline__ uint32_t operator()(uint32_t d) const
    {
        const float x = (float)(d - dmin) * scale_c;
        const uint32_t kc = min((uint32_t)(kCoarse - 1), (uint32_t)x);
        const float below = (float)cstart[kc], in_bin = (float)(cstart[kc + 1] - cstart[kc]);
        return min(n_b - 1u, (uint32_t)((below + (x - (float)kc) * in_bin) * ratio));
    }
};
// exclusive scan of s_c[0..kCoarse) into s_cstart[0..kCoarse], by warp 0 (call with all threads, then sync)
__device__ __forceinline__ void coarse_scan(const uint32_t* __restrict__ s_c, uint32_t* __restrict__ s_cstart)
{
    if (threadIdx.x < 32) {
        constexpr int kPerLane = kCoarse / 32;
        const int lane = threadIdx.x;
        uint32_t v[kPerLane], sum = 0;
#pragma unroll
        for (int q = 0; q < kPerLane; ++q) { v[q] = s_c[lane * kPerLane + q]; sum += v[q]; }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        uint32_t run = incl - sum;
#pragma unroll
        for (int q = 0; q < kPerLane; ++q) { s_cstart[lane * kPerLane + q] = run; run += v[q]; }
        if (lane == 31) s_cstart[kCoarse] = run;
    }
}

// largest c in [1, kCoarse] with cstart[c] <= limit (cstart: the exclusive prefix above, non-decreasing); 1 if there is none
__device__ __forceinline__ uint32_t bins_below(const uint32_t* __restrict__ cstart, uint32_t limit)
{
    uint32_t lo = 1, hi = kCoarse;
    while (lo < hi) {
        const uint32_t mid = (lo + hi + 1) >> 1;
        if (cstart[mid] <= limit) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// kFromShared: the n words wait in s_out[0..n) (gathered there by the caller) instead of words[begin..]
template <int E, bool kFromShared = false>
__device__ __forceinline__ void dist_sort_tile(uint64_t* __restrict__ s_out, uint32_t* __restrict__ s_hist,
                                               const uint64_t* __restrict__ words, uint32_t begin, uint32_t n, uint32_t tile,
                                               uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    constexpr uint32_t n_b = kTileSortThreads * E;
    __shared__ uint32_t s_red[2][kTileSortThreads / 32];
    __shared__ uint32_t s_part[kTileSortThreads / 32];
    __shared__ uint32_t s_c[kCoarse], s_cstart[kCoarse + 1];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    uint64_t w[E];
    uint32_t dmin = 0xffffffffu, dmax = 0u;
    if (t < kCoarse) s_c[t] = 0;
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint32_t i = t + r * kTileSortThreads;
        w[r] = i < n ? (kFromShared ? s_out[i] : __ldg(words + begin + i)) : ~0ull;
        if (i < n) { dmin = min(dmin, (uint32_t)(w[r] >> 32)); dmax = max(dmax, (uint32_t)(w[r] >> 32)); }
        s_hist[i] = 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        dmin = min(dmin, __shfl_xor_sync(0xffffffffu, dmin, o));
        dmax = max(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    }
    if (lane == 0) { s_red[0][warp] = dmin; s_red[1][warp] = dmax; }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kTileSortThreads / 32; ++k) { dmin = min(dmin, s_red[0][k]); dmax = max(dmax, s_red[1][k]); }
    BucketMap map{dmin, (float)kCoarse / ((float)(dmax - dmin) + 1.0f), (float)n_b / (float)n, s_cstart, n_b};
#pragma unroll
    for (int r = 0; r < E; ++r)
        if (t + r * kTileSortThreads < n) atomicAdd(&s_c[map.coarse((uint32_t)(w[r] >> 32))], 1u);
    __syncthreads();
    coarse_scan(s_c, s_cstart);
    __syncthreads();
    uint32_t bk[E];   // bucket | rank inside the bucket << 16
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const uint32_t i = t + r * kTileSortThreads;
        bk[r] = 0;
        if (i < n) {
            const uint32_t b = map((uint32_t)(w[r] >> 32));
            bk[r] = b | (atomicAdd(&s_hist[b], 1u) << 16);
        }
    }
    __syncthreads();
    // exclusive scan of the bucket counts; thread t owns buckets [t E, t E + E)
    uint32_t cnt[E], sum = 0, big = 0;
#pragma unroll
    for (int q = 0; q < E; ++q) { cnt[q] = s_hist[t * E + q]; sum += cnt[q]; big = max(big, cnt[q]); }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_part[warp] = incl;
    const bool overfull = __syncthreads_or(big > kMaxBucket);
    uint32_t start = incl - sum;
#pragma unroll
    for (int k = 0; k < kTileSortThreads / 32; ++k) if (k < warp) start += s_part[k];
    if (overfull) {   // uniform for the CTA
        sort_tile<E, kFromShared>(s_out, words, begin, n, tile, keys, vals);
        return;
    }
    {
        uint32_t run = start;
#pragma unroll
        for (int q = 0; q < E; ++q) { s_hist[t * E + q] = run; run += cnt[q]; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < E; ++r)
        if (t + r * kTileSortThreads < n) s_out[s_hist[bk[r] & 0xffffu] + (bk[r] >> 16)] = w[r];
    __syncthreads();
    {
        uint32_t b0 = start;
#pragma unroll
        for (int q = 0; q < E; ++q) {
            const uint32_t m = cnt[q];
            for (uint32_t i = 1; i < m; ++i) {   // insertion sort of bucket [b0, b0 + m)
                const uint64_t x = s_out[b0 + i];
                uint32_t j = i;
                while (j > 0 && s_out[b0 + j - 1] > x) { s_out[b0 + j] = s_out[b0 + j - 1]; --j; }
                s_out[b0 + j] = x;
            }
            b0 += m;
        }
    }
    __syncthreads();
    for (uint32_t i = t; i < n; i += kTileSortThreads) {
        const uint64_t x = s_out[i];
        vals[begin + i] = (uint32_t)x;
        keys[begin + i] = ((uint64_t)tile << 32) | (x >> 32);
    }
}

__global__ void __launch_bounds__(kTileSortThreads)
tile_sort_kernel(const uint2* __restrict__ ranges, const uint64_t* __restrict__ words,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, const uint32_t* __restrict__ counters)
{
    if (__ldg(counters + 4) == 0u) return;
    extern __shared__ __align__(16) uint8_t s_dyn[];
    uint64_t* s = reinterpret_cast<uint64_t*>(s_dyn);                              // [kTileSortMax]
    uint32_t* h = reinterpret_cast<uint32_t*>(s_dyn + sizeof(uint64_t) * kTileSortMax);   // [kTileSortMax]
    const uint32_t tile = blockIdx.x;
    const uint2 range = ranges[tile];
    const uint32_t n = range.y - range.x;
    if (n == 0 || n > kTileSortMax) return;   // long lists: tile_sort_large_kernel
    if (n <= 256) dist_sort_tile<1>(s, h, words, range.x, n, tile, keys, vals);
    else if (n <= 512) dist_sort_tile<2>(s, h, words, range.x, n, tile, keys, vals);
    else if (n <= 1024) dist_sort_tile<4>(s, h, words, range.x, n, tile, keys, vals);
    else if (n <= 2048) dist_sort_tile<8>(s, h, words, range.x, n, tile, keys, vals);
    else dist_sort_tile<16>(s, h, words, range.x, n, tile, keys, vals);
}

// First round of a lazily sorted frame (lean pipelines on large images, see tile_sort_large_kernel's `lazy`): for every long
// list, the nearest depth chunk — a depth prefix of about kLazyWords words — is gathered into shared memory and sorted there
// with the small lists' routine; its length goes to word 7 of the tile's record.  ONE pass over the whole list:
//   1. a systematic sample (32 consecutive words out of every 32 S, at most kTileSortMax of them) gives the depth interval
//      and a coarse histogram, hence a depth bound below which about kLazyGather words are expected;
//   2. the pass over the list gathers every word below that bound into shared memory (room for kTileSortMax);
//   3. if more than kLazyWords arrived, an exact coarse histogram of the gathered words (shared memory only) trims them to
//      the leading bins that hold at most kLazyWords.
// Both cuts are monotone in the depth ("depth < bound", "coarse bin of the depth < c"), so the chunk is a prefix of the sorted
// list, ties included.  (Before: three passes over the list — interval, histogram, gather — and one shared-memory atomic per
// gathered word: 1.26 ms on the 2048^2 / 2 M frame, now 1.0.)
// 256-thread CTAs, four per SM, so that several lists are in flight on an SM.  A list whose gather overflows the buffer, or
// whose first bin alone holds more than kLazyWords words, gets length 0: the blend then marks the tile and the second round
// sorts it completely.
constexpr uint32_t kLazyWords = 2048;      // 2048^2 / 2 M frame: 1024 -> 12.0 ms (too many second rounds), 2048 -> 7.1, 4096 -> 7.3, 8192 -> 7.9
constexpr uint32_t kLazyGather = 2816;     // expected words of the gather: kLazyWords + head room for the trim, 1280 below the buffer's end for the sample's error
static_assert(kLazyWords == 8 * kTileSortThreads, "tile_sort_lazy_kernel sorts its chunk with dist_sort_tile<8>");
static_assert(2 * kLazyWords == kTileSortMax, "the trim compacts into the histogram array, which holds kTileSortMax / 2 words");
__global__ void __launch_bounds__(kTileSortThreads, 4)
tile_sort_lazy_kernel(const uint2* __restrict__ ranges, const uint64_t* __restrict__ words, uint64_t* __restrict__ keys,
                      uint32_t* __restrict__ vals, const uint32_t* __restrict__ counters, uint32_t* __restrict__ recs)
{
    if (__ldg(counters + 4) == 0u) return;
    extern __shared__ __align__(16) uint8_t s_dyn[];
    uint64_t* s = reinterpret_cast<uint64_t*>(s_dyn);                              // [kTileSortMax]
    uint32_t* h = reinterpret_cast<uint32_t*>(s_dyn + sizeof(uint64_t) * kTileSortMax);   // [kTileSortMax]
    uint64_t* h64 = reinterpret_cast<uint64_t*>(h);                                // [kTileSortMax / 2] during the trim
    __shared__ uint32_t s_red[2][kTileSortThreads / 32];
    __shared__ uint32_t s_c[kCoarse], s_cstart[kCoarse + 1];
    __shared__ uint32_t s_n;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t n_large = __ldg(counters + 7);
    constexpr int kWalk = 8;                 // loads in flight per thread
    // CTA-wide min / max of a depth pair (every thread gets the result); also clears the coarse bins and the gather counter
    auto interval = [&](uint32_t& dmin, uint32_t& dmax) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            dmin = min(dmin, __shfl_xor_sync(0xffffffffu, dmin, o));
            dmax = max(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
        }
        if (lane == 0) { s_red[0][warp] = dmin; s_red[1][warp] = dmax; }
        if (t < kCoarse) s_c[t] = 0;
        if (t == 0) s_n = 0;
        __syncthreads();
        for (int k = 0; k < kTileSortThreads / 32; ++k) { dmin = min(dmin, s_red[0][k]); dmax = max(dmax, s_red[1][k]); }
    };
    for (uint32_t li = blockIdx.x; li < n_large; li += gridDim.x) {
        const uint32_t tile = recs[(size_t)li * kTileCountStride + 1];
        const uint2 range = ranges[tile];
        const uint32_t n = range.y - range.x;
        const uint64_t* src = words + range.x;
        auto fail = [&]() {                           // (uniform) left to the second round
            if (t == 0) recs[(size_t)tile * kTileCountStride + 7] = 0u;
            __syncthreads();
        };
        // ---- 1. the sample: depth bits of words [32 S j, 32 S j + 32), j = 0, 1, ... into h ----
        const uint32_t S = max(4u, (n + kTileSortMax - 1) / kTileSortMax);
        const uint32_t full = n / (32u * S), rest = min(32u, n - full * 32u * S);
        const uint32_t n_s = full * 32u + rest;       // <= kTileSortMax
        uint32_t dmin = 0xffffffffu, dmax = 0u;
        for (uint32_t v0 = t; v0 < n_s; v0 += kWalk * kTileSortThreads) {
            uint32_t d[kWalk];
#pragma unroll
            for (int k = 0; k < kWalk; ++k) {
                const uint32_t v = v0 + (uint32_t)k * kTileSortThreads;
                d[k] = v < n_s ? (uint32_t)(__ldg(src + (size_t)(v >> 5) * 32u * S + (v & 31u)) >> 32) : 0u;
            }
#pragma unroll
            for (int k = 0; k < kWalk; ++k) {
                const uint32_t v = v0 + (uint32_t)k * kTileSortThreads;
                if (v < n_s) { h[v] = d[k]; dmin = min(dmin, d[k]); dmax = max(dmax, d[k]); }
            }
        }
        interval(dmin, dmax);
        const float span = (float)(dmax - dmin) + 1.0f, scale_c = (float)kCoarse / span;
        for (uint32_t v = t; v < n_s; v += kTileSortThreads)
            atomicAdd(&s_c[min((uint32_t)(kCoarse - 1), (uint32_t)((float)(h[v] - dmin) * scale_c))], 1u);
        __syncthreads();
        coarse_scan(s_c, s_cstart);
        __syncthreads();
        // sampled words wanted below the bound -> number of leading bins -> the bound in depth bits (any bound cuts a prefix)
        const uint32_t c1 = bins_below(s_cstart, (uint32_t)((float)kLazyGather * (float)n_s / (float)n));
        const uint32_t bound = c1 == (uint32_t)kCoarse ? 0xffffffffu : dmin + (uint32_t)((float)c1 * span * (1.0f / (float)kCoarse));
        // ---- 2. one pass over the list: gather the words below the bound (places claimed once per warp and step) ----
        for (uint32_t base = 0; base < n; base += kWalk * kTileSortThreads) {
            uint64_t w[kWalk];
#pragma unroll
            for (int k = 0; k < kWalk; ++k) {
                const uint32_t i = base + (uint32_t)(k * kTileSortThreads + t);
                w[k] = i < n ? __ldg(src + i) : ~0ull;
            }
            uint32_t cnt = 0;
#pragma unroll
            for (int k = 0; k < kWalk; ++k) cnt += (uint32_t)(w[k] >> 32) < bound ? 1u : 0u;
            uint32_t incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += u;
            }
            uint32_t at = 0;
            if (lane == 31) at = atomicAdd(&s_n, incl);
            at = __shfl_sync(0xffffffffu, at, 31) + incl - cnt;
#pragma unroll
            for (int k = 0; k < kWalk; ++k)
                if ((uint32_t)(w[k] >> 32) < bound) {
                    if (at < kTileSortMax) s[at] = w[k];
                    ++at;
                }
        }
        __syncthreads();
        uint32_t m = s_n;
        __syncthreads();                              // (interval() clears s_n: below, or at the next list)
        if (m > kTileSortMax || m == 0) { fail(); continue; }
        if (m > kLazyWords) {
            // ---- 3. trim to the leading bins of an exact histogram of the gathered words ----
            uint32_t gmin = 0xffffffffu, gmax = 0u;
            for (uint32_t i = t; i < m; i += kTileSortThreads) { const uint32_t d = (uint32_t)(s[i] >> 32); gmin = min(gmin, d); gmax = max(gmax, d); }
            interval(gmin, gmax);
            const float scale_g = (float)kCoarse / ((float)(gmax - gmin) + 1.0f);
            auto coarse_g = [&](uint64_t w) { return min((uint32_t)(kCoarse - 1), (uint32_t)((float)((uint32_t)(w >> 32) - gmin) * scale_g)); };
            for (uint32_t i = t; i < m; i += kTileSortThreads) atomicAdd(&s_c[coarse_g(s[i])], 1u);
            __syncthreads();
            coarse_scan(s_c, s_cstart);
            __syncthreads();
            const uint32_t c2 = bins_below(s_cstart, kLazyWords), m2 = s_cstart[c2];
            if (m2 > kLazyWords) { fail(); continue; }         // the first bin alone is too long for this kernel
            for (uint32_t base = 0; base < m; base += kTileSortThreads) {
                const uint32_t i = base + t;
                const uint64_t w = i < m ? s[i] : 0ull;
                const bool keep = i < m && coarse_g(w) < c2;
                const uint32_t mask = __ballot_sync(0xffffffffu, keep);
                uint32_t at = 0;
                if (lane == 0 && mask) at = atomicAdd(&s_n, (uint32_t)__popc(mask));
                at = __shfl_sync(0xffffffffu, at, 0);
                if (keep) h64[at + __popc(mask & ((1u << lane) - 1u))] = w;
            }
            __syncthreads();
            for (uint32_t i = t; i < m2; i += kTileSortThreads) s[i] = h64[i];
            m = m2;
            __syncthreads();
        }
        if (m <= 256) dist_sort_tile<1, true>(s, h, nullptr, range.x, m, tile, keys, vals);
        else if (m <= 512) dist_sort_tile<2, true>(s, h, nullptr, range.x, m, tile, keys, vals);
        else if (m <= 1024) dist_sort_tile<4, true>(s, h, nullptr, range.x, m, tile, keys, vals);
        else dist_sort_tile<8, true>(s, h, nullptr, range.x, m, tile, keys, vals);       // m <= kLazyWords = 2048
        if (t == 0) recs[(size_t)tile * kTileCountStride + 7] = m;
        __syncthreads();
    }
}

// Long tile lists (kTileSortMax < n <= kTileSortLargeMax): same distribution sort, but the list stays in
// global memory (L2) for two streaming passes instead of registers, one CTA of 1024 threads owns a whole
// SM's shared memory, and an overfull bucket is sorted in place by the CTA with a comparison network.
constexpr int kLargeThreads = 1024;
constexpr uint32_t kLargeBuckets = 8192;
constexpr uint32_t kLargeInsertMax = 32;   // buckets up to this size are ordered by their owner thread
// Lists up to this length are copied into the upper half of the CTA's shared memory first, so that the four passes
// over the list (depth interval, coarse histogram, bucket histogram, distribution) read shared memory instead of
// making four dependent round trips to L2 (the trained-like 512^2 frame: ~1000 lists of 5-10k words, 0.15 ms -> see
// DESIGN.md); longer lists stream from global memory as before.
constexpr uint32_t kLargeStagedMax = kTileSortLargeMax / 2;

// ascending bitonic network on s[0..m) by the whole CTA ("normalised" form: every compare-exchange puts the
// smaller word first, the first step of a phase mirrors, so positions >= m can be treated as +infinity)
__device__ void cta_sort_range(uint64_t* __restrict__ s, uint32_t m)
{
    uint32_t m_pad = 2;
    while (m_pad < m) m_pad <<= 1;
    for (uint32_t k = 2; k <= m_pad; k <<= 1) {
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < (m_pad >> 1); i += kLargeThreads) {
                const uint32_t lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                const uint32_t hi = (j == (k >> 1)) ? (lo ^ (k - 1)) : (lo | j);   // mirrored partner in the first step
                if (hi < m) {
                    const uint64_t a = s[lo], b = s[hi];
                    if (a > b) { s[lo] = b; s[hi] = a; }
                }
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(kLargeThreads, 1)
tile_sort_large_kernel(const uint2* __restrict__ ranges, const uint64_t* __restrict__ words, uint64_t* __restrict__ keys,
                       uint32_t* __restrict__ vals, const uint32_t* __restrict__ counters, uint32_t* __restrict__ tile_list,
                       uint32_t* __restrict__ status_flags, const int lazy)
{
    // lazy = 0: every long list.  lazy = 2: second round of a lazily sorted frame (lean pipelines on large images; the first
    // round is tile_sort_lazy_kernel): only the tiles the blend marked (word 8 of the tile's record) because it reached the
    // end of the sorted chunk with live pixels; they are then blended again.
    if (__ldg(counters + 4) == 0u) return;
    extern __shared__ __align__(16) uint8_t s_dyn[];
    uint64_t* s_out = reinterpret_cast<uint64_t*>(s_dyn);                                      // [kTileSortLargeMax]
    uint32_t* s_hist = reinterpret_cast<uint32_t*>(s_dyn + sizeof(uint64_t) * kTileSortLargeMax);   // [kLargeBuckets]
    __shared__ uint32_t s_red[2][kLargeThreads / 32];
    __shared__ uint32_t s_part[kLargeThreads / 32];
    __shared__ uint32_t s_big[64], s_nbig;
    __shared__ uint32_t s_c[kCoarse], s_cstart[kCoarse + 1];
    constexpr int kPer = kLargeBuckets / kLargeThreads;   // buckets per thread
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t n_large = __ldg(counters + 7);
    for (uint32_t li = blockIdx.x; li < n_large; li += gridDim.x) {
        const uint32_t tile = tile_list[(size_t)li * kTileCountStride + 1];
        if (lazy == 2 && tile_list[(size_t)tile * kTileCountStride + 8] == 0u) continue;
        const uint2 range = ranges[tile];
        const uint32_t n = range.y - range.x;
        const uint64_t* gsrc = words + range.x;
        bool staged = n <= kLargeStagedMax;
        uint32_t n_walk = n;          // words the passes walk: the list, or (lazy round) its nearest chunk once gathered
        uint64_t* s_in = s_out + kLargeStagedMax;
        // One pass over the list: a thread's words are fetched kWalk at a time BEFORE any of them is used.  The passes
        // below update shared memory (atomics) per word; written as load-use-load-use a streamed list cost one L2 round
        // trip per word and thread — 17 to 24 in a row for the 2048^2 frame's lists — instead of three.
        constexpr int kWalk = 8;
        auto walk = [&](auto&& use) {
            if (staged) {      // the list sits in shared memory: nothing to batch
                for (uint32_t i = t; i < n_walk; i += kLargeThreads) use(s_in[i]);
                return;
            }
            for (uint32_t i0 = t; i0 < n_walk; i0 += kWalk * kLargeThreads) {
                uint64_t w[kWalk];
#pragma unroll
                for (int k = 0; k < kWalk; ++k) {
                    const uint32_t i = i0 + (uint32_t)k * kLargeThreads;
                    w[k] = i < n_walk ? __ldg(gsrc + i) : 0ull;
                }
#pragma unroll
                for (int k = 0; k < kWalk; ++k)
                    if (i0 + (uint32_t)k * kLargeThreads < n_walk) use(w[k]);
            }
        };
        // depth interval of the list
        uint32_t dmin = 0xffffffffu, dmax = 0u;
        if (staged) {
            for (uint32_t i0 = t; i0 < n; i0 += kWalk * kLargeThreads) {
                uint64_t w[kWalk];
#pragma unroll
                for (int k = 0; k < kWalk; ++k) {
                    const uint32_t i = i0 + (uint32_t)k * kLargeThreads;
                    w[k] = i < n ? __ldg(gsrc + i) : 0ull;
                }
#pragma unroll
                for (int k = 0; k < kWalk; ++k) {
                    const uint32_t i = i0 + (uint32_t)k * kLargeThreads;
                    if (i < n) {
                        s_in[i] = w[k];
                        const uint32_t d = (uint32_t)(w[k] >> 32);
                        dmin = min(dmin, d); dmax = max(dmax, d);
                    }
                }
            }
        } else {
            walk([&](uint64_t w) { const uint32_t d = (uint32_t)(w >> 32); dmin = min(dmin, d); dmax = max(dmax, d); });
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            dmin = min(dmin, __shfl_xor_sync(0xffffffffu, dmin, o));
            dmax = max(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
        }
        if (lane == 0) { s_red[0][warp] = dmin; s_red[1][warp] = dmax; }
        if (t < kCoarse) s_c[t] = 0;
        __syncthreads();
        for (int k = 0; k < kLargeThreads / 32; ++k) { dmin = min(dmin, s_red[0][k]); dmax = max(dmax, s_red[1][k]); }
        const float scale_c = (float)kCoarse / ((float)(dmax - dmin) + 1.0f);
        auto coarse_of = [&](uint64_t w) { return min((uint32_t)(kCoarse - 1), (uint32_t)((float)((uint32_t)(w >> 32) - dmin) * scale_c)); };
        walk([&](uint64_t w) { atomicAdd(&s_c[coarse_of(w)], 1u); });
        __syncthreads();
        coarse_scan(s_c, s_cstart);
        __syncthreads();
        // A list longer than the shared-memory buffer is sorted in depth chunks: runs of coarse bins [c0, c1) holding at
        // most kTileSortLargeMax words each, one after the other (a coarse bin is monotone in depth and equal depths share
        // a bin, so the chunks' sorted runs concatenate to the sorted list).  Every chunk re-reads the list from L2.
        uint32_t c0 = 0;
        while (c0 < (uint32_t)kCoarse) {
            uint32_t c1 = kCoarse;                      // a list that fits the buffer is one chunk: no walk over the bins
            if (n > kTileSortLargeMax) {
                c1 = c0 + 1;
                while (c1 < (uint32_t)kCoarse && s_cstart[c1 + 1] - s_cstart[c0] <= kTileSortLargeMax) ++c1;
            }
            const uint32_t first = s_cstart[c0], m_chunk = s_cstart[c1] - first;
            const bool whole = (c0 == 0 && c1 == (uint32_t)kCoarse);
            if (m_chunk == 0) { c0 = c1; continue; }
            if (m_chunk > kTileSortLargeMax) {
                // one coarse bin (1/128 of the list's depth interval) alone overflows the buffer: depths this clustered are
                // not sorted here.  Status flag 8 tells the caller to render the frame with the radix path (mode 0).
                // The chunk is still copied out (unsorted) so that the kernels queued behind this one read valid indices.
                if (t == 0) { atomicOr(status_flags, 8u); s_nbig = 0; }
                __syncthreads();
                walk([&](uint64_t w) {
                    const uint32_t kc = coarse_of(w);
                    if (kc >= c0 && kc < c1) {
                        const uint32_t pos = atomicAdd(&s_nbig, 1u);
                        vals[range.x + first + pos] = (uint32_t)w;
                        keys[range.x + first + pos] = ((uint64_t)tile << 32) | (w >> 32);
                    }
                });
                __syncthreads();
                c0 = c1;
                continue;
            }
            for (uint32_t i = t; i < kLargeBuckets; i += kLargeThreads) s_hist[i] = 0;
            if (t == 0) s_nbig = 0;
            __syncthreads();
            // buckets even out the chunk's load through the same piecewise-linear CDF estimate as BucketMap
            const float ratio = (float)kLargeBuckets / (float)m_chunk;
            auto bucket = [&](uint64_t w, uint32_t& kc) {      // one evaluation of the depth's position serves bin and bucket
                const float x = (float)((uint32_t)(w >> 32) - dmin) * scale_c;
                kc = min((uint32_t)(kCoarse - 1), (uint32_t)x);
                const float below = (float)(s_cstart[kc] - first), in_bin = (float)(s_cstart[kc + 1] - s_cstart[kc]);
                return min(kLargeBuckets - 1u, (uint32_t)((below + (x - (float)kc) * in_bin) * ratio));
            };
            walk([&](uint64_t w) {
                uint32_t kc;
                const uint32_t bk = bucket(w, kc);
                if (whole || (kc >= c0 && kc < c1)) atomicAdd(&s_hist[bk], 1u);
            });
            __syncthreads();
            // exclusive scan; thread t owns buckets [t kPer, (t + 1) kPer)
            uint32_t cnt[kPer], sum = 0;
#pragma unroll
            for (int q = 0; q < kPer; ++q) { cnt[q] = s_hist[t * kPer + q]; sum += cnt[q]; }
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) s_part[warp] = incl;
            __syncthreads();
            uint32_t start = incl - sum;
            for (int k = 0; k < warp; ++k) start += s_part[k];
            {
                uint32_t run = start;
#pragma unroll
                for (int q = 0; q < kPer; ++q) { s_hist[t * kPer + q] = run; run += cnt[q]; }
            }
            __syncthreads();
            walk([&](uint64_t w) {
                uint32_t kc;
                const uint32_t bk = bucket(w, kc);
                if (whole || (kc >= c0 && kc < c1)) s_out[atomicAdd(&s_hist[bk], 1u)] = w;
            });
            __syncthreads();
            {
                uint32_t b0 = start;
#pragma unroll
                for (int q = 0; q < kPer; ++q) {
                    const uint32_t m = cnt[q];
                    if (m > kLargeInsertMax) {
                        const uint32_t slot = atomicAdd(&s_nbig, 1u);
                        if (slot < 32) { s_big[2 * slot] = b0; s_big[2 * slot + 1] = m; }
                    } else {
                        for (uint32_t i = 1; i < m; ++i) {
                            const uint64_t x = s_out[b0 + i];
                            uint32_t j = i;
                            while (j > 0 && s_out[b0 + j - 1] > x) { s_out[b0 + j] = s_out[b0 + j - 1]; --j; }
                            s_out[b0 + j] = x;
                        }
                    }
                    b0 += m;
                }
            }
            __syncthreads();
            const uint32_t nbig = s_nbig;
            if (nbig > 32) {
                cta_sort_range(s_out, m_chunk);               // hopelessly clustered: sort the whole chunk
            } else {
                for (uint32_t k = 0; k < nbig; ++k) cta_sort_range(s_out + s_big[2 * k], s_big[2 * k + 1]);
            }
            for (uint32_t i = t; i < m_chunk; i += kLargeThreads) {
                const uint64_t x = s_out[i];
                vals[range.x + first + i] = (uint32_t)x;
                keys[range.x + first + i] = ((uint64_t)tile << 32) | (x >> 32);
            }
            __syncthreads();
            c0 = c1;
        }
        if (t == 0 && lazy != 0) tile_list[(size_t)tile * kTileCountStride + 7] = n;   // sorted completely now
        __syncthreads();
    }
}

}  // namespace

int launch_tile_scan(cudaStream_t s, const GeomState& g, const ImgState& im, int grid_x, int grid_y, uint64_t capacity, int P, bool strict)
{
    const int tiles = grid_x * grid_y;
    const size_t smem = tiles <= kTileDiffMaxTiles ? sizeof(uint32_t) * (size_t)tiles : 0;   // larger grids: preprocess does not use the array
    if (smem > 48 * 1024) GGD_CUDA_CHECK(cudaFuncSetAttribute(tile_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tile_scan_kernel<<<1, kScanThreadsT, smem, s>>>(g.tile_counts, im.ranges, grid_x, grid_y, (uint32_t)capacity, g.point_offsets, P,
                                                    g.counters, strict ? 1 : 0);
    GGD_LAUNCH_CHECK("tile_scan_kernel");
    return GGD_OK;
}

int launch_tile_scatter(cudaStream_t s, int P, int grid_x, int grid_y, const GeomState& g, const int* radii,
                        const BinState& b, const ImgState& im, uint64_t capacity, bool ref_order_arrays)
{
    (void)capacity; (void)im;
    const int blocks = (P + kScatterThreads - 1) / kScatterThreads;
    const int tiles = grid_x * grid_y;
    const int hist_tiles = tiles <= 4096 ? tiles : 0;
    // images beyond 1024^2 (more than 4096 tiles): two-level scatter over super-tiles (see super_emit_kernel)
    const int ns = ((grid_x + kSuper - 1) / kSuper) * ((grid_y + kSuper - 1) / kSuper);
    const char* force = getenv("GGD_SCATTER_SUPER");      // "1" / "0": force the choice (tests cover small images with it)
    const bool super = ns <= kSuperMax && (force ? force[0] == '1' : tiles > 4096);
    if (super) {
        super_bin_kernel<true><<<blocks, kSuperThreads, 0, s>>>(P, grid_x, grid_y, g.means2D, radii, g.tile_counts, b.vals_tmp, g.counters);
        GGD_LAUNCH_CHECK("super_bin_kernel<count>");
        super_scan_kernel<<<1, kSuperMax, 0, s>>>(ns, g.tile_counts, g.counters);
        GGD_LAUNCH_CHECK("super_scan_kernel");
        super_bin_kernel<false><<<blocks, kSuperThreads, 0, s>>>(P, grid_x, grid_y, g.means2D, radii, g.tile_counts, b.vals_tmp, g.counters);
        GGD_LAUNCH_CHECK("super_bin_kernel<write>");
        auto emit = ref_order_arrays ? super_emit_kernel<true> : super_emit_kernel<false>;
        emit<<<device_sm_count() * 8, kSuperThreads, 0, s>>>(grid_x, grid_y, g.means2D, g.depths, g.point_offsets, radii, b.keys_unsorted,
                                                             b.vals_unsorted, b.keys_tmp, g.tile_counts, b.vals_tmp, g.counters);
        GGD_LAUNCH_CHECK("super_emit_kernel");
        return GGD_OK;
    }
    auto kern = ref_order_arrays ? tile_scatter_kernel<true> : tile_scatter_kernel<false>;
    kern<<<blocks, kScatterThreads, sizeof(uint32_t) * 2 * (size_t)hist_tiles, s>>>(
        P, grid_x, grid_y, g.means2D, g.depths, g.point_offsets, radii, b.keys_unsorted, b.vals_unsorted, b.keys_tmp,
        g.tile_counts, g.counters, hist_tiles);
    GGD_LAUNCH_CHECK("tile_scatter_kernel");
    return GGD_OK;
}

int launch_tile_sort(cudaStream_t s, const GeomState& g, const BinState& b, const ImgState& im, int tiles, int lazy)
{
    if (lazy == 2) {      // second round: the marked long lists only
        int sms2 = device_sm_count();
        constexpr int smem2 = (int)(sizeof(uint64_t) * kTileSortLargeMax + sizeof(uint32_t) * kLargeBuckets);
        tile_sort_large_kernel<<<std::min(tiles, sms2), kLargeThreads, smem2, s>>>(im.ranges, b.keys_tmp, b.keys, b.vals, g.counters,
                                                                                  g.tile_counts, g.counters + 1, 2);
        GGD_LAUNCH_CHECK("tile_sort_large_kernel<round 2>");
        return GGD_OK;
    }
    constexpr int smem = (int)((sizeof(uint64_t) + sizeof(uint32_t)) * kTileSortMax);
    GGD_CUDA_CHECK(cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tile_sort_kernel<<<tiles, kTileSortThreads, smem, s>>>(im.ranges, b.keys_tmp, b.keys, b.vals, g.counters);
    GGD_LAUNCH_CHECK("tile_sort_kernel");
    constexpr int smem_large = (int)(sizeof(uint64_t) * kTileSortLargeMax + sizeof(uint32_t) * kLargeBuckets);
    GGD_CUDA_CHECK(cudaFuncSetAttribute(tile_sort_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_large));
    int sms = 148;
    { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    // (the flag word is counters[1]; passed as its own pointer because `counters` is read through the read-only path)
    if (lazy == 1) {      // first round of a lazily sorted frame: the nearest depth chunk of every long list
        GGD_CUDA_CHECK(cudaFuncSetAttribute(tile_sort_lazy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        tile_sort_lazy_kernel<<<std::min(tiles, 4 * sms), kTileSortThreads, smem, s>>>(im.ranges, b.keys_tmp, b.keys, b.vals, g.counters,
                                                                                      g.tile_counts);
        GGD_LAUNCH_CHECK("tile_sort_lazy_kernel");
        return GGD_OK;
    }
    tile_sort_large_kernel<<<std::min(tiles, sms), kLargeThreads, smem_large, s>>>(im.ranges, b.keys_tmp, b.keys, b.vals, g.counters,
                                                                                  g.tile_counts, g.counters + 1, 0);
    GGD_LAUNCH_CHECK("tile_sort_large_kernel");
    return GGD_OK;
}

}  // namespace ggd
