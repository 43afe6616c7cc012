// Stage 2a of the rasteriser: key duplication, stable LSD radix sort
// ("onesweep": one read + one write of the pairs per digit, chained-scan
// decoupled look-back across tiles), and tile range detection.
//
// Replaces duplicateWithKeys (DGR/cuda_rasterizer/rasterizer_impl.cu:70-111),
// the call into cub::DeviceRadixSort::SortPairs (:303-308; CUB is a CUDA
// toolkit dependency, not vendored — the contract is "stable sort on key bits
// [0, 32+bit)", which is what this file implements), and identifyTileRanges
// (:116-138).  Key layout is the reference's: (tile_id << 32) | float_bits(depth).
//
// What differs from the reference's structure:
//  * duplication is slot-centric: a CTA owns a contiguous range of output
//    slots (its 256 Gaussians' prefix-sum interval) and each thread fills slots
//    s, s+256, ... after a binary search for the owning Gaussian.  Writes are
//    fully coalesced and a huge splat is spread over the whole CTA instead of
//    being looped over by one thread.
//  * the digit histograms of ALL sort passes are accumulated while the keys
//    are produced (shared-memory atomics), so the sort never re-reads the keys
//    for a separate histogram kernel.
//  * the element count lives in device memory; grids are sized by the buffer
//    capacity and surplus CTAs exit, so no host round-trip is needed.
#include "common.cuh"
#include <atomic>
#include <cstdlib>

namespace ggd {

namespace {

constexpr int kDupThreads = 256;
static_assert(kSortThreads == kRadixBins, "one thread per digit in the look-back");

// ---------------------------------------------------------------------------
// duplicate
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kDupThreads)
duplicate_kernel(int P, int grid_x, int grid_y, const float2* __restrict__ means2D,
                 const float* __restrict__ depths, const uint32_t* __restrict__ offsets,
                 const int* __restrict__ radii, uint64_t* __restrict__ keys,
                 uint32_t* __restrict__ vals, uint32_t capacity, uint32_t* __restrict__ counters,
                 uint32_t* __restrict__ hist, int num_passes, const uint32_t* __restrict__ skip)
{
    if (skip && __ldg(skip)) return;   // frame binned per tile (tile_binning.cu)
    __shared__ uint32_t s_end[kDupThreads];
    __shared__ uint32_t s_depth[kDupThreads];
    __shared__ uint32_t s_rect[kDupThreads];  // x0 | y0 << 16 (tile coordinates, each < 65536)
    __shared__ uint32_t s_w[kDupThreads];
    __shared__ uint32_t s_hist[kMaxSortPasses][kRadixBins];

    const int t = threadIdx.x;
    const int base = blockIdx.x * kDupThreads;
    const int idx = base + t;
    for (int i = t; i < num_passes * kRadixBins; i += kDupThreads) (&s_hist[0][0])[i] = 0;

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
    const int last = min(kDupThreads, P - base) - 1;
    const uint32_t block_end = s_end[last];

    if (blockIdx.x == 0 && t == 0) {
        const uint32_t R = __ldg(offsets + P - 1);
        counters[3] = min(R, capacity);  // element count the sort and the range kernel use
        counters[1] = (R > capacity) ? 1u : 0u;
    }

    for (uint32_t s = block_start + t; s < block_end; s += kDupThreads) {
        // first g with s_end[g] > s
        int lo = 0, hi = last;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int mid = (lo + hi) >> 1;
            if (s_end[mid] > s) hi = mid; else lo = mid + 1;
        }
        const int g = lo;
        const uint32_t g_start = (g == 0) ? block_start : s_end[g - 1];
        const uint32_t j = s - g_start;
        const uint32_t w = s_w[g];
        const uint32_t rc = s_rect[g];
        const uint32_t y = (rc >> 16) + j / w;
        const uint32_t x = (rc & 0xffffu) + j % w;
        uint64_t key = (uint64_t)(y * (uint32_t)grid_x + x);
        key <<= 32;
        key |= s_depth[g];
        if (s < capacity) {
            keys[s] = key;
            vals[s] = (uint32_t)(base + g);
#pragma unroll
            for (int p = 0; p < kMaxSortPasses; ++p)
                if (p < num_passes) atomicAdd(&s_hist[p][(key >> (p * kRadixBits)) & (kRadixBins - 1)], 1u);
        }
    }
    __syncthreads();
    for (int i = t; i < num_passes * kRadixBins; i += kDupThreads) {
        const uint32_t c = (&s_hist[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
}

// ---------------------------------------------------------------------------
// stand-alone histogram (ggd_sort_pairs entry point, keys not produced by us)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
histogram_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ n_ptr,
                 uint32_t capacity, int begin_bit, int end_bit, int num_passes,
                 uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_hist[kMaxSortPasses][kRadixBins];
    const uint32_t n = min(__ldg(n_ptr), capacity);
    for (int i = threadIdx.x; i < num_passes * kRadixBins; i += 256) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < n; i += gridDim.x * 256u) {
        const uint64_t key = __ldg(keys + i);
        for (int p = 0; p < num_passes; ++p) {
            const int shift = begin_bit + p * kRadixBits;
            const int bits = min(kRadixBits, end_bit - shift);
            atomicAdd(&s_hist[p][(key >> shift) & ((1u << bits) - 1u)], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < num_passes * kRadixBins; i += 256) {
        const uint32_t c = (&s_hist[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
}

// ---------------------------------------------------------------------------
// onesweep digit pass
// ---------------------------------------------------------------------------
constexpr uint32_t kLbAgg = 1u << 30;
constexpr uint32_t kLbPre = 2u << 30;
constexpr uint32_t kLbMask = (1u << 30) - 1u;
constexpr int kSortWarps = kSortThreads / 32;

// exclusive scan of one value per thread over the 256-thread CTA
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp /*[8]*/)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    __syncthreads();  // protect s_warp reuse
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w)
        if (w < warp) off += s_warp[w];
    return off + incl - v;
}

// One CTA: bit q of *plan = pass q's digit is the same for all n keys.
__global__ void __launch_bounds__(kRadixBins)
sort_plan_kernel(const uint32_t* __restrict__ hist, const uint32_t* __restrict__ n_ptr, uint32_t capacity, int num_passes,
                 uint32_t* __restrict__ plan)
{
    const uint32_t n = min(__ldg(n_ptr), capacity);
    uint32_t m = 0;
    for (int q = 0; q < num_passes; ++q)
        if (__syncthreads_or(n != 0 && hist[q * kRadixBins + threadIdx.x] == n)) m |= 1u << q;
    if (threadIdx.x == 0) *plan = m;
}

// kMinCtas: resident CTAs per SM the register allocation aims at.  2 -> 128 registers, no spills: fastest when a pass
// is a few hundred CTAs and latency-bound (the headline frame); 4 -> 64 registers with some spilled words, but
// twice the warps in flight: fastest when a pass streams tens of MB or more (trained-like and 2048^2 frames).
template <int kMinCtas>
__global__ void __launch_bounds__(kSortThreads, kMinCtas)
onesweep_kernel(const uint64_t* __restrict__ keys_src, uint64_t* __restrict__ keys_a, uint64_t* __restrict__ keys_b,
                const uint32_t* __restrict__ vals_src, uint32_t* __restrict__ vals_a, uint32_t* __restrict__ vals_b,
                const uint32_t* __restrict__ n_ptr, uint32_t capacity, int shift, uint32_t mask,
                const uint32_t* __restrict__ hist /*[256] digit counts of this pass*/,
                uint32_t* __restrict__ lookback /*[tiles][256]*/, uint32_t* __restrict__ ticket,
                const uint32_t* __restrict__ plan /*bit q: pass q is the identity*/, int pass, int num_passes)
{
    __shared__ __align__(16) uint64_t s_keys[kSortTile];       // 32 KB, reused for values
    __shared__ uint32_t s_whist[kSortWarps][kRadixBins];        // 8 KB
    __shared__ uint32_t s_gbase[kRadixBins];
    __shared__ uint32_t s_loff[kRadixBins];
    __shared__ uint32_t s_scan[kSortWarps];
    __shared__ uint32_t s_tile;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) s_whist[w][t] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t n = min(__ldg(n_ptr), capacity);
    const uint64_t tile_base = (uint64_t)tile * kSortTile;
    if (tile_base >= n) return;
    const uint32_t n_valid = (uint32_t)min((uint64_t)kSortTile, (uint64_t)n - tile_base);

    // ---- which buffers this pass works on.  A pass whose digit is the same for every key (the exponent byte of
    // depths that all lie in [2,4), tile-id bits above the tile count) is the identity permutation and is skipped
    // outright; sort_plan_kernel found those from the histograms.  The remaining passes ping-pong between the two
    // buffer sets such that the LAST of them lands in set a (the caller's output): counted from the back.
    const uint32_t ident = __ldg(plan);
    const int real_total = num_passes - __popc(ident & ((1u << num_passes) - 1u));
    const uint64_t* keys_in; const uint32_t* vals_in;
    uint64_t* keys_out; uint32_t* vals_out;
    if ((ident >> pass) & 1u) {
        if (real_total != 0 || pass != num_passes - 1) return;
        // every pass is the identity (one key, or all keys equal in the sorted bits): the output is the input
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t li = t + k * kSortThreads;
            if (li < n_valid) {
                keys_a[tile_base + li] = __ldg(keys_src + tile_base + li);
                vals_a[tile_base + li] = __ldg(vals_src + tile_base + li);
            }
        }
        return;
    }
    {
        const int j = __popc(~ident & ((1u << pass) - 1u));          // real passes before this one
        const bool to_a = ((real_total - 1 - j) & 1) == 0;
        keys_out = to_a ? keys_a : keys_b; vals_out = to_a ? vals_a : vals_b;
        keys_in = j == 0 ? keys_src : (to_a ? keys_b : keys_a);
        vals_in = j == 0 ? vals_src : (to_a ? vals_b : vals_a);
    }

    // ---- load (warp-striped: item k of lane l sits at warp_base + k*32 + l)
    uint64_t key[kSortItems];
    const uint32_t warp_base = warp * (32 * kSortItems);
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t li = warp_base + k * 32 + lane;
        key[k] = (li < n_valid) ? __ldg(keys_in + tile_base + li) : ~0ull;
    }

    // ---- stable in-warp ranking, digit by digit position (k outer)
    uint16_t rank[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t li = warp_base + k * 32 + lane;
        const uint32_t d = (li < n_valid) ? (uint32_t)((key[k] >> shift) & mask) : (kRadixBins - 1);
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t prior = s_whist[warp][d];
        __syncwarp();
        if ((peers & ((1u << lane) - 1u)) == 0) s_whist[warp][d] = prior + __popc(peers);
        __syncwarp();
        rank[k] = (uint16_t)(prior + __popc(peers & ((1u << lane) - 1u)));
    }
    __syncthreads();

    // ---- per digit (thread t == digit t): warp counts -> warp-exclusive offsets, CTA total
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
        const uint32_t c = s_whist[w][t];
        s_whist[w][t] = total;
        total += c;
    }
    uint32_t count = total;
    if (t == kRadixBins - 1) count -= (kSortTile - n_valid);  // padding items are not real

    // ---- publish aggregate, look back for the exclusive prefix of this digit
    uint32_t* my_lb = lookback + (size_t)tile * kRadixBins + t;
    uint32_t excl = 0;
    if (tile == 0) {
        atomicExch(my_lb, kLbPre | count);
    } else {
        atomicExch(my_lb, kLbAgg | count);
        // windowed look-back: kWin predecessor words are fetched as independent loads per round, so the
        // walk costs one L2 round trip per kWin tiles instead of one per tile
        constexpr int kWin = 8;
        int look = (int)tile - 1;
        bool found = false;
        while (!found) {
            uint32_t w[kWin];
#pragma unroll
            for (int i = 0; i < kWin; ++i)
                w[i] = (look - i >= 0) ? *reinterpret_cast<volatile uint32_t*>(lookback + (size_t)(look - i) * kRadixBins + t)
                                       : (2u << 30);      // virtual inclusive prefix 0 in front of tile 0
#pragma unroll
            for (int i = 0; i < kWin; ++i) {
                if (found) break;
                const uint32_t f = w[i] >> 30;
                if (f == 0) break;                        // not published yet: re-read from here
                excl += w[i] & kLbMask;
                --look;
                if (f == 2) found = true;
            }
        }
        atomicExch(my_lb, kLbPre | (excl + count));
    }

    // ---- digit bases: global (from the whole-array histogram) and CTA-local
    const uint32_t g_excl = block_excl_scan(__ldg(hist + t), s_scan);
    const uint32_t l_excl = block_excl_scan(total, s_scan);
    s_loff[t] = l_excl;
    s_gbase[t] = g_excl + excl - l_excl;
    __syncthreads();

    // ---- local reorder through shared memory, then coalesced scatter
    uint16_t lpos[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t li = warp_base + k * 32 + lane;
        const uint32_t d = (li < n_valid) ? (uint32_t)((key[k] >> shift) & mask) : (kRadixBins - 1);
        lpos[k] = (uint16_t)(s_loff[d] + s_whist[warp][d] + rank[k]);
        s_keys[lpos[k]] = key[k];
    }
    __syncthreads();
    uint32_t gpos[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t i = t + k * kSortThreads;
        if (i < n_valid) {
            const uint64_t kk = s_keys[i];
            const uint32_t d = (uint32_t)((kk >> shift) & mask);
            gpos[k] = s_gbase[d] + i;
            keys_out[gpos[k]] = kk;
        }
    }
    __syncthreads();
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t li = warp_base + k * 32 + lane;
        if (li < n_valid) s_vals[lpos[k]] = __ldg(vals_in + tile_base + li);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t i = t + k * kSortThreads;
        if (i < n_valid) vals_out[gpos[k]] = s_vals[i];
    }
}

// ---------------------------------------------------------------------------
// tile ranges
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t* __restrict__ n_ptr, const uint64_t* __restrict__ keys,
                   uint2* __restrict__ ranges)
{
    const uint32_t L = __ldg(n_ptr);
    for (uint32_t idx = blockIdx.x * 256u + threadIdx.x; idx < L; idx += gridDim.x * 256u) {
        const uint32_t cur = (uint32_t)(__ldg(keys + idx) >> 32);
        if (idx == 0) {
            ranges[cur].x = 0;
        } else {
            const uint32_t prev = (uint32_t)(__ldg(keys + idx - 1) >> 32);
            if (cur != prev) {
                ranges[prev].y = idx;
                ranges[cur].x = idx;
            }
        }
        if (idx == L - 1) ranges[cur].y = L;
    }
}

}  // namespace

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
SortScratch carve_sort_scratch(void* base, uint64_t capacity)
{
    SortScratch sc;
    const size_t tiles = (size_t)((capacity + kSortTile - 1) / kSortTile) + 1;
    char* p = static_cast<char*>(base);
    size_t off = 0;
    sc.hist = reinterpret_cast<uint32_t*>(p + off);
    off += align_up(sizeof(uint32_t) * kMaxSortPasses * kRadixBins, 256);
    sc.tickets = reinterpret_cast<uint32_t*>(p + off);
    off += 256;
    sc.lookback = reinterpret_cast<uint32_t*>(p + off);
    off += align_up(sizeof(uint32_t) * kMaxSortPasses * tiles * kRadixBins, 256);
    sc.bytes = off;
    sc.zero_bytes = off;
    return sc;
}

int sort_pairs(cudaStream_t s, const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in,
               uint32_t* vals_out, const uint32_t* n_dev, uint64_t capacity, int begin_bit,
               int end_bit, uint64_t* keys_tmp, uint32_t* vals_tmp, const SortScratch& sc,
               bool hist_precomputed)
{
    if (capacity >= (1ull << 30)) {
        set_error("sort_pairs: capacity %llu exceeds 2^30-1", (unsigned long long)capacity);
        return GGD_E_INVALID;
    }
    if (begin_bit < 0 || end_bit > 64 || end_bit <= begin_bit) {
        set_error("sort_pairs: bad bit range [%d,%d)", begin_bit, end_bit);
        return GGD_E_INVALID;
    }
    const int num_passes = (end_bit - begin_bit + kRadixBits - 1) / kRadixBits;
    const uint32_t cap = (uint32_t)capacity;
    const uint32_t tiles = (uint32_t)((capacity + kSortTile - 1) / kSortTile);
    if (tiles == 0) return GGD_OK;
    if (!hist_precomputed) {
        const int blocks = (int)std::min<uint64_t>((capacity + 255) / 256, 148 * 8);
        histogram_kernel<<<blocks, 256, 0, s>>>(keys_in, n_dev, cap, begin_bit, end_bit, num_passes, sc.hist);
        GGD_LAUNCH_CHECK("histogram_kernel");
    }
    const size_t lb_stride = (size_t)(tiles + 1) * kRadixBins;
    uint32_t* plan = sc.tickets + 16;
    sort_plan_kernel<<<1, kRadixBins, 0, s>>>(sc.hist, n_dev, cap, num_passes, plan);
    GGD_LAUNCH_CHECK("sort_plan_kernel");
    for (int p = 0; p < num_passes; ++p) {
        const int shift = begin_bit + p * kRadixBits;
        const int bits = std::min(kRadixBits, end_bit - shift);
        const uint32_t mask = (1u << bits) - 1u;
        // measured on B200: R = 0.8 M -> 0.117 ms (2 CTAs/SM) vs 0.148 (4); R = 6.5 M -> 0.58 vs 0.44; R = 280 M -> 21 vs 13.5 ms
        if (capacity >= (4ull << 20))
            onesweep_kernel<4><<<tiles, kSortThreads, 0, s>>>(keys_in, keys_out, keys_tmp, vals_in, vals_out, vals_tmp, n_dev, cap, shift,
                                                              mask, sc.hist + p * kRadixBins, sc.lookback + p * lb_stride,
                                                              sc.tickets + p, plan, p, num_passes);
        else
            onesweep_kernel<2><<<tiles, kSortThreads, 0, s>>>(keys_in, keys_out, keys_tmp, vals_in, vals_out, vals_tmp, n_dev, cap, shift,
                                                              mask, sc.hist + p * kRadixBins, sc.lookback + p * lb_stride,
                                                              sc.tickets + p, plan, p, num_passes);
        GGD_LAUNCH_CHECK("onesweep_kernel");
    }
    return GGD_OK;
}

int launch_duplicate(cudaStream_t s, int P, int grid_x, int grid_y, const GeomState& g,
                     const int* radii, const BinState& b, uint64_t capacity,
                     const SortScratch& sc, int num_passes, const uint32_t* skip)
{
    const int blocks = (P + kDupThreads - 1) / kDupThreads;
    duplicate_kernel<<<blocks, kDupThreads, 0, s>>>(P, grid_x, grid_y, g.means2D, g.depths,
                                                    g.point_offsets, radii, b.keys_unsorted,
                                                    b.vals_unsorted, (uint32_t)capacity, g.counters,
                                                    sc.hist, num_passes, skip);
    GGD_LAUNCH_CHECK("duplicate_kernel");
    return GGD_OK;
}

int launch_tile_ranges(cudaStream_t s, const GeomState& g, const BinState& b, uint64_t capacity,
                       const ImgState& im, int tiles, const uint32_t* skip)
{
    (void)tiles;
    const int blocks = (int)std::max<uint64_t>(1, std::min<uint64_t>((capacity + 255) / 256, 148 * 8));
    // with `skip` the element count comes from counters[6], which is 0 for a frame binned per tile
    tile_ranges_kernel<<<blocks, 256, 0, s>>>(skip ? g.counters + 6 : g.counters + 3, b.keys, im.ranges);
    GGD_LAUNCH_CHECK("tile_ranges_kernel");
    return GGD_OK;
}

}  // namespace ggd

// ---------------------------------------------------------------------------
// Sort entry point for callers that may not allocate device memory
// ---------------------------------------------------------------------------
// Some hosts of this library own all device memory themselves: a renderer or
// serving process that carves the GPU into fixed per-tenant budgets at start-up,
// or an engine that embeds this code as a plug-in and forbids it to touch the
// device allocator (and cudaMalloc would stall every stream of a running frame
// loop anyway).  Such callers can hand us keys, values and ping-pong buffers out
// of their own arena, but they have nowhere to get the sort's bookkeeping memory
// (digit histograms, tile tickets, look-back words) from.  For them the
// bookkeeping lives in the array below: it is part of the loaded module, so every
// GPU that loads the library holds these ~4 MB until the process exits.
namespace {
constexpr uint64_t kStaticSortMaxPairs = GGD_SORT_STATIC_MAX_PAIRS;
constexpr size_t kStaticSortTiles = (size_t)((kStaticSortMaxPairs + ggd::kSortTile - 1) / ggd::kSortTile) + 1;
constexpr size_t kStaticSortBytes =
    ((sizeof(uint32_t) * ggd::kMaxSortPasses * ggd::kRadixBins + 255) / 256) * 256 + 256 +
    ((sizeof(uint32_t) * ggd::kMaxSortPasses * kStaticSortTiles * ggd::kRadixBins + 255) / 256) * 256;
constexpr int kMaxDevices = 64;
}  // namespace

static __device__ __align__(256) unsigned char g_static_sort_scratch[kStaticSortBytes];

extern "C" int ggd_sort_pairs_static(void* stream, const uint64_t* keys_in, uint64_t* keys_out,
                                     const uint32_t* vals_in, uint32_t* vals_out,
                                     const uint32_t* n_dev, uint64_t capacity, int begin_bit, int end_bit,
                                     uint64_t* keys_tmp, uint32_t* vals_tmp)
{
    using namespace ggd;
    if (!keys_in || !keys_out || !vals_in || !vals_out || !n_dev || !keys_tmp || !vals_tmp) {
        set_error("ggd_sort_pairs_static: null pointer");
        return GGD_E_INVALID;
    }
    if (capacity > kStaticSortMaxPairs) {
        set_error("ggd_sort_pairs_static: capacity %llu exceeds the built-in scratch limit of %llu pairs; "
                  "use ggd_sort_pairs with caller-provided scratch",
                  (unsigned long long)capacity, (unsigned long long)kStaticSortMaxPairs);
        return GGD_E_CAPACITY;
    }
    // Set-up, once per device, with that device current: in host code the array's name is only a
    // stand-in for the symbol, not a device address, so the address each GPU gave it is looked up
    // and remembered here and THAT pointer is what the kernels receive.
    static std::atomic<void*> scratch_on_device[kMaxDevices];
    int dev = -1;
    GGD_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("ggd_sort_pairs_static: device ordinal %d out of range", dev);
        return GGD_E_INVALID;
    }
    // The address belongs to the CURRENT device; a stream of another device would hand its kernels a
    // pointer into the wrong GPU.  (The legacy and per-thread default streams always follow the current device.)
    if (stream != nullptr && stream != (void*)cudaStreamLegacy && stream != (void*)cudaStreamPerThread) {
        int sdev = -1;
        GGD_CUDA_CHECK(cudaStreamGetDevice(static_cast<cudaStream_t>(stream), &sdev));
        if (sdev != dev) {
            set_error("ggd_sort_pairs_static: stream belongs to device %d but device %d is current; "
                      "make the stream's device current before the call", sdev, dev);
            return GGD_E_INVALID;
        }
    }
    void* scratch = scratch_on_device[dev].load(std::memory_order_acquire);
    if (!scratch) {   // two threads may both get here: they look up and store the same address
        GGD_CUDA_CHECK(cudaGetSymbolAddress(&scratch, g_static_sort_scratch));
        scratch_on_device[dev].store(scratch, std::memory_order_release);
    }
    SortScratch sc = carve_sort_scratch(scratch, capacity);
    if (sc.bytes > kStaticSortBytes) {
        set_error("ggd_sort_pairs_static: internal scratch size mismatch (%zu > %zu)", sc.bytes, kStaticSortBytes);
        return GGD_E_CAPACITY;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // The array still holds the previous call's histograms, tickets and look-back flags:
    // everything a sort reads before writing must be zero again.
    GGD_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sc.zero_bytes, s));
    return sort_pairs(s, keys_in, keys_out, vals_in, vals_out, n_dev, capacity, begin_bit, end_bit,
                      keys_tmp, vals_tmp, sc, false);
}

