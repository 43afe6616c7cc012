// Point-set producer: density grid -> iso-surface -> random points on the surface.
//
// Replaces what main/decoder_utils/target_dataloader.py does between the density grid and
// `result.vertices` (:96-118, :185-190): skimage.measure.marching_cubes(sigmas, level) on the CPU, a
// trimesh round trip, one random barycentric point per face repeated until n points exist, truncation to
// n, and the radial scale clip(randn * thickness + 1, 0, 1).  Here the grid stays on the GPU:
//   mt_count_kernel   triangles per cell; the cell is split into the six tetrahedra around its main
//                     diagonal (no case tables: a tetrahedron has 0, 1 or 2 triangles), which tessellates
//                     the same piecewise-linear iso-surface family as marching cubes, watertight across cells
//   scan_kernel       (preprocess.cu) inclusive scan of the counts, total in counters[0]
//   mt_sample_kernel  point k belongs to triangle k mod n_triangles (the reference's repeat-and-truncate
//                     order); the thread finds the cell by binary search in the scanned counts, rebuilds
//                     that cell's triangles (cheaper than storing them: no triangle buffer, no capacity
//                     to guess) and draws the point with a counter-based generator
// Vertex coordinates follow the reference: index / res - 0.5.  The triangulation differs from skimage's
// (Lewiner) and the random stream from torch's, so parity with the reference is statistical, not bitwise:
// the checker is oracle/surface_oracle.py, a numpy restatement of exactly this algorithm (same cells,
// same generator), plus geometric properties (points lie on the iso-surface).
#include "common.cuh"

namespace ggd {
namespace {

// the six tetrahedra of the Kuhn subdivision: corner index bit0 = +x, bit1 = +y, bit2 = +z
__device__ __constant__ int8_t kTet[6][4] = {{0, 1, 3, 7}, {0, 1, 5, 7}, {0, 2, 3, 7}, {0, 2, 6, 7}, {0, 4, 5, 7}, {0, 4, 6, 7}};

struct Cell {
    float v[8];
    int x, y, z;
};

__device__ __forceinline__ Cell load_cell(const float* __restrict__ sigma, int res, uint32_t cell)
{
    const int n = res - 1;
    Cell c;
    c.x = (int)(cell % n); c.y = (int)((cell / n) % n); c.z = (int)(cell / ((uint32_t)n * n));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        // sigma is indexed [i0][i1][i2] like the numpy array the reference hands to marching_cubes; vertex
        // coordinate j of a mesh vertex runs along axis j, so corner bit j advances index j
        const int i0 = c.x + (i & 1), i1 = c.y + ((i >> 1) & 1), i2 = c.z + ((i >> 2) & 1);
        c.v[i] = __ldg(sigma + ((size_t)i0 * res + i1) * res + i2);
    }
    return c;
}

__device__ __forceinline__ int tet_triangles(uint32_t mask)   // mask: which of the 4 vertices are above the level
{
    const int inside = __popc(mask);
    return (inside == 0 || inside == 4) ? 0 : (inside == 2 ? 2 : 1);
}

__global__ void __launch_bounds__(256)
mt_count_kernel(const float* __restrict__ sigma, int res, float level, uint32_t cells, uint32_t* __restrict__ counts)
{
    const uint32_t cell = blockIdx.x * 256u + threadIdx.x;
    if (cell >= cells) return;
    const Cell c = load_cell(sigma, res, cell);
    uint32_t above = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) above |= (c.v[i] > level ? 1u : 0u) << i;
    uint32_t n = 0;
    if (above != 0u && above != 0xffu) {
#pragma unroll
        for (int t = 0; t < 6; ++t) {
            uint32_t m = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) m |= ((above >> kTet[t][k]) & 1u) << k;
            n += tet_triangles(m);
        }
    }
    counts[cell] = n;
}

// ---- counter-based random numbers: splitmix64 of (seed, stream, index) ----
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ float uniform01(uint64_t seed, uint32_t k, uint32_t j)   // in (0, 1)
{
    const uint64_t r = mix64(seed ^ mix64(((uint64_t)k << 8) | j));
    return ((float)(uint32_t)(r >> 40) + 0.5f) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float3 edge_point(const Cell& c, int a, int b, float level)
{
    const float va = c.v[a], vb = c.v[b];
    const float t = (level - va) / (vb - va);
    const float ax = (float)(c.x + (a & 1)), ay = (float)(c.y + ((a >> 1) & 1)), az = (float)(c.z + ((a >> 2) & 1));
    const float bx = (float)(c.x + (b & 1)), by = (float)(c.y + ((b >> 1) & 1)), bz = (float)(c.z + ((b >> 2) & 1));
    return make_float3(ax + t * (bx - ax), ay + t * (by - ay), az + t * (bz - az));
}

__global__ void __launch_bounds__(256)
mt_sample_kernel(const float* __restrict__ sigma, int res, float level, uint32_t cells, const uint32_t* __restrict__ offsets,
                 const uint32_t* __restrict__ counters, uint64_t seed, int n_points, float thickness, float* __restrict__ points)
{
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= n_points) return;
    const uint32_t n_tri = __ldg(counters);
    if (n_tri == 0) { points[3 * k] = points[3 * k + 1] = points[3 * k + 2] = 0.f; return; }
    const uint32_t tri = (uint32_t)k % n_tri;
    // first cell whose inclusive offset exceeds tri
    uint32_t lo = 0, hi = cells - 1;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(offsets + mid) > tri) hi = mid; else lo = mid + 1;
    }
    const uint32_t cell = lo;
    uint32_t local = tri - (cell ? __ldg(offsets + cell - 1) : 0u);
    const Cell c = load_cell(sigma, res, cell);
    uint32_t above = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) above |= (c.v[i] > level ? 1u : 0u) << i;
    float3 p0 = make_float3(0.f, 0.f, 0.f), p1 = p0, p2 = p0;
    for (int t = 0; t < 6; ++t) {
        uint32_t m = 0;
        for (int q = 0; q < 4; ++q) m |= ((above >> kTet[t][q]) & 1u) << q;
        const uint32_t nt = (uint32_t)tet_triangles(m);
        if (local >= nt) { local -= nt; continue; }
        const int8_t* tv = kTet[t];
        const int inside = __popc(m);
        if (inside == 2) {
            // a, b on one side, cc, d on the other: quad (ac, ad, bd, bc) -> (ac, ad, bd), (ac, bd, bc)
            int in[2], out[2], ni = 0, no = 0;
            for (int q = 0; q < 4; ++q) { if ((m >> q) & 1u) in[ni++] = tv[q]; else out[no++] = tv[q]; }
            const float3 ac = edge_point(c, in[0], out[0], level), ad = edge_point(c, in[0], out[1], level);
            const float3 bd = edge_point(c, in[1], out[1], level), bc = edge_point(c, in[1], out[0], level);
            p0 = ac;
            if (local == 0) { p1 = ad; p2 = bd; } else { p1 = bd; p2 = bc; }
        } else {
            // the lone vertex (inside == 1: the one above; inside == 3: the one below) and its three edges
            const uint32_t lone_mask = inside == 1 ? m : (~m & 0xfu);
            const int lq = __ffs(lone_mask) - 1;
            int others[3], no = 0;
            for (int q = 0; q < 4; ++q) if (q != lq) others[no++] = tv[q];
            p0 = edge_point(c, tv[lq], others[0], level);
            p1 = edge_point(c, tv[lq], others[1], level);
            p2 = edge_point(c, tv[lq], others[2], level);
        }
        break;
    }
    // the reference's weights: three uniforms divided by their sum (target_dataloader.py:104-106)
    const float u0 = uniform01(seed, (uint32_t)k, 0), u1 = uniform01(seed, (uint32_t)k, 1), u2 = uniform01(seed, (uint32_t)k, 2);
    const float inv = 1.0f / (u0 + u1 + u2);
    const float w0 = u0 * inv, w1 = u1 * inv, w2 = u2 * inv;
    float px = p0.x * w0 + p1.x * w1 + p2.x * w2, py = p0.y * w0 + p1.y * w1 + p2.y * w2, pz = p0.z * w0 + p1.z * w1 + p2.z * w2;
    // index coordinates -> the generator's cube (:98-99), then the radial jitter (:110-113)
    const float r = 1.0f / (float)res;
    px = px * r - 0.5f; py = py * r - 0.5f; pz = pz * r - 0.5f;
    const float g0 = uniform01(seed, (uint32_t)k, 3), g1 = uniform01(seed, (uint32_t)k, 4);
    const float normal = sqrtf(-2.0f * logf(g0)) * cosf(6.283185307179586f * g1);
    const float scale = fminf(fmaxf(normal * thickness + 1.0f, 0.0f), 1.0f);
    points[3 * k] = px * scale; points[3 * k + 1] = py * scale; points[3 * k + 2] = pz * scale;
}

struct SurfaceScratch {
    uint32_t* counts;
    uint32_t* offsets;
    uint32_t* counters;     // [64]: [0] total, [2] scan ticket
    uint64_t* scan_status;
    size_t bytes, zero_off, zero_bytes;
};
SurfaceScratch carve_surface(void* base, size_t cells)
{
    SurfaceScratch sc;
    char* p = static_cast<char*>(base);
    auto take = [&](size_t bytes) { char* r = p; p += align_up(bytes, 256); return r; };
    sc.counts = reinterpret_cast<uint32_t*>(take(cells * 4));
    sc.offsets = reinterpret_cast<uint32_t*>(take(cells * 4));
    char* z0 = p;
    sc.counters = reinterpret_cast<uint32_t*>(take(64 * 4));
    sc.scan_status = reinterpret_cast<uint64_t*>(take(((cells + kScanTile - 1) / kScanTile + 1) * 8));
    sc.zero_off = (size_t)(z0 - static_cast<char*>(base));
    sc.zero_bytes = (size_t)(p - z0);
    sc.bytes = (size_t)(p - static_cast<char*>(base));
    return sc;
}

}  // namespace
}  // namespace ggd

using namespace ggd;

extern "C" {

size_t ggd_surface_points_scratch_bytes(int res)
{
    if (res < 2) return 0;
    const size_t cells = (size_t)(res - 1) * (res - 1) * (res - 1);
    return carve_surface(nullptr, cells).bytes;
}

int ggd_surface_points(void* stream, const float* sigma, int res, float level, uint64_t seed, int n_points, float thickness,
                       void* scratch, size_t scratch_bytes, float* points, uint32_t* num_triangles_host)
{
    if (!sigma || !scratch || (n_points > 0 && !points)) { set_error("ggd_surface_points: null required pointer"); return GGD_E_INVALID; }
    if (res < 2 || res > 1024 || n_points < 0) { set_error("ggd_surface_points: bad size (res %d, n_points %d)", res, n_points); return GGD_E_INVALID; }
    const size_t cells = (size_t)(res - 1) * (res - 1) * (res - 1);
    if (cells >= (1ull << 31)) { set_error("ggd_surface_points: grid too large"); return GGD_E_INVALID; }
    SurfaceScratch sc = carve_surface(scratch, cells);
    if (scratch_bytes < sc.bytes) { set_error("ggd_surface_points: scratch too small (%zu < %zu)", scratch_bytes, sc.bytes); return GGD_E_INVALID; }
    if (reinterpret_cast<uintptr_t>(scratch) & 255) { set_error("ggd_surface_points: scratch must be 256-byte aligned"); return GGD_E_INVALID; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    GGD_CUDA_CHECK(cudaMemsetAsync(static_cast<char*>(scratch) + sc.zero_off, 0, sc.zero_bytes, s));
    mt_count_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, s>>>(sigma, res, level, (uint32_t)cells, sc.counts);
    GGD_LAUNCH_CHECK("mt_count_kernel");
    GeomState g{};   // the scan only touches these four fields
    g.tiles_touched = sc.counts; g.point_offsets = sc.offsets; g.counters = sc.counters; g.scan_status = sc.scan_status;
    if (int e = launch_scan(s, (int)cells, g)) return e;
    if (n_points > 0) {
        mt_sample_kernel<<<(n_points + 255) / 256, 256, 0, s>>>(sigma, res, level, (uint32_t)cells, sc.offsets, sc.counters, seed,
                                                                n_points, thickness, points);
        GGD_LAUNCH_CHECK("mt_sample_kernel");
    }
    if (num_triangles_host)
        GGD_CUDA_CHECK(cudaMemcpyAsync(num_triangles_host, sc.counters, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    return GGD_OK;
}

}  // extern "C"
