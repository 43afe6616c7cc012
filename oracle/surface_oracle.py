"""TEST INFRASTRUCTURE ONLY - numpy restatement of csrc/surface_points.cu (point-set producer).

Follows the same steps as the CUDA kernels: the (res-1)^3 cells of the density grid in x-fastest order, six
tetrahedra per cell around the main diagonal, 0/1/2 triangles per tetrahedron, point k on triangle
k mod n_triangles with splitmix64-derived uniforms, v / res - 0.5, radial scale clip(normal * thickness + 1, 0, 1).

What it stands in for in the reference: main/decoder_utils/target_dataloader.py:96-118, 185-190
(skimage.measure.marching_cubes + trimesh + torch.rand).  skimage and trimesh are not installed here and the
reference's random stream is torch's, so this oracle is NOT pinned against reference outputs: "parity unpinned"
for this row; tests check this restatement against the CUDA path exactly (triangle counts) / to float tolerance
(points), and check geometric properties of the result (points on the iso-surface) that hold for the reference
too.
"""
import numpy as np

TETS = np.array([[0, 1, 3, 7], [0, 1, 5, 7], [0, 2, 3, 7], [0, 2, 6, 7], [0, 4, 5, 7], [0, 4, 6, 7]])
M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def mix64(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def uniform01(seed, k, j):
    k = np.asarray(k, dtype=np.uint64)
    r = mix64(np.uint64(seed) ^ mix64((k << np.uint64(8)) | np.uint64(j)))
    return ((r >> np.uint64(40)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)


def corner_values(sigma):
    """[cells, 8] corner values, cells in x-fastest order (x = axis 0 of sigma)."""
    res = sigma.shape[0]
    n = res - 1
    out = np.empty((n, n, n, 8), dtype=np.float32)            # indexed [z][y][x]
    for i in range(8):
        dx, dy, dz = i & 1, (i >> 1) & 1, (i >> 2) & 1
        blk = sigma[dx:dx + n, dy:dy + n, dz:dz + n]          # [x][y][z]
        out[..., i] = np.transpose(blk, (2, 1, 0))
    return out.reshape(-1, 8)


def triangle_counts(sigma, level):
    v = corner_values(np.asarray(sigma, dtype=np.float32))
    above = v > np.float32(level)
    cnt = np.zeros(v.shape[0], dtype=np.int64)
    for t in TETS:
        inside = above[:, t].sum(1)
        cnt += np.where((inside == 0) | (inside == 4), 0, np.where(inside == 2, 2, 1))
    return cnt


def _edge_point(cx, cy, cz, v, a, b, level):
    va, vb = v[a], v[b]
    t = (np.float32(level) - va) / (vb - va)
    pa = np.array([cx + (a & 1), cy + ((a >> 1) & 1), cz + ((a >> 2) & 1)], dtype=np.float32)
    pb = np.array([cx + (b & 1), cy + ((b >> 1) & 1), cz + ((b >> 2) & 1)], dtype=np.float32)
    return pa + t * (pb - pa)


def cell_triangles(sigma, level, cell):
    res = sigma.shape[0]
    n = res - 1
    cx, cy, cz = cell % n, (cell // n) % n, cell // (n * n)
    v = np.array([sigma[cx + (i & 1), cy + ((i >> 1) & 1), cz + ((i >> 2) & 1)] for i in range(8)], dtype=np.float32)
    tris = []
    for tv in TETS:
        m = [bool(v[c] > np.float32(level)) for c in tv]
        inside = sum(m)
        if inside in (0, 4):
            continue
        if inside == 2:
            i_in = [tv[q] for q in range(4) if m[q]]
            i_out = [tv[q] for q in range(4) if not m[q]]
            ac, ad = _edge_point(cx, cy, cz, v, i_in[0], i_out[0], level), _edge_point(cx, cy, cz, v, i_in[0], i_out[1], level)
            bd, bc = _edge_point(cx, cy, cz, v, i_in[1], i_out[1], level), _edge_point(cx, cy, cz, v, i_in[1], i_out[0], level)
            tris += [(ac, ad, bd), (ac, bd, bc)]
        else:
            lone = [q for q in range(4) if m[q] == (inside == 1)][0]
            others = [tv[q] for q in range(4) if q != lone]
            tris.append(tuple(_edge_point(cx, cy, cz, v, tv[lone], o, level) for o in others))
    return tris


def surface_points(sigma, level, seed, n_points, thickness):
    """Returns (points [n,3] float32, n_triangles). Small grids only (python loop over points)."""
    sigma = np.asarray(sigma, dtype=np.float32)
    res = sigma.shape[0]
    cnt = triangle_counts(sigma, level)
    off = np.cumsum(cnt)
    n_tri = int(off[-1]) if len(off) else 0
    pts = np.zeros((n_points, 3), dtype=np.float32)
    if n_tri == 0:
        return pts, 0
    cache = {}
    for k in range(n_points):
        tri = k % n_tri
        cell = int(np.searchsorted(off, tri, side="right"))
        local = tri - (int(off[cell - 1]) if cell else 0)
        if cell not in cache:
            cache[cell] = cell_triangles(sigma, level, cell)
        p0, p1, p2 = cache[cell][local]
        u = [uniform01(seed, k, j) for j in range(3)]
        inv = np.float32(1.0) / (u[0] + u[1] + u[2])
        w = [x * inv for x in u]
        p = p0 * w[0] + p1 * w[1] + p2 * w[2]
        p = p * np.float32(1.0 / res) - np.float32(0.5)
        g0, g1 = uniform01(seed, k, 3), uniform01(seed, k, 4)
        normal = np.sqrt(np.float32(-2.0) * np.log(g0)) * np.cos(np.float32(6.283185307179586) * g1)
        scale = np.clip(normal * np.float32(thickness) + np.float32(1.0), 0.0, 1.0)
        pts[k] = p * np.float32(scale)
    return pts, n_tri
