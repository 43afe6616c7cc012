"""Point-set producer: numpy oracle properties on CPU, CUDA path against the oracle on GPU."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import surface_oracle as so  # noqa: E402


def sphere_grid(res, radius=0.3, sharp=200.0, level=10.0):
    """density = level + sharp * (radius - |p|), p = index / res - 0.5  ->  iso-surface |p| = radius"""
    ax = np.arange(res, dtype=np.float32) / res - 0.5
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    return (level + sharp * (radius - np.sqrt(x * x + y * y + z * z))).astype(np.float32)


def bumpy_grid(res, seed):
    rng = np.random.RandomState(seed)
    g = sphere_grid(res)
    return (g + 8.0 * rng.randn(res, res, res)).astype(np.float32)


def test_oracle_points_lie_on_the_sphere():
    res = 24
    n_tri = int(so.triangle_counts(sphere_grid(res), 10.0).sum())
    # point k sits on triangle k mod n_triangles (cells in grid order), as in the reference's repeat-and-truncate:
    # the whole surface is only covered once n_points >= n_triangles
    pts, n_tri2 = so.surface_points(sphere_grid(res), 10.0, seed=5, n_points=n_tri + 100, thickness=0.0)
    assert n_tri == n_tri2 and n_tri > 500
    r = np.linalg.norm(pts, axis=1)
    assert np.all(np.abs(r - 0.3) < 0.6 / res), (r.min(), r.max())      # within a fraction of a cell of the true surface
    for o in range(8):                                                     # every octant is populated
        s = np.array([1 if o & 1 else -1, 1 if o & 2 else -1, 1 if o & 4 else -1])
        assert np.sum(np.all(pts * s > 0, axis=1)) > 30


def test_oracle_triangle_count_and_jitter():
    res = 16
    g = sphere_grid(res)
    cnt = so.triangle_counts(g, 10.0)
    assert cnt.sum() == sum(len(so.cell_triangles(g, 10.0, c)) for c in np.nonzero(cnt)[0])
    assert so.triangle_counts(np.zeros((8, 8, 8), np.float32), 10.0).sum() == 0
    a, _ = so.surface_points(g, 10.0, seed=1, n_points=300, thickness=0.0)
    b, _ = so.surface_points(g, 10.0, seed=1, n_points=300, thickness=0.1)
    scale = np.linalg.norm(b, axis=1) / np.linalg.norm(a, axis=1)          # clip(N(1, 0.1), 0, 1): never outward
    assert np.all(scale <= 1.0 + 1e-6) and 0.3 < np.mean(scale < 0.999) < 0.7


@pytest.mark.gpu
@pytest.mark.parametrize("res,make", [(16, sphere_grid), (20, lambda r: bumpy_grid(r, 3)), (33, sphere_grid)])
def test_cuda_matches_oracle(cuda, res, make):
    from gaussian_gan_decoder_b200.decoder_utils.surface_points import surface_points
    g = make(res)
    n = 2500
    ref, n_tri_ref = so.surface_points(g, 10.0, seed=77, n_points=n, thickness=0.1)
    pts, n_tri = surface_points(torch.from_numpy(g).to(cuda), 10.0, n, 0.1, seed=77, return_num_triangles=True)
    assert n_tri == n_tri_ref                                              # integer stage: exact
    assert np.abs(pts.cpu().numpy() - ref).max() < 2e-5                   # float stage: rounding of a few operations


@pytest.mark.gpu
def test_cuda_full_size_properties(cuda):
    """The reference's size: 128^3 grid, 500k points (target_dataloader.py:31, :102)."""
    from gaussian_gan_decoder_b200.decoder_utils.surface_points import surface_points
    res = 128
    g = torch.from_numpy(sphere_grid(res)).to(cuda)
    pts, n_tri = surface_points(g, 10.0, 500_000, 0.0, seed=1, return_num_triangles=True)
    assert n_tri > 50_000
    r = pts.norm(dim=1)
    assert float((r - 0.3).abs().max()) < 0.6 / res
    again = surface_points(g, 10.0, 500_000, 0.0, seed=1)
    assert torch.equal(pts, again)                                          # reproducible
    other = surface_points(g, 10.0, 500_000, 0.0, seed=2)
    assert not torch.equal(pts, other)
    empty, n0 = surface_points(torch.zeros(32, 32, 32, device=cuda), 10.0, 100, 0.1, return_num_triangles=True)
    assert n0 == 0 and float(empty.abs().max()) == 0.0
    with pytest.raises(RuntimeError):
        surface_points(torch.zeros(8, 8, 8), 10.0, 10)
