"""CPU restatement of the reference rasteriser (oracle/raster_oracle.cpp): structural
invariants, and agreement with golden vectors captured from the reference's own CUDA
build on a B200 (tests/golden/raster_*.npz, made by tests/golden/make_raster_golden.py)."""
import glob
import os

import numpy as np
import pytest

from _harness import CpuOracle
from gaussian_gan_decoder_b200 import synthetic

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def oracle():
    return CpuOracle()


def test_higher_msb_matches_bit_length(oracle):
    # reference: getHigherMsb, rasterizer_impl.cu:35-50 == bit_length for n >= 1
    for n in list(range(1, 5000)) + [2 ** k + d for k in range(12, 31) for d in (-1, 0, 1)]:
        assert oracle.lib.oracle_higher_msb(n) == int(n).bit_length()


@pytest.mark.parametrize("P,size", [(1, 64), (3000, 96), (20000, 200)])
def test_forward_invariants(oracle, P, size):
    s = synthetic.raster_scene(P, size=size, seed_offset=P)
    R = oracle.forward(s)
    tt, off, rad = oracle.get("tiles_touched"), oracle.get("point_offsets"), oracle.get("radii")
    assert (np.cumsum(tt, dtype=np.uint32) == off).all() and (off[-1] if P else 0) == R
    assert ((rad > 0) == (tt > 0)).all()
    keys, vals = oracle.get("point_list_keys"), oracle.get("point_list")
    ku, vu = oracle.get("point_list_keys_unsorted"), oracle.get("point_list_unsorted")
    gx = (size + 15) // 16
    tiles = gx * ((size + 15) // 16)
    assert (np.diff(keys.astype(np.int64)) >= 0).all()                      # sorted
    order = np.argsort(ku, kind="stable")
    assert (ku[order] == keys).all() and (vu[order] == vals).all()         # stable
    dep = oracle.get("depths")
    assert ((keys & 0xffffffff).astype(np.uint32) == dep.view(np.uint32)[vals]).all()
    assert ((keys >> 32) < tiles).all()
    rng = oracle.get("ranges").reshape(-1, 2)
    tile_of = (keys >> 32).astype(np.int64)
    cnt = np.bincount(tile_of, minlength=tiles)
    assert ((rng[:, 1] - rng[:, 0]) == cnt).all()
    nz = cnt > 0
    assert (rng[nz, 0] == (np.cumsum(cnt) - cnt)[nz]).all() and (rng[~nz] == 0).all()
    nc = oracle.get("n_contrib").reshape(size, size)
    T = oracle.get("accum_alpha").reshape(size, size)
    len_px = np.repeat(np.repeat(cnt.reshape(-1, gx), 16, 0), 16, 1)[:size, :size]
    assert (nc <= len_px).all() and (T <= 1).all() and (T >= 0).all()
    img = oracle.get("out_color")
    assert np.isfinite(img).all()


def test_all_culled_gives_background(oracle):
    s = synthetic.raster_scene(100, size=64)
    s["means3D"] = s["means3D"] + 100.0       # far behind / outside
    R = oracle.forward(s)
    img = oracle.get("out_color").reshape(3, 64, 64)
    if R == 0:
        assert np.allclose(img, np.asarray(s["bg"])[:, None, None])


GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "raster_*.npz")))


@pytest.mark.parametrize("path", GOLDEN or [None])
def test_oracle_matches_reference_cuda_golden(oracle, path):
    if path is None:
        pytest.skip("no golden vectors captured yet")
    z = np.load(path)
    meta = {k: z[k].item() for k in ("W", "H", "tanfovx", "tanfovy", "sh_degree", "scale_modifier")}
    scene = {k: z["in_" + k] for k in ("bg", "means3D", "shs", "opacities", "scales", "rotations",
                                         "viewmatrix", "projmatrix", "campos") if "in_" + k in z}
    scene.update(meta)
    R = oracle.forward(scene)
    P = scene["means3D"].shape[0]
    # integers: equal except where a float sits within an ulp of a rounding boundary
    rad_bad = (oracle.get("radii") != z["radii"]).sum()
    assert rad_bad <= max(2, P // 20000), f"radii mismatches {rad_bad}"
    assert abs(R - int(z["R"])) <= max(8, int(z["R"]) // 5000)
    vis = z["radii"] > 0
    for name, tol in (("depths", 1e-6), ("means2D", 2e-4), ("conic_opacity", 1e-4), ("rgb", 1e-6)):
        a, b = oracle.get(name), z[name]
        w = a.size // P
        m = np.repeat(vis & (oracle.get("radii") > 0), w)
        rel = np.abs(a[m] - b[m]) / (np.abs(b[m]) + 1e-3)
        assert rel.max() <= tol * 50 and np.median(rel) <= tol, (name, rel.max())
    a, b = oracle.get("out_color"), z["out_color"].reshape(-1)
    mse = float(((a - b) ** 2).mean())
    psnr = 10 * np.log10(1.0 / max(mse, 1e-20))
    assert psnr >= 50.0, psnr
    nc_bad = (oracle.get("n_contrib") != z["n_contrib"]).mean()
    assert nc_bad <= 0.02, nc_bad


@pytest.mark.parametrize("path", GOLDEN or [None])
def test_oracle_backward_matches_reference_cuda_golden(oracle, path):
    """Gradients of the CPU restatement vs the reference CUDA backward (float atomics there,
    double accumulation here -> tolerance, stated per tensor)."""
    if path is None:
        pytest.skip("no golden vectors captured yet")
    z = np.load(path)
    meta = {k: z[k].item() for k in ("W", "H", "tanfovx", "tanfovy", "sh_degree", "scale_modifier")}
    scene = {k: z["in_" + k] for k in ("bg", "means3D", "shs", "opacities", "scales", "rotations",
                                         "viewmatrix", "projmatrix", "campos")}
    scene.update(meta)
    oracle.forward(scene)
    if (oracle.get("n_contrib") != z["n_contrib"]).mean() > 0.02:
        pytest.skip("forward differs too much for a gradient comparison")
    oracle.backward(z["in_dL_dpix"])
    for name in ("dL_dmean2D", "dL_dconic", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dsh",
                 "dL_dscale", "dL_drot"):
        a, b = oracle.get(name).reshape(-1), z[name].reshape(-1)
        assert a.shape == b.shape, name
        scale = np.abs(b).max() + 1e-30
        err = np.abs(a - b) / scale
        # a handful of pixels whose n_contrib differs by one move single Gaussians' gradients
        assert np.quantile(err, 0.999) < 2e-3 and np.median(err) < 1e-6, (name, err.max(), np.quantile(err, 0.999))
