"""Parity of the sm_100a rasteriser (through the reference-facing binding and the C ABI)
against (1) the UNMODIFIED reference CUDA rasteriser, bit for bit, and (2) the CPU oracle."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from _harness import CpuOracle, RefCuda, ref_cuda_available, run_ours
from gaussian_gan_decoder_b200 import _lib, synthetic

pytestmark = pytest.mark.gpu

INT_FIELDS = ["radii", "tiles_touched", "point_offsets", "point_list_keys_unsorted", "point_list_unsorted",
              "point_list_keys", "point_list", "ranges", "n_contrib"]
VIS_FIELDS = [("depths", 1), ("means2D", 2), ("conic_opacity", 4), ("rgb", 3), ("cov3D", 6)]


def to_dev(scene, dev):
    return {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in scene.items()}


def random_init_like(scene, seed):
    """attributes in the range a random-init decoder emits: tiny scales, opacity ~0.5"""
    g = synthetic.gen(seed)
    P = scene["means3D"].shape[0]
    scene["scales"] = torch.exp(-7.0 + 0.1 * torch.randn(P, 3, generator=g))
    scene["opacities"] = torch.sigmoid(0.2 * torch.randn(P, 1, generator=g))
    return scene


def assert_bit_exact(ours, ref, P):
    assert ours["R"] == ref["R"]
    for n in INT_FIELDS:
        if n in ref:
            a, b = ours[n].view(-1), ref[n].view(-1)
            assert a.shape == b.shape, n
            bad = int((a != b).sum())
            assert bad == 0, f"{n}: {bad} / {a.numel()} differ"
    vis = ref["radii"] > 0
    for n, w in VIS_FIELDS:
        a = ours[n].view(P, w)[vis].view(torch.int32)
        b = ref[n].view(P, w)[vis].view(torch.int32)
        bad = int((a != b).sum())
        assert bad == 0, f"{n}: {bad} / {a.numel()} float bit patterns differ"
    for n in ("accum_alpha", "out_color"):
        a, b = ours[n].contiguous().view(-1).view(torch.int32), ref[n].contiguous().view(-1).view(torch.int32)
        bad = int((a != b).sum())
        assert bad == 0, f"{n}: {bad} / {a.numel()} float bit patterns differ"


SCENES = [
    # P, size, fov, trained_like
    (1, 64, 12.0, True),
    (257, 100, 12.0, True),          # ragged: image not a multiple of 16, P not a multiple of 256
    (5000, 256, 12.0, True),
    (50000, 512, 12.0, False),
    (100000, 512, 17.0, True),
]


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("P,size,fov,trained", SCENES)
def test_bit_exact_vs_reference_cuda(cuda, P, size, fov, trained):
    s = synthetic.raster_scene(P, size=size, fov_deg=fov, seed_offset=P % 97)
    if not trained:
        s = random_init_like(s, P)
    s = to_dev(s, cuda)
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    assert_bit_exact(ours, ref, P)
    # second call takes the sync-free path (capacity hint) and must give the same answer
    ours2 = run_ours(s, cuda)
    assert_bit_exact(ours2, ref, P)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("W,H", [(200, 120), (120, 200), (333, 17), (16, 250)])
def test_bit_exact_non_square_images(cuda, W, H):
    """image_width != image_height (the settings carry both; every scene above is square): the same matrices with another
    pixel grid are a distorted but valid camera for both implementations."""
    s = synthetic.raster_scene(3000, size=max(W, H), fov_deg=14.0, seed_offset=W + H)
    s = to_dev(s, cuda)
    s["W"], s["H"] = W, H
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    assert ref["R"] > 0
    assert_bit_exact(ours, ref, 3000)
    assert ours["out_color"].shape == (3, H, W)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
def test_bit_exact_full_size_config2(cuda):
    """BASELINE config 2 shape: 500k Gaussians, 512x512 — both attribute distributions."""
    for trained in (False, True):
        s = synthetic.raster_scene(500_000, size=512, seed_offset=2)
        if not trained:
            s = random_init_like(s, 7)
        s = to_dev(s, cuda)
        ref = RefCuda().forward(s, cuda)
        ours = run_ours(s, cuda)
        assert_bit_exact(ours, ref, 500_000)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("mode", [0, 1])
def test_bit_exact_config5_2048_2M(cuda, mode):
    """BASELINE configs[4]: 2M trained-like Gaussians at 2048x2048 (16384 tiles -> 47-bit keys, ~2.8e8 instances,
    tile lists beyond the 24576 words the long-list sort holds at once, the large-capacity digit pass) — every integer
    and float stage bit for bit against the reference CUDA build, with the radix sort pinned (mode 0) and with the
    device-side choice (mode 1: per-tile binning, the longest lists sorted in depth chunks).  Follows rasterizer_impl.cu:300-308 (getHigherMsb, sort bit range) and
    forward.cu:261-374."""
    lib = _lib.load()
    P = 2_000_000
    s = to_dev(synthetic.raster_scene(P, size=2048, fov_deg=12.0, seed_offset=50), cuda)
    ref = RefCuda().forward(s, cuda)
    assert ref["R"] > 100_000_000
    from gaussian_gan_decoder_b200 import rasterizer as rz
    rz._binning_hint.pop((cuda.index, P, 2048, 2048), None)     # forget what an earlier run learnt about this shape
    prev = lib.ggd_raster_binning(mode)
    try:
        ours = run_ours(s, cuda)
        info = (C.c_uint32 * 2)()
        _lib.check(lib.ggd_raster_binning_info(None, C.c_void_p(ours["_buffers"][0].data_ptr()), P, info))
    finally:
        lib.ggd_raster_binning(prev)
    assert_bit_exact(ours, ref, P)
    rng = ours["ranges"].view(-1, 2).long()
    assert rng.shape[0] == 128 * 128 and int((rng[:, 1] - rng[:, 0]).max()) > 24576
    if mode == 1:    # lists beyond the 24576-word shared-memory buffer are sorted per tile in depth chunks
        assert info[0] == 1 and 24576 < info[1] <= 98304
    del ours, ref
    torch.cuda.empty_cache()


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
def test_bit_exact_many_random_gaussians_preprocess(cuda):
    """>= 10^7 random Gaussians through the projection stage: zero integer mismatches."""
    ref_lib = RefCuda()
    total = 0
    for rep in range(5):
        P = 2_000_000
        s = synthetic.raster_scene(P, size=512, fov_deg=5 + 3 * rep, seed_offset=100 + rep)
        g = synthetic.gen(200 + rep)
        s["scales"] = torch.exp(-6.5 + 1.2 * torch.randn(P, 3, generator=g))
        s["means3D"] = s["means3D"] * (1.0 + 0.5 * rep)   # push some outside the frustum / image
        s = to_dev(s, cuda)
        ref = ref_lib.forward(s, cuda)
        ours = run_ours(s, cuda)
        for n in ("radii", "tiles_touched", "point_offsets"):
            assert int((ours[n].view(-1) != ref[n].view(-1)).sum()) == 0, n
        vis = ref["radii"] > 0
        for n, w in VIS_FIELDS:
            a = ours[n].view(P, w)[vis].view(torch.int32); b = ref[n].view(P, w)[vis].view(torch.int32)
            assert int((a != b).sum()) == 0, n
        total += P
    assert total >= 10_000_000


@pytest.mark.parametrize("P,size", [(3000, 96), (20000, 200)])
def test_vs_cpu_oracle(cuda, P, size):
    s = synthetic.raster_scene(P, size=size, seed_offset=P)
    o = CpuOracle()
    R = o.forward(s)
    ours = run_ours(to_dev(s, cuda), cuda)
    assert abs(ours["R"] - R) <= max(4, R // 5000)
    bad = int((ours["radii"].cpu().numpy() != o.get("radii")).sum())
    assert bad <= max(2, P // 20000)
    a, b = ours["out_color"].cpu().numpy().reshape(-1), o.get("out_color")
    psnr = 10 * np.log10(1.0 / max(float(((a - b) ** 2).mean()), 1e-20))
    assert psnr >= 50.0, psnr
    nc_bad = (ours["n_contrib"].cpu().numpy() != o.get("n_contrib")).mean()
    assert nc_bad <= 0.02


def test_empty_and_all_culled(cuda):
    s = to_dev(synthetic.raster_scene(64, size=64), cuda)
    s0 = dict(s)
    for k in ("means3D", "scales", "rotations", "opacities", "shs"):
        s0[k] = s[k][:0]
    out = run_ours(s0, cuda)
    assert out["R"] == 0 and out["out_color"].shape == (3, 64, 64) and float(out["out_color"].abs().sum()) == 0.0
    s1 = dict(s)
    s1["means3D"] = s["means3D"] + 100.0
    out = run_ours(s1, cuda)
    assert out["R"] == 0 and int((out["radii"] != 0).sum()) == 0
    bg = s["bg"].view(3, 1, 1).expand(3, 64, 64)
    assert torch.equal(out["out_color"], bg)
    assert int(out["ranges"].abs().sum()) == 0


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
def test_alternative_inputs_bit_exact(cuda):
    """colors_precomp, cov3D_precomp and SH degree 3 branches (on the API surface)."""
    P = 4000
    base = to_dev(synthetic.raster_scene(P, size=160, seed_offset=5), cuda)
    g = synthetic.gen(9)
    # (a) precomputed colours
    s = dict(base); s["colors_precomp"] = torch.rand(P, 3, generator=g).to(cuda); s["shs"] = None
    ref = RefCuda().forward(s, cuda); ours = run_ours(s, cuda)
    for n in INT_FIELDS + ["out_color"]:
        assert torch.equal(ours[n].view(-1), ref[n].view(-1)), n
    # (b) precomputed 3D covariance (taken from a scale/rotation run)
    ref0 = RefCuda().forward(base, cuda)
    s = dict(base); s["cov3D_precomp"] = ref0["cov3D"].view(P, 6).clone(); s["scales"] = None; s["rotations"] = None
    ref = RefCuda().forward(s, cuda); ours = run_ours(s, cuda)
    for n in INT_FIELDS + ["out_color"]:
        assert torch.equal(ours[n].view(-1), ref[n].view(-1)), n
    # (c) SH degree 3, 16 coefficients
    s = dict(base); s["shs"] = (0.3 * torch.randn(P, 16, 3, generator=g)).to(cuda); s["sh_degree"] = 3
    ref = RefCuda().forward(s, cuda); ours = run_ours(s, cuda)
    vis = ref["radii"] > 0
    for n in INT_FIELDS + ["out_color"]:
        assert torch.equal(ours[n].view(-1), ref[n].view(-1)), n
    assert torch.equal(ours["rgb"].view(P, 3)[vis], ref["rgb"].view(P, 3)[vis])
    assert torch.equal(ours["clamped"].view(P, 3)[vis], ref["clamped"].view(P, 3)[vis])


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("scale_modifier", [0.6, 1.7])
@pytest.mark.parametrize("deg,M", [(0, 16), (1, 4), (1, 16), (2, 9), (2, 16)])
def test_scale_modifier_and_partial_sh_degrees(cuda, scale_modifier, deg, M):
    """GaussianRasterizationSettings.scale_modifier != 1 (the viewers' splat-size slider) and an active SH degree below
    the stored one (3DGS raises it during training), as well as M = (deg + 1)^2 exactly."""
    P = 3000
    s = to_dev(synthetic.raster_scene(P, size=144, seed_offset=int(10 * scale_modifier) + deg + M), cuda)
    g = synthetic.gen(100 + deg + M)
    s["shs"] = (0.3 * torch.randn(P, M, 3, generator=g)).to(cuda)
    s["sh_degree"] = deg
    s["scale_modifier"] = scale_modifier
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    assert ref["R"] > 0
    assert_bit_exact(ours, ref, P)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("spread,size", [(1.5, 256), (2.5, 192)])
def test_bit_exact_elongated_rotated_splats(cuda, spread, size):
    """The blend skips (8x4 pixel block, instance) pairs whose pixels all fall below alpha = 1/255 — decided from the splat's
    box and, for larger splats, from the minimum of its quadratic form over the block.  Needle-like, rotated splats are
    where that test has to be conservative: per-axis log-scales N(-5, spread) give axis ratios beyond 100:1 (the most
    ill-conditioned ones take the 'never culled' branch).  Image, T and contributor counts must keep the reference's bits."""
    P = 20_000
    s = synthetic.raster_scene(P, size=size, fov_deg=14.0, seed_offset=int(10 * spread))
    g = synthetic.gen(500 + int(10 * spread))
    s["scales"] = torch.exp(-5.0 + spread * torch.randn(P, 3, generator=g)).clamp(max=0.5)
    s["opacities"] = torch.sigmoid(1.0 + 2.5 * torch.randn(P, 1, generator=g))
    s = to_dev(s, cuda)
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    assert ref["R"] > 10 * P
    assert_bit_exact(ours, ref, P)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("P,W,H,trained", [(20_000, 256, 256, True), (50_000, 512, 512, False), (3000, 333, 100, True), (257, 40, 24, True)])
def test_super_tile_scatter_bit_exact(cuda, monkeypatch, P, W, H, trained):
    """Images beyond 1024^2 bin their instances in two levels (8x8-tile super-tiles first: super_bin / super_emit kernels);
    GGD_SCATTER_SUPER=1 forces that path on small images, where the reference comparison is cheap (the 2048^2 test takes it
    by itself).  Everything the reference returns must keep its bits, the emission-order arrays included; a second call and
    a re-run of stage 2 on the same geometry state must give the same frame."""
    monkeypatch.setenv("GGD_SCATTER_SUPER", "1")
    s = synthetic.raster_scene(P, size=max(W, H), fov_deg=13.0, seed_offset=P % 53)
    if not trained:
        s = random_init_like(s, P)
    s = to_dev(s, cuda)
    s["W"], s["H"] = W, H
    ref = RefCuda().forward(s, cuda)
    for _ in range(2):
        ours = run_ours(s, cuda)
        assert_bit_exact(ours, ref, P)
    monkeypatch.setenv("GGD_SCATTER_SUPER", "0")
    other = run_ours(s, cuda)
    assert torch.equal(other["out_color"], ours["out_color"]) and torch.equal(other["point_list"], ours["point_list"])


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("size,offset", [(256, 0.05), (528, 0.6), (1104, 0.3)])
def test_bit_exact_camera_inside_the_point_cloud(cuda, size, offset):
    """Every other scene looks at the head from 2.7 units away.  Here the camera sits inside it: about half of the Gaussians
    are behind the near plane (culled at z <= 0.2, auxiliary.h:141-164), the ones just in front of it project to splats
    of hundreds of pixels with centres far outside the image."""
    P = 40_000
    s = synthetic.raster_scene(P, size=size, fov_deg=40.0, seed_offset=size)
    V = s["viewmatrix"].double()
    proj = torch.linalg.inv(V) @ s["projmatrix"].double()
    V2 = V.clone()
    V2[3, :3] = torch.tensor([0.02, -0.03, offset], dtype=torch.float64)         # row-vector convention: translation in the last row
    s["viewmatrix"] = V2.float().contiguous()
    s["projmatrix"] = (V2 @ proj).float().contiguous()
    s["campos"] = torch.linalg.inv(V2)[3, :3].float().contiguous()
    s = to_dev(s, cuda)
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    culled = int((ref["radii"] == 0).sum())
    assert 0.2 * P < culled < 0.97 * P, culled
    assert int(ref["radii"].max()) > size // 2
    assert_bit_exact(ours, ref, P)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("size", [256, 1104])
def test_non_finite_and_absurd_gaussians_like_the_reference(cuda, size):
    """NaN / infinite means and opacities, and splats of astronomical size (their radius saturates the int it is stored in):
    the reference culls the former through an empty tile rectangle and spreads the latter over every tile.  Same bits
    here, whichever binning path the image size selects.
    NaN / infinite SCALES are left out of the comparison: the reference gives them radius (int)NaN = 0 but one touched
    tile, and duplicateWithKeys then skips them (radius 0) — their key slots keep whatever the buffer held.  Here they are
    culled: the frame is the frame without them (second half)."""
    P = 30_000
    s = synthetic.raster_scene(P, size=size, fov_deg=14.0, seed_offset=size)
    nan, inf = float("nan"), float("inf")
    m, sc, op = s["means3D"].clone(), s["scales"].clone(), s["opacities"].clone()
    m[10] = nan; m[11, 0] = inf; m[12, 2] = -inf; m[13, 1] = 1e30; m[14] = torch.tensor([1e20, -1e20, 1e20])
    sc[22] = 1e10; sc[23, 1] = 1e5; sc[24] = 0.0; sc[25] = torch.tensor([3e3, 1e-9, 2.0]); sc[26] = 40.0
    op[30] = nan; op[31] = inf; op[32] = -1.0; op[33] = 0.0; op[22] = 0.05; op[23] = 0.05; op[25] = 0.05; op[26] = 0.03
    s["means3D"], s["scales"], s["opacities"] = m, sc, op
    s = to_dev(s, cuda)
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    assert ref["R"] > 0
    assert_bit_exact(ours, ref, P)
    sc = s["scales"].clone()
    sc[20] = nan; sc[21, 0] = inf
    s["scales"] = sc
    worse = run_ours(s, cuda)
    assert worse["R"] == ours["R"] - int(ours["tiles_touched"][20]) - int(ours["tiles_touched"][21])
    assert int(worse["radii"][20]) == 0 and int(worse["tiles_touched"][20]) == 0 and int(worse["tiles_touched"][21]) == 0
    gone = torch.ones(P, dtype=torch.bool, device=cuda)
    gone[20:22] = False
    without = {k: (v[gone].contiguous() if torch.is_tensor(v) and v.dim() > 0 and v.shape[0] == P else v) for k, v in s.items()}
    expect = run_ours(without, cuda)
    assert torch.equal(worse["out_color"], expect["out_color"]) and torch.equal(worse["n_contrib"], expect["n_contrib"])


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("W,H", [(1100, 1050), (2050, 520), (2100, 2100), (4100, 600)])      # the third: more than 16 384 tiles
def test_bit_exact_large_image_mixed_splat_sizes(cuda, W, H):
    """Beyond 4096 tiles the frame takes the difference-array tile counting, the super-tile scatter and (lean binding) the
    lazy sort by itself.  Image sides that are no multiple of 16 (nor of the 128-pixel super-tiles), a few hundred splats
    that cover most of the image next to needle-like and ordinary ones: every output keeps the reference's bits, and the
    lean binding renders the same frame."""
    from gaussian_gan_decoder_b200 import rasterizer as rz
    P = 300_000
    s = synthetic.raster_scene(P, size=max(W, H), fov_deg=13.0, seed_offset=W)
    g = synthetic.gen(W + H)
    scales = s["scales"].clone()
    scales[:300] = torch.exp(-1.2 + 0.4 * torch.randn(300, 3, generator=g))                  # cover hundreds to thousands of tiles
    scales[300:20_300] = torch.exp(-5.0 + 2.0 * torch.randn(20_000, 3, generator=g)).clamp(max=0.3)   # needles
    s["scales"] = scales
    op = s["opacities"].clone()
    op[:300] = 0.02 + 0.1 * torch.rand(300, 1, generator=g)                                   # faint: the lists stay long
    s["opacities"] = op
    s = to_dev(s, cuda)
    s["W"], s["H"] = W, H
    assert ((W + 15) // 16) * ((H + 15) // 16) > 4096
    ref = RefCuda().forward(s, cuda)
    ours = run_ours(s, cuda)
    info = (C.c_uint32 * 2)()
    _lib.check(_lib.load().ggd_raster_binning_info(None, C.c_void_p(ours["_buffers"][0].data_ptr()), P, info))
    assert info[0] == 1 and info[1] > 4096, tuple(info)        # per-tile path with long lists
    assert ref["R"] > 20 * P
    assert_bit_exact(ours, ref, P)
    old = rz.keep_reference_order_arrays
    rz.keep_reference_order_arrays = False
    try:
        e = torch.empty(0, device=cuda)
        lean = rz._C.rasterize_gaussians(s["bg"], s["means3D"], e, s["opacities"], s["scales"], s["rotations"], 1.0, e, s["viewmatrix"],
                                         s["projmatrix"], s["tanfovx"], s["tanfovy"], H, W, s["shs"], 0, s["campos"], False, False)
    finally:
        rz.keep_reference_order_arrays = old
    assert lean[0] == ours["R"] and torch.equal(lean[1], ours["out_color"]) and torch.equal(lean[2], ours["radii"])
    cap = getattr(lean[4], "ggd_capacity", 0)
    for name, dt in (("accum_alpha", torch.float32), ("n_contrib", torch.int32)):
        assert torch.equal(rz.field(lean[5], 2, name, P, W, H, cap, dt), ours[name]), name


@pytest.mark.parametrize("backward,scene", [(False, "head"), (True, "head"), (False, "depth_layers"), (True, "depth_layers"), (False, "index_by_depth"),
                                            (False, "head_longer_lists"), (True, "head_chunked_lists")])
def test_lean_large_image_sorts_lazily_same_frame(cuda, backward, scene):
    """GGD_BINNING_LEAN on an image beyond 1024^2: long tile lists are sorted only in their nearest depth chunk, the blend
    marks the tiles that did not saturate inside it, those are sorted completely and blended again.  Image, final T and
    contributor counts must be the bits of the complete path, and the backward (which replays the lists from the last
    contributor) must see the same gradients.
    The chunk's depth bound comes from a sample of the list: `depth_layers` puts every Gaussian on one of three depths (a
    bound then takes a whole layer or nothing: the chunk overflows or is empty and the tile goes to the second round),
    `index_by_depth` orders the Gaussians, hence the unsorted lists, by depth (a sample taken at regular places of the list
    must still see the whole interval), `head_longer_lists` zooms in until lists pass 4 x 4096 words (the sample then takes
    less than a quarter of a list), `head_chunked_lists` adds Gaussians until the longest pass the 24 576 words the complete sort holds at
    once (the second round then sorts a marked tile in depth chunks)."""
    from gaussian_gan_decoder_b200 import rasterizer as rz
    _C = rz._C
    P, S = (2_200_000 if scene == "head_chunked_lists" else 1_500_000), 1040      # 65 x 65 tiles, ~13k instances per tile
    s = synthetic.raster_scene(P, size=S, fov_deg={"head_longer_lists": 9.0, "head_chunked_lists": 9.0}.get(scene, 12.0), seed_offset=77)
    if scene in ("depth_layers", "index_by_depth"):
        V = s["viewmatrix"]
        d = V[:3, 2] / V[:3, 2].norm()
        z = s["means3D"] @ d
        if scene == "depth_layers":
            layer = torch.bucketize(z, torch.quantile(z[::97], torch.tensor([1 / 3, 2 / 3])))
            target = torch.stack([z[layer == k].mean() for k in range(3)])[layer]
            s["means3D"] = (s["means3D"] + (target - z)[:, None] * d[None, :]).contiguous()
        else:
            order = torch.argsort(z)
            s = {k: (v[order].contiguous() if torch.is_tensor(v) and v.dim() > 0 and v.shape[0] == P else v) for k, v in s.items()}
    s = to_dev(s, cuda)
    e = torch.empty(0, device=cuda)

    def run(keep):
        old = rz.keep_reference_order_arrays
        rz.keep_reference_order_arrays = keep
        try:
            out = rz._C.rasterize_gaussians(s["bg"], s["means3D"], e, s["opacities"], s["scales"], s["rotations"], 1.0, e, s["viewmatrix"],
                                            s["projmatrix"], s["tanfovx"], s["tanfovy"], S, S, s["shs"], 0, s["campos"], False, False)
        finally:
            rz.keep_reference_order_arrays = old
        return out

    full, lean = run(True), run(False)
    R, cap = full[0], getattr(full[4], "ggd_capacity", 0)
    info = (C.c_uint32 * 2)()
    _lib.check(_lib.load().ggd_raster_binning_info(None, _lib.ptr(lean[3]), P, info))
    assert info[0] == 1 and info[1] > {"head_longer_lists": 16384, "head_chunked_lists": 24576}.get(scene, 12288), tuple(info)   # binned per tile, lists long enough for the lazy path
    assert lean[0] == R and torch.equal(lean[1], full[1]) and torch.equal(lean[2], full[2])
    for name, dt in (("accum_alpha", torch.float32), ("n_contrib", torch.int32), ("ranges", torch.int32)):
        a = rz.field(full[5], 2, name, P, S, S, cap, dt)
        b = rz.field(lean[5], 2, name, P, S, S, cap, dt)
        assert torch.equal(a, b), name
    if backward:
        dL = torch.randn(3, S, S, generator=synthetic.gen(5)).to(cuda)
        grads = []
        for o in (full, full, full, lean):     # the backward sums with float atomics: its own run-to-run spread is the yardstick
            grads.append(_C.rasterize_gaussians_backward(s["bg"], s["means3D"], o[2], e, s["scales"], s["rotations"], 1.0, e, s["viewmatrix"],
                                                         s["projmatrix"], s["tanfovx"], s["tanfovy"], dL, s["shs"], 0, s["campos"], o[3], R,
                                                         o[4], o[5], False))
        torch.cuda.synchronize()
        for ga, gb, gb2, gc in zip(*grads):
            if ga.numel():
                scale = float(ga.abs().max()) + 1e-30
                spread = max(float((ga - gb).abs().max()), float((ga - gb2).abs().max()), float((gb - gb2).abs().max()))
                assert float((gc - ga).abs().max()) <= max(8.0 * spread, 3e-4 * scale), (float((gc - ga).abs().max()), spread, scale)
                assert spread <= 5e-3 * scale      # (the yardstick itself stays small: 1e-3 of the largest gradient was seen)


def test_capacity_overflow_recovers(cuda):
    from gaussian_gan_decoder_b200 import rasterizer as rz
    s = to_dev(synthetic.raster_scene(20000, size=256, seed_offset=3), cuda)
    a = run_ours(s, cuda)
    key = (cuda.index, 20000, 256, 256)
    rz._capacity_hint[key] = 4096          # far too small: forces the overflow path
    b = run_ours(s, cuda)
    assert a["R"] == b["R"] and a["R"] > 4096
    assert torch.equal(a["out_color"], b["out_color"]) and torch.equal(a["point_list"], b["point_list"])


def test_size_independent_properties_full_size(cuda):
    """500k Gaussians @ 512^2: sortedness, stability, range partition, contributor bounds."""
    s = to_dev(synthetic.raster_scene(500_000, size=512, seed_offset=11), cuda)
    o = run_ours(s, cuda)
    R = o["R"]
    keys, vals = o["point_list_keys"], o["point_list"].long()
    assert R == int(o["point_offsets"].view(-1)[-1]) and R > 0
    assert bool((keys[1:] >= keys[:-1]).all())
    ku, vu = o["point_list_keys_unsorted"], o["point_list_unsorted"].long()
    # stability: among equal keys Gaussian ids ascend (emission order is ascending id)
    same = keys[1:] == keys[:-1]
    assert bool((vals[1:][same] > vals[:-1][same]).all())
    # multiset of (key, value) is preserved
    mix = lambda k, v: (k * 1000003 + v).sum().item()
    assert mix(ku, vu) == mix(keys, vals)
    assert torch.equal(torch.sort(vu).values, torch.sort(vals).values)
    tiles = 32 * 32
    cnt = torch.bincount((keys >> 32), minlength=tiles)
    rng = o["ranges"].view(-1, 2).long()
    assert torch.equal(rng[:, 1] - rng[:, 0], cnt)
    assert bool((o["n_contrib"].view(512, 512).long() <= cnt.view(32, 32).repeat_interleave(16, 0).repeat_interleave(16, 1)).all())
    assert bool(torch.isfinite(o["out_color"]).all())


def _sort_case(cuda, n, begin, end, seed, cap=None):
    lib = _lib.load()
    cap = cap if cap is not None else max(n, 1)
    g = torch.Generator(device="cpu"); g.manual_seed(seed)
    keys = torch.randint(0, 2 ** 62, (cap,), generator=g, dtype=torch.int64)
    if seed % 2:                                   # many duplicates -> exercises stability
        keys = keys % 1000
    vals = torch.arange(cap, dtype=torch.int32)
    kd, vd = keys.to(cuda), vals.to(cuda)
    ko, vo, kt, vt = torch.empty_like(kd), torch.empty_like(vd), torch.empty_like(kd), torch.empty_like(vd)
    nd = torch.tensor([n], dtype=torch.int32, device=cuda)
    sb = lib.ggd_sort_scratch_bytes(cap)
    scratch = torch.empty(sb, dtype=torch.uint8, device=cuda)
    p = lambda t: C.c_void_p(t.data_ptr())
    _lib.check(lib.ggd_sort_pairs(C.c_void_p(torch.cuda.current_stream().cuda_stream), p(kd), p(ko), p(vd), p(vo),
                                  p(nd), cap, begin, end, p(kt), p(vt), p(scratch), sb))
    torch.cuda.synchronize()
    k = keys[:n].numpy().astype(np.uint64)
    mask = np.uint64((1 << (end - begin)) - 1)
    digit = (k >> np.uint64(begin)) & mask
    order = np.argsort(digit, kind="stable")
    assert (ko[:n].cpu().numpy().astype(np.uint64) == k[order]).all()
    assert (vo[:n].cpu().numpy() == vals[:n].numpy()[order]).all()
    assert torch.equal(kd.cpu(), keys)             # input preserved


@pytest.mark.parametrize("n", [0, 1, 31, 4095, 4096, 4097, 100_000, 3_000_001])
def test_sort_pairs_sizes(cuda, n):
    _sort_case(cuda, n, 0, 43, seed=n)


@pytest.mark.parametrize("n,cap,end", [(4_194_304, 4_194_304, 47), (5_000_001, 6_000_000, 47), (4_500_000, 4_500_000, 43)])
def test_sort_pairs_large_capacity_variant(cuda, n, cap, end):
    """capacity >= 4 Mi pairs selects the 4-CTAs-per-SM digit pass (onesweep_kernel<4>, binning.cu); 47 key bits
    are what a 2048^2 frame sorts on (32 depth bits + getHigherMsb(16384 tiles) = 15)."""
    _sort_case(cuda, n, 0, end, seed=n % 1000 + 2, cap=cap)


@pytest.mark.parametrize("begin,end", [(0, 8), (0, 64), (5, 17), (32, 47), (0, 1), (3, 44)])
def test_sort_pairs_bit_ranges(cuda, begin, end):
    _sort_case(cuda, 50_000, begin, end, seed=begin * 64 + end)
    _sort_case(cuda, 50_001, begin, end, seed=begin * 64 + end + 1)


def test_sort_pairs_count_below_capacity(cuda):
    _sort_case(cuda, 70_000, 0, 43, seed=4, cap=200_000)


def test_sort_rejects_oversize(cuda):
    lib = _lib.load()
    assert lib.ggd_sort_pairs(None, *([C.c_void_p(8)] * 5), 1 << 30, 0, 43, C.c_void_p(8), C.c_void_p(8),
                              C.c_void_p(8), 0) == _lib.GGD_E_INVALID


def test_mark_visible(cuda):
    from gaussian_gan_decoder_b200.rasterizer import GaussianRasterizer, GaussianRasterizationSettings
    s = to_dev(synthetic.raster_scene(5000, size=64), cuda)
    pts = torch.cat([s["means3D"], s["means3D"] + torch.tensor([0.0, 0.0, 50.0], device=cuda)])
    rs = GaussianRasterizationSettings(64, 64, s["tanfovx"], s["tanfovy"], s["bg"], 1.0, s["viewmatrix"],
                                       s["projmatrix"], 0, s["campos"], False, False)
    vis = GaussianRasterizer(rs).markVisible(pts)
    view = s["viewmatrix"]
    z = pts @ view[:3, 2] + view[3, 2]
    expect = z > 0.2
    assert vis.dtype == torch.bool and int((vis != expect).sum()) <= 2


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
def test_mark_visible_bit_exact_vs_reference_cuda(cuda):
    """markVisible against the reference's own kernel (rasterizer_impl.cu:54-66, 141-152), including
    points behind the camera, on the z = 0.2 cut and far outside the frustum."""
    from gaussian_gan_decoder_b200.rasterizer import GaussianRasterizer, GaussianRasterizationSettings
    s = to_dev(synthetic.raster_scene(300_001, size=512, seed_offset=21), cuda)
    g = synthetic.gen(22)
    spread = (torch.randn(300_001, 3, generator=g) * torch.tensor([1.0, 1.0, 4.0])).to(cuda)
    pts = torch.cat([s["means3D"], s["means3D"] + spread])
    rs = GaussianRasterizationSettings(512, 512, s["tanfovx"], s["tanfovy"], s["bg"], 1.0, s["viewmatrix"],
                                       s["projmatrix"], 0, s["campos"], False, False)
    ours = GaussianRasterizer(rs).markVisible(pts)
    ref = RefCuda().mark_visible(pts, s["viewmatrix"], s["projmatrix"])
    assert 0 < int(ref.sum()) < pts.shape[0]
    assert torch.equal(ours, ref)


@pytest.mark.parametrize("P,size,trained,expect_per_tile", [
    (257, 100, True, True), (50000, 512, False, True), (500_000, 512, False, True),
    (500_000, 512, True, True),      # trained-like splats: tile lists of 5-10k instances (long-list kernel)
    (400_000, 64, True, True),       # 16 tiles only: lists of ~50k, sorted per tile in depth chunks of <= 24576
    (1_200_000, 32, True, False),    # 4 tiles: lists far longer than the per-tile path takes at all (> 98304)
])
def test_binning_modes_agree(cuda, P, size, trained, expect_per_tile):
    """Per-tile binning (tile_binning.cu) and the global radix sort give the same bits; the per-frame choice
    is made on the device from the longest tile list."""
    lib = _lib.load()
    s = synthetic.raster_scene(P, size=size, seed_offset=11)
    if not trained:
        s = random_init_like(s, 5)
    s = to_dev(s, cuda)
    prev = lib.ggd_raster_binning(1)
    try:
        auto = run_ours(s, cuda)
        info = (C.c_uint32 * 2)()
        _lib.check(lib.ggd_raster_binning_info(None, C.c_void_p(auto["_buffers"][0].data_ptr()), P, info))
        assert bool(info[0]) == expect_per_tile, (info[0], info[1])
        assert (info[1] <= 98304) == expect_per_tile      # kTileSortHugeMax
        lib.ggd_raster_binning(0)
        radix = run_ours(s, cuda)
    finally:
        lib.ggd_raster_binning(prev)
    assert_bit_exact(auto, radix, P)
    counts = (auto["ranges"].view(-1, 2)[:, 1] - auto["ranges"].view(-1, 2)[:, 0])
    assert int(counts.max()) == info[1]
    assert int(counts.sum()) == auto["R"]


@pytest.mark.parametrize("P,longest_at_least", [(60_000, 64), (400_000, 4097)])
def test_binning_clustered_depths(cuda, P, longest_at_least):
    """All splats at (nearly) one depth: every tile list collapses into a few buckets of the distribution
    sort, which must hand over to its comparison network; ties in depth are ordered by Gaussian index.
    The second size puts the lists into the long-list kernel."""
    lib = _lib.load()
    s = synthetic.raster_scene(P, size=256, seed_offset=3)
    V = s["viewmatrix"]
    d = V[:3, 2] / V[:3, 2].norm()
    s["means3D"] = (s["means3D"] - (s["means3D"] @ d)[:, None] * d[None, :]).contiguous()
    s = to_dev(s, cuda)
    prev = lib.ggd_raster_binning(1)
    try:
        auto = run_ours(s, cuda)
        info = (C.c_uint32 * 2)()
        _lib.check(lib.ggd_raster_binning_info(None, C.c_void_p(auto["_buffers"][0].data_ptr()), P, info))
        assert info[0] == 1 and info[1] >= longest_at_least, (info[0], info[1])
        lib.ggd_raster_binning(0)
        radix = run_ours(s, cuda)
    finally:
        lib.ggd_raster_binning(prev)
    assert auto["R"] > 0
    depth_bits = auto["point_list_keys"] & 0xffffffff
    assert depth_bits.unique().numel() < 64            # the scene really is degenerate in depth
    assert_bit_exact(auto, radix, P)


@pytest.mark.parametrize("super_scatter", ["0", "1"])
def test_stage2_rerun_on_same_geometry_state(cuda, monkeypatch, super_scatter):
    """ggd_raster_render may be called again on the state ggd_raster_preprocess left (the binding does that
    after a capacity overflow; a caller may do it to change the background): same image, in both binning modes
    (and with the two-level scatter of large images forced, whose super-tile counters live in the same state)."""
    monkeypatch.setenv("GGD_SCATTER_SUPER", super_scatter)
    lib = _lib.load()
    P, size = 20_000, 160
    s = to_dev(synthetic.raster_scene(P, size=size, seed_offset=4), cuda)
    p = _lib.ptr
    for mode in (1, 0):
        prev = lib.ggd_raster_binning(mode)
        try:
            first = run_ours(s, cuda)
            geom, binning, img = first["_buffers"]
            again = torch.zeros_like(first["out_color"])
            _lib.check(lib.ggd_raster_render(None, P, size, size, p(s["bg"]), None, p(first["radii"]), p(geom), p(binning),
                                             first["capacity"], p(img), p(again), mode))
            torch.cuda.synchronize()
        finally:
            lib.ggd_raster_binning(prev)
        assert torch.equal(again, first["out_color"])


def _split_frame(lib, s, cuda, pre_mode, render_mode):
    """Stage 1 and stage 2 through the C ABI with separately chosen binning modes."""
    from gaussian_gan_decoder_b200 import rasterizer as rz
    P, W, H = s["means3D"].shape[0], s["W"], s["H"]
    p = _lib.ptr
    cap = 16 * P
    gb, bb, ib = rz.workspace_sizes(P, W, H, cap)
    u8 = dict(dtype=torch.uint8, device=cuda)
    geom, binning, img = torch.empty(gb, **u8), torch.empty(bb, **u8), torch.empty(ib, **u8)
    radii = torch.zeros(P, dtype=torch.int32, device=cuda)
    image = torch.zeros(3, H, W, device=cuda)
    _lib.check(lib.ggd_raster_preprocess(None, P, 0, 1, W, H, p(s["means3D"]), p(s["shs"]), None, p(s["opacities"]),
                                         p(s["scales"]), 1.0, p(s["rotations"]), None, p(s["viewmatrix"]),
                                         p(s["projmatrix"]), p(s["campos"]), s["tanfovx"], s["tanfovy"], 0,
                                         p(geom), p(radii), None, pre_mode))
    _lib.check(lib.ggd_raster_render(None, P, W, H, p(s["bg"]), None, p(radii), p(geom), p(binning), cap, p(img),
                                     p(image), render_mode))
    st = (C.c_uint32 * 2)()
    _lib.check(lib.ggd_raster_status(None, p(geom), P, st))
    info = (C.c_uint32 * 2)()
    _lib.check(lib.ggd_raster_binning_info(None, p(geom), P, info))
    return image, st[0], st[1], info[0]


def test_binning_mode_mismatch_between_stages_is_safe(cuda):
    """The mode is an argument of each stage.  A geometry state prepared WITHOUT per-tile counts (mode 0)
    and rendered in mode 1 must take the radix sort (same image), not read the zeroed counts as an empty
    frame; rendered in mode 2 it must raise status flag 2.  (this used to render background only.)"""
    lib = _lib.load()
    s = to_dev(synthetic.raster_scene(30_000, size=256, seed_offset=8), cuda)
    ref, R, flags, per_tile = _split_frame(lib, s, cuda, 1, 1)
    assert R > 0 and flags == 0 and per_tile == 1
    assert float((ref - s["bg"].view(3, 1, 1)).abs().max()) > 0.05     # something was drawn
    img, R2, flags, per_tile = _split_frame(lib, s, cuda, 0, 1)
    assert R2 == R and flags == 0 and per_tile == 0
    assert torch.equal(img, ref)
    img, R2, flags, per_tile = _split_frame(lib, s, cuda, 0, 0)
    assert flags == 0 and torch.equal(img, ref)
    _, _, flags, per_tile = _split_frame(lib, s, cuda, 0, 2)
    assert flags & 2 and per_tile == 0
    img, _, flags, per_tile = _split_frame(lib, s, cuda, 2, 2)
    assert flags == 0 and per_tile == 1 and torch.equal(img, ref)
    # GGD_BINNING_LEAN: per-tile binning without the reference-order debug arrays — same image
    for m in (1 | _lib.BINNING_LEAN, 2 | _lib.BINNING_LEAN, 0 | _lib.BINNING_LEAN):
        img, _, flags, _ = _split_frame(lib, s, cuda, m, m)
        assert flags == 0 and torch.equal(img, ref), m
    with pytest.raises(_lib.GGDError):
        _split_frame(lib, s, cuda, 3, 1)


def test_clustered_depths_beyond_the_chunked_sort_fall_back_to_radix(cuda):
    """Tile lists longer than the 24576-word shared-memory buffer are sorted per tile in depth chunks; when one
    1/128 slice of a list's depth interval alone overflows the buffer (here: every splat at nearly one depth, 16 tiles
    of ~40k instances) the kernel raises status flag 8 and the binding renders the frame again with the radix sort.
    The result must be the radix path's bits."""
    from gaussian_gan_decoder_b200 import rasterizer as rz
    lib = _lib.load()
    P, size = 700_000, 64
    s = synthetic.raster_scene(P, size=size, seed_offset=13)
    V = s["viewmatrix"]
    d = V[:3, 2] / V[:3, 2].norm()
    s["means3D"] = (s["means3D"] - (s["means3D"] @ d)[:, None] * d[None, :]).contiguous()      # one plane facing the camera
    s = to_dev(s, cuda)
    rz._binning_hint.pop((cuda.index, P, size, size), None)
    prev = lib.ggd_raster_binning(1)
    try:
        auto = run_ours(s, cuda)
        st = (C.c_uint32 * 2)()
        _lib.check(lib.ggd_raster_status(None, C.c_void_p(auto["_buffers"][0].data_ptr()), P, st))
        lib.ggd_raster_binning(0)
        radix = run_ours(s, cuda)
    finally:
        lib.ggd_raster_binning(prev)
    counts = auto["ranges"].view(-1, 2)[:, 1] - auto["ranges"].view(-1, 2)[:, 0]
    assert int(counts.max()) > 24576
    assert rz._binning_hint[(cuda.index, P, size, size)][0] == 0       # the shape is remembered as radix-only
    assert_bit_exact(auto, radix, P)
