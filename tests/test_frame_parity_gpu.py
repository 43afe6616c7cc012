"""End-to-end parity of the FRAME at BASELINE configs[1] (500k Gaussians, PanoHead tri-grids, 512x512):

    ours      : decode_tc_kernel (fp16 operands) or decode_fwd_kernel (fp32)  ->  our rasteriser
    reference : fp32 decoder oracle (torch CPU ops = the ATen operators the reference lowers to,
                pinned to the reference classes by tests/golden/decoder_*.npz)
                ->  the UNMODIFIED reference CUDA rasteriser (oracle/_ref)

The rasteriser alone is bit-exact given equal inputs (test_raster_gpu.py); here the inputs are NOT equal —
they carry the decoder's rounding — and radii, tile rectangles and per-pixel contributor counts are integer
functions of them.  The contract (BASELINE.json north_star): rendered image PSNR >= 50 dB; decoded attributes
max-abs 1e-3 (fp32) / relative 1e-2 (reduced precision).  The integer disagreements caused by the attribute
rounding are counted and bounded, once with default-init weights (what the reference has at iteration 0) and
once with every weight matrix scaled x3 so hidden pre-activations reach |x| ~ 5 and the heads emit O(1)
values with trained-like splat footprints.  Measured numbers land in gpurun_out/frame_parity_*.json.
"""
import json
import math
import os
import sys

import numpy as np
import pytest
import torch

from _harness import RefCuda, ref_cuda_available, run_ours
from _stubs import StubG
from gaussian_gan_decoder_b200 import synthetic
from gaussian_gan_decoder_b200.decoder_models import SequentialDecoderReverse

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import decoder_oracle as do  # noqa: E402

pytestmark = pytest.mark.gpu
HEADS = ("xyz", "scale", "rotation", "opacity", "color")
P, SIZE, DEPTH = 500_000, 512, 3

# (weights, precision) -> bounds: image PSNR (dB, contract 50), attribute max-abs error (absolute: the fp32
# contract is 1e-3; relative to the attribute's largest magnitude: the reduced-precision contract is 1e-2),
# fraction of Gaussians whose radius / tile count differs, and the mean |difference| of the per-pixel contributor
# count relative to its mean (the count is a running index: one splat flipping across the 1/255 alpha threshold
# shifts it by one for the rest of the pixel's list, so "fraction of pixels that differ" says little).
# The fp32 rows show that the anchor path reproduces the reference up to summation order (a handful of ceil()
# flips in 500k radii); the fp16 rows are the production path.  Measured (B200, end of r2): PSNR 144.6 / 111.6 dB
# (default fp32 / fp16), 108.3 / 58.7 dB (scaled); radii differing 0 / 6e-6 and 8e-6 / 1.6e-2; contributor
# count off by 2e-8 / 1e-4 and 7e-7 / 2.8e-3 of its mean.  The scaled scene's PSNR is the noisy one: with weights x3 a
# few splats are larger than the image (scale up to e^11), and the fp16 error of THEIR opacity and colour tints every
# pixel — three GELU fits with the same per-activation accuracy gave 65.0, 58.7 and 57.5 dB while every attribute and
# index statistic moved the other way (the shipped fit has the smallest attribute errors and half the radius flips of
# the 65 dB one).  Its bound is therefore 55 dB, 5 above the contract.
BOUNDS = {
    ("default", "fp32"): dict(psnr=120.0, attr_abs=5e-6, attr_rel=1e-5, radii=2e-5, tiles=2e-5, n_contrib=1e-5),
    ("default", "fp16"): dict(psnr=100.0, attr_abs=3e-4, attr_rel=1e-2, radii=1e-4, tiles=1e-3, n_contrib=1e-3),
    ("scaled", "fp32"): dict(psnr=95.0, attr_abs=1e-4, attr_rel=1e-5, radii=1e-4, tiles=1e-4, n_contrib=1e-5),
    ("scaled", "fp16"): dict(psnr=55.0, attr_abs=None, attr_rel=1e-2, radii=5e-2, tiles=3e-2, n_contrib=1e-2),
}


def _decoder(planes, dev, weights):
    torch.manual_seed(303)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", DEPTH), hidden_dim=128, use_xyz_embedding=False,
                                   use_gen_finetune=False, device=dev)
    if weights == "scaled":      # stand-in for a trained decoder: hidden pre-activations up to |x| ~ 5
        with torch.no_grad():
            for h in HEADS:
                for lin in getattr(dec, f"{h}_decoder").linears():
                    lin.weight.mul_(3.0)
    return dec


def _scene(attrs, cam_host, bg, dev):
    cam = synthetic.camera(SIZE, 12.0)
    return dict(means3D=attrs["xyz"].to(dev), scales=attrs["act_scale"].to(dev), rotations=attrs["act_rotation"].to(dev),
                opacities=attrs["act_opacity"].to(dev), shs=attrs["color"][:, None, :].contiguous().to(dev),
                viewmatrix=cam_host["world_view_transform"].to(dev), projmatrix=cam_host["full_proj_transform"].to(dev),
                campos=cam_host["camera_center"].to(dev), bg=bg.to(dev), W=SIZE, H=SIZE,
                tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), sh_degree=0, scale_modifier=1.0)


@pytest.mark.skipif(not ref_cuda_available(), reason="oracle/_ref/libggd_ref_cuda.so not present")
@pytest.mark.parametrize("weights", ["default", "scaled"])
def test_frame_parity_config2(cuda, weights):
    g = synthetic.gen(2)
    planes = synthetic.planes("panohead", g, res=256, depth=DEPTH)
    xyz = synthetic.head_shell_points(P, g)
    cam = synthetic.camera(SIZE, 12.0)
    cam_host = {k: getattr(cam, k).contiguous().clone() for k in ("world_view_transform", "full_proj_transform", "camera_center")}
    bg = torch.tensor(synthetic.BG["panohead"])
    dec = _decoder(planes.to(cuda), cuda, weights)
    flat = torch.cat([p.detach().reshape(-1).cpu() for p in dec.get_params_custom()])

    # reference side: fp32 decode on the CPU, reference CUDA rasteriser
    with torch.no_grad():
        ref_attr = do.decode("panohead", "sequential_reverse", 0, flat, planes, xyz, 1.0, DEPTH)
    ref = RefCuda().forward(_scene(ref_attr, cam_host, bg, cuda), cuda)
    ref_img = ref["out_color"].clone()
    ref_radii, ref_tiles, ref_nc = ref["radii"].clone(), ref["tiles_touched"].clone(), ref["n_contrib"].clone()
    assert ref["R"] > 0
    act_max = {h: float(ref_attr[h].abs().max()) for h in HEADS}
    if weights == "scaled":
        assert max(act_max["rotation"], act_max["color"], act_max["opacity"]) > 2.0     # really O(1) outputs

    stats = {"weights": weights, "P": P, "image": SIZE, "R_ref": int(ref["R"]), "attr_max": act_max}
    for precision in ("fp32", "fp16"):
        dec.precision = precision
        with torch.no_grad():
            out = dec(torch.zeros(1, 512, device=cuda), torch.zeros(1, 25, device=cuda), xyz.to(cuda), 1.0)
            act = dec.activated()
        ours_attr = {h: out[h] for h in HEADS}
        ours_attr.update(act)
        ours = run_ours(_scene(ours_attr, cam_host, bg, cuda), cuda)
        mse = float(((ours["out_color"] - ref_img) ** 2).mean())
        psnr = 10 * math.log10(1.0 / max(mse, 1e-20))
        err_abs = {h: float((out[h].cpu() - ref_attr[h]).abs().max()) for h in HEADS}
        err_rel = {h: err_abs[h] / max(act_max[h], 1e-12) for h in HEADS}
        radii_bad = float((ours["radii"] != ref_radii).float().mean())
        tiles_bad = float((ours["tiles_touched"].view(-1) != ref_tiles.view(-1)).float().mean())
        nc_bad = float((ours["n_contrib"].view(-1) != ref_nc.view(-1)).float().mean())
        nc_abs = (ours["n_contrib"].view(-1).long() - ref_nc.view(-1).long()).abs().float()
        row = dict(psnr_db=round(psnr, 2), attr_max_abs_err=err_abs, attr_rel_err=err_rel, R=int(ours["R"]),
                   radii_differ=radii_bad, tiles_touched_differ=tiles_bad, n_contrib_differ=nc_bad,
                   n_contrib_mean_abs_diff=float(nc_abs.mean()), n_contrib_max_abs_diff=float(nc_abs.max()),
                   n_contrib_mean=float(ref_nc.float().mean()),
                   image_max_abs_err=float((ours["out_color"] - ref_img).abs().max()))
        if ours["R"] == ref["R"] and tiles_bad == 0.0:      # same emission pattern: keys are comparable one to one
            row["sorted_keys_differ"] = float((ours["point_list_keys"] != ref["point_list_keys"]).float().mean())
            row["point_list_differ"] = float((ours["point_list"] != ref["point_list"]).float().mean())
            row["ranges_differ"] = float((ours["ranges"].view(-1) != ref["ranges"].view(-1)).float().mean())
        stats[precision] = row
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"frame_parity_{weights}.json"), "w") as f:
        json.dump(stats, f, indent=1)
    print(json.dumps(stats))

    for precision in ("fp32", "fp16"):
        b, r = BOUNDS[(weights, precision)], stats[precision]
        assert r["psnr_db"] >= b["psnr"], (weights, precision, r)
        for h in HEADS:
            assert r["attr_rel_err"][h] <= b["attr_rel"], (weights, precision, h, r["attr_rel_err"])
            if b["attr_abs"] is not None:
                assert r["attr_max_abs_err"][h] <= b["attr_abs"], (weights, precision, h, r["attr_max_abs_err"])
        assert r["radii_differ"] <= b["radii"], (weights, precision, r)
        assert r["tiles_touched_differ"] <= b["tiles"], (weights, precision, r)
        assert r["n_contrib_mean_abs_diff"] <= b["n_contrib"] * r["n_contrib_mean"], (weights, precision, r)
