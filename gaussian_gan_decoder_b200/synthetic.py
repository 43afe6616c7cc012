"""Seeded synthetic inputs of the shapes BASELINE.json names.

Everything is generated on the CPU with a ``torch.Generator`` so that the same
seed gives the same bits on every machine, then moved to the target device.
There is no network for checkpoints or datasets; benchmarks say ``"data":
"synthetic"``.
"""
import math

import torch

from .cameras import CustomCam, look_at_cam2world

SEED = 303  # the reference's default seed (main/train_pano2gaussian_decoder.py:31)


def gen(seed_offset=0):
    g = torch.Generator(device="cpu")
    g.manual_seed(SEED + seed_offset)
    return g


def head_shell_points(P, g):
    """P points on a head-like shell inside the box_warp=1 cube: unit direction x 0.3 x
    clip(1 + 0.1 randn, 0, 1) (mimics surface sampling + radial jitter,
    main/decoder_utils/target_dataloader.py:104-118)."""
    d = torch.randn(P, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    r = 0.3 * torch.clip(1 + 0.1 * torch.randn(P, 1, generator=g), 0, 1)
    return (d * r).contiguous()


def trained_like_attributes(P, g):
    """Raw (pre-activation) decoder outputs with a trained-like footprint: log-scale
    ~ N(-5.5, 0.7), rotation randn, opacity logit ~ N(2, 2), colour ~ U(-1, 1)."""
    return dict(
        scale=(-5.5 + 0.7 * torch.randn(P, 3, generator=g)),
        rotation=torch.randn(P, 4, generator=g),
        opacity=(2 + 2 * torch.randn(P, 1, generator=g)),
        color=(torch.rand(P, 3, generator=g) * 2 - 1),
    )


def activated(attrs):
    """GaussianModel getters (gaussian_splatting/scene/gaussian_model.py:99-121), as seeded INPUT data: evaluated in float64
    and rounded once.  torch's vectorised fp32 exp / sigmoid on the CPU may differ in the last bit between its vector body
    and scalar tail, i.e. with the alignment of the tensor, i.e. from one process to the next — and one splat whose radius
    changes by a pixel changes the instance count of a "fixed" benchmark scene (seen: R = 6 438 932 vs 6 438 940)."""
    return dict(
        scales=torch.exp(attrs["scale"].double()).float(),
        rotations=torch.nn.functional.normalize(attrs["rotation"].double()).float(),
        opacities=torch.sigmoid(attrs["opacity"].double()).float(),
        shs=attrs["color"][:, None, :].contiguous(),
    )


def camera(size=512, fov_deg=12.0, h=math.pi / 2 + 0.3, v=math.pi / 2 - 0.2, radius=2.7):
    return CustomCam(size=size, fov=fov_deg * math.pi / 180.0, extr=look_at_cam2world(h, v, radius))


def planes(arch, g, res=256, channels=32, depth=3):
    """Random tri-planes [1,3,32,res,res] (eg3d) or tri-grids [1,3,32*depth,res,res] (panohead)."""
    c = channels * (depth if arch == "panohead" else 1)
    return torch.randn(1, 3, c, res, res, generator=g)


BG = {"panohead": [0.55717, 0.52256, 0.51045], "eg3d": [0.0, 0.0, 0.0]}


def raster_scene(P, size=512, fov_deg=12.0, seed_offset=0, device="cpu", arch="panohead"):
    """A complete rasteriser input set: activated attributes + camera matrices."""
    g = gen(seed_offset)
    xyz = head_shell_points(P, g)
    act = activated(trained_like_attributes(P, g))
    cam = camera(size, fov_deg)
    scene = dict(
        means3D=xyz, **act,
        viewmatrix=cam.world_view_transform.contiguous(),
        projmatrix=cam.full_proj_transform.contiguous(),
        campos=cam.camera_center.contiguous(),
        bg=torch.tensor(BG[arch], dtype=torch.float32),
    )
    scene = {k: v.to(device) for k, v in scene.items()}
    scene.update(W=size, H=size, tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5),
                 sh_degree=0, scale_modifier=1.0)
    return scene


class _Renderer:
    plane_axes = None      # the fused kernel hard-wires the per-arch axis permutation


class SyntheticGenerator(torch.nn.Module):
    """Stand-in for the upstream EG3D / PanoHead generator (out of scope):
    ``mapping`` returns zeros, ``synthesis`` returns the given ``feature_planes``."""

    def __init__(self, planes, arch="eg3d", depth=1, box_warp=1.0):
        super().__init__()
        self.planes = planes
        self.rendering_kwargs = {"box_warp": box_warp}
        if arch == "panohead":
            self.rendering_kwargs["triplane_depth"] = depth
        self.renderer = _Renderer()
        self.calls = {}          # last camera tensor each entry point received (no host sync here)
        self._ws = None

    def mapping(self, z, c, truncation_psi=1):
        self.calls["mapping"] = c
        if self._ws is None or self._ws.device != z.device:
            self._ws = torch.zeros(1, 14, 512, device=z.device)
        return self._ws

    def synthesis(self, ws, c, noise_mode="const"):
        self.calls["synthesis"] = c
        return {"feature_planes": self.planes}
