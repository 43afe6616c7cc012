"""Scene / render glue on the path: the attribute container with the reference's
activations and ``render_simple``.

Mirrors ``GaussianModel``'s getters (gaussian_splatting/scene/gaussian_model.py:35-45,
99-121; densification, PLY I/O and the simple_knn import are out of scope) and
``render_simple`` (gaussian_splatting/gaussian_renderer/__init__.py:105-186), including its
placeholder ``alpha`` / ``depth`` keys (reference quirk q7).
"""
import math

import torch

from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer


class GaussianModel:
    def __init__(self, sh_degree: int = 0):
        self.active_sh_degree = 0
        self.max_sh_degree = sh_degree
        self._xyz = torch.empty(0)
        self._features_dc = torch.empty(0)
        self._features_rest = torch.empty(0)
        self._scaling = torch.empty(0)
        self._rotation = torch.empty(0)
        self._opacity = torch.empty(0)
        self._activated = None      # optional (scales, rotations, opacities) from the fused decoder

    @classmethod
    def from_decoder(cls, result, activated=None):
        """The assignment block of the training / demo loops
        (main/train_pano2gaussian_decoder.py:223-227, main/load_decoder.py:17-21)."""
        g = cls(0)
        g._xyz = result.xyz
        g._scaling = result.scale
        g._rotation = result.rotation
        g._opacity = result.opacity
        g._features_dc = result.color.unsqueeze(1)
        if activated is not None:
            g._activated = (activated["act_scale"], activated["act_rotation"], activated["act_opacity"])
        return g

    # ---- the reference's .ply wire format (gaussian_splatting/scene/gaussian_model.py:259-303) ----
    def construct_list_of_attributes(self):
        l = ['x', 'y', 'z', 'nx', 'ny', 'nz']
        l += [f'f_dc_{i}' for i in range(self._features_dc.shape[1] * self._features_dc.shape[2])]
        if self.active_sh_degree > 0:
            l += [f'f_rest_{i}' for i in range(self._features_rest.shape[1] * self._features_rest.shape[2])]
        l.append('opacity')
        l += [f'scale_{i}' for i in range(self._scaling.shape[1])]
        l += [f'rot_{i}' for i in range(self._rotation.shape[1])]
        return l

    def save_ply(self, path):
        """Binary little-endian PLY with the reference's property names and order (what the external
        viewers load).  Written directly; no plyfile dependency."""
        import os
        import numpy as np
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        cols = [self._xyz.detach().cpu().numpy(), np.zeros((self._xyz.shape[0], 3), np.float32),
                self._features_dc.detach().transpose(1, 2).flatten(start_dim=1).contiguous().cpu().numpy()]
        if self.active_sh_degree > 0:
            cols.append(self._features_rest.detach().transpose(1, 2).flatten(start_dim=1).contiguous().cpu().numpy())
        cols += [self._opacity.detach().cpu().numpy(), self._scaling.detach().cpu().numpy(),
                 self._rotation.detach().cpu().numpy()]
        data = np.ascontiguousarray(np.concatenate(cols, axis=1), dtype='<f4')
        names = self.construct_list_of_attributes()
        assert data.shape[1] == len(names)
        header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % data.shape[0]
        header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
        with open(path, "wb") as f:
            f.write(header.encode("ascii"))
            f.write(data.tobytes())

    def load_ply(self, path, device="cpu"):
        import numpy as np
        with open(path, "rb") as f:
            names, n = [], 0
            while True:
                line = f.readline().decode("ascii").strip()
                if line.startswith("element vertex"):
                    n = int(line.split()[-1])
                elif line.startswith("property float"):
                    names.append(line.split()[-1])
                elif line == "end_header":
                    break
            data = np.frombuffer(f.read(n * len(names) * 4), dtype='<f4').reshape(n, len(names))
        col = lambda prefix: torch.from_numpy(np.stack([data[:, names.index(k)] for k in names if k.startswith(prefix)], 1).copy())
        self._xyz = torch.from_numpy(data[:, :3].copy()).to(device)
        self._features_dc = col("f_dc_").view(n, 3, -1).transpose(1, 2).contiguous().to(device)
        self._opacity = col("opacity").to(device)
        self._scaling = col("scale_").to(device)
        self._rotation = col("rot_").to(device)
        self._activated = None
        return self

    @property
    def get_scaling(self):
        return self._activated[0] if self._activated is not None else torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return self._activated[1] if self._activated is not None else torch.nn.functional.normalize(self._rotation)

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_features(self):
        if self.active_sh_degree == 0:
            return self._features_dc
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_opacity(self):
        return self._activated[2] if self._activated is not None else torch.sigmoid(self._opacity)


def render_simple(viewpoint_camera, pc, bg_color, xyz_offset=None, scaling_modifier=1.0, override_color=None,
                  debug=False):
    """Render the scene.  Background tensor (bg_color) must be on the GPU."""
    xyz = pc.get_xyz
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height),
        image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5),
        tanfovy=math.tan(viewpoint_camera.FoVy * 0.5),
        bg=bg_color,
        scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center,
        prefiltered=False,
        debug=debug,
    )
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D = xyz if xyz_offset is None else xyz + xyz_offset
    shs, colors_precomp = (pc.get_features, None) if override_color is None else (None, override_color)
    rendered_image, radii = rasterizer(means3D=means3D, means2D=screenspace_points, shs=shs,
                                       colors_precomp=colors_precomp, opacities=pc.get_opacity,
                                       scales=pc.get_scaling, rotations=pc.get_rotation, cov3D_precomp=None)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "alpha": radii, "depth": radii}
