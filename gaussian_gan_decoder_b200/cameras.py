"""Camera glue on the path: EG3D cam2world -> 3DGS view / projection matrices.

Mirrors ``CustomCam`` (gaussian_splatting/scene/cameras.py:75-92) and
``getProjectionMatrix`` (gaussian_splatting/utils/graphics_utils.py:52-75) of the
reference, minus their hard-coded ``.cuda()`` calls: matrices are
built on the device of ``extr`` so the same fp32 bits can be fed to the
reference and to this implementation.
"""
import math

import torch


def getProjectionMatrix(znear, zfar, fovX, fovY):
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    z_sign = 1.0
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = z_sign
    P[2, 2] = z_sign * (zfar + znear) / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


class CustomCam:
    """``extr`` is the EG3D cam2world matrix; ``fov`` is in radians (both axes)."""

    def __init__(self, size, fov, extr, znear=0.01, zfar=10.0):
        self.image_width = size
        self.image_height = size
        self.FoVy = fov
        self.FoVx = fov
        self.znear = znear
        self.zfar = zfar
        # invert and transpose: EG3D column-vector c2w -> 3DGS row-vector w2c
        self.world_view_transform = extr.T.inverse()
        self.projection_matrix = getProjectionMatrix(
            znear=self.znear, zfar=self.zfar, fovX=self.FoVx, fovY=self.FoVy).transpose(0, 1).to(extr.device)
        self.full_proj_transform = (self.world_view_transform.unsqueeze(0).bmm(
            self.projection_matrix.unsqueeze(0))).squeeze(0)
        # NB: this is the w2c translation, not the camera position (reference quirk q6);
        # it only matters for SH degree > 0, which the decoder path never uses.
        self.camera_center = self.world_view_transform[3, :3]


def look_at_cam2world(h, v, radius=2.7):
    """cam2world of a camera on a sphere looking at the origin, y up, no roll.
    Same geometry as the reference's LookAtPoseSampler / create_cam2world_matrix
    (main/camera_utils.py:63-95, 131-150) for zero stddev."""
    v = min(max(v, 1e-5), math.pi - 1e-5)
    theta = h
    phi = math.acos(1 - 2 * (v / math.pi))
    o = torch.tensor([radius * math.sin(phi) * math.cos(math.pi - theta),
                      radius * math.cos(phi),
                      radius * math.sin(phi) * math.sin(math.pi - theta)], dtype=torch.float32)
    fwd = -o / o.norm()
    up = torch.tensor([0.0, 1.0, 0.0])
    right = -torch.linalg.cross(up, fwd)
    right = right / right.norm()
    up = torch.linalg.cross(fwd, right)
    up = up / up.norm()
    rot = torch.eye(4)
    rot[:3, :3] = torch.stack((right, up, fwd), dim=-1)
    trans = torch.eye(4)
    trans[:3, 3] = o
    return trans @ rot
