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
    """OpenGL-style perspective matrix (column-vector convention, z_sign = +1, symmetric frustum) with the
    reference's ``(f + n) / (f - n)`` depth term.  The entries are evaluated in Python doubles in the same order
    as the reference does (near-plane half extents first) and rounded to fp32 once, so the matrix is bit-identical
    to the reference's — pinned by tests/test_cameras.py against tests/golden/cameras.npz."""
    half_w = math.tan(fovX / 2) * znear          # right = -left
    half_h = math.tan(fovY / 2) * znear          # top = -bottom
    depth = zfar - znear
    return torch.tensor([[2.0 * znear / (half_w + half_w), 0.0, 0.0, 0.0],
                         [0.0, 2.0 * znear / (half_h + half_h), 0.0, 0.0],
                         [0.0, 0.0, (zfar + znear) / depth, -(zfar * znear) / depth],
                         [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32)


class CustomCam:
    """3DGS camera from an EG3D pose.  ``extr`` is the EG3D cam2world matrix (column vectors); ``fov`` is in
    radians and applies to both axes.  Attribute names are the ones ``render_simple`` reads."""

    def __init__(self, size, fov, extr, znear=0.01, zfar=10.0):
        self.image_width = self.image_height = size
        self.FoVx = self.FoVy = fov
        self.znear, self.zfar = znear, zfar
        # 3DGS kernels index matrices as m[4*col + row] of a row-vector world->camera transform:
        # that is the inverse of the transposed cam2world
        view = extr.T.inverse()
        proj = getProjectionMatrix(znear, zfar, fov, fov).T.to(extr.device)
        self.world_view_transform = view
        self.projection_matrix = proj
        self.full_proj_transform = torch.bmm(view[None], proj[None])[0]
        # NB: row 3 of the view matrix is the w2c translation, not the camera position (reference quirk q6);
        # it only matters for SH degree > 0, which the decoder path never uses.
        self.camera_center = view[3, :3]


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
