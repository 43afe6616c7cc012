"""CustomCam / getProjectionMatrix against matrices produced by the reference's own code
(tests/golden/cameras.npz, made by tests/golden/make_camera_golden.py from scene/cameras.py:75-92,
utils/graphics_utils.py:52-75 and main/camera_utils.py:63-150).  Runs without a GPU."""
import math
import os

import numpy as np
import torch

from gaussian_gan_decoder_b200.cameras import CustomCam, getProjectionMatrix, look_at_cam2world

Z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cameras.npz"))


def _bits(t):
    return torch.as_tensor(np.ascontiguousarray(t)).contiguous().view(torch.int32)


def test_projection_matrix_bit_identical_to_reference():
    for i, (size, fov_deg, h, v, radius) in enumerate(Z["cases"]):
        fov = float(fov_deg) * math.pi / 180.0
        P = getProjectionMatrix(znear=0.01, zfar=10.0, fovX=fov, fovY=fov)
        assert P.dtype == torch.float32 and torch.equal(_bits(P), _bits(Z[f"P_{i}"])), i


def test_custom_cam_matches_reference():
    for i, (size, fov_deg, h, v, radius) in enumerate(Z["cases"]):
        fov = float(fov_deg) * math.pi / 180.0
        cam = CustomCam(size=int(size), fov=fov, extr=torch.from_numpy(Z[f"c2w_{i}"]))
        assert cam.image_width == cam.image_height == int(size) and cam.FoVx == cam.FoVy == fov
        assert (cam.znear, cam.zfar) == (0.01, 10.0)
        assert torch.equal(_bits(cam.projection_matrix), _bits(Z[f"proj_{i}"])), i
        # the 4x4 inverse goes through LAPACK: identical on this image, allow 1-2 ulp for another CPU's code path
        for name, ref in (("world_view_transform", f"view_{i}"), ("full_proj_transform", f"full_{i}"),
                          ("camera_center", f"center_{i}")):
            got, want = getattr(cam, name), torch.from_numpy(Z[ref])
            assert got.shape == want.shape and got.dtype == torch.float32
            assert torch.allclose(got, want, rtol=3e-7, atol=3e-7), (i, name, (got - want).abs().max())
        # quirk q6 kept: camera_center is row 3 of the view matrix
        assert torch.equal(cam.camera_center, cam.world_view_transform[3, :3])


def test_look_at_pose_matches_reference_sampler():
    for i, (size, fov_deg, h, v, radius) in enumerate(Z["cases"]):
        c2w = look_at_cam2world(float(h), float(v), float(radius))
        assert torch.allclose(c2w, torch.from_numpy(Z[f"c2w_{i}"]), rtol=1e-6, atol=1e-6), i
