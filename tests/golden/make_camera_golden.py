"""Generates tests/golden/cameras.npz from the REFERENCE's own camera code (run in the build container,
where /root/reference exists; the fixture travels, the reference does not):

    main/camera_utils.py:63-95,131-150          LookAtPoseSampler.sample / create_cam2world_matrix
    gaussian_splatting/scene/cameras.py:75-92   CustomCam
    gaussian_splatting/utils/graphics_utils.py:52-75  getProjectionMatrix

CustomCam calls ``.cuda()`` on the projection matrix; there is no GPU here, so ``Tensor.cuda`` is patched to
the identity for the duration of the script (the arithmetic is CPU fp32 either way).

    python tests/golden/make_camera_golden.py
"""
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path[:0] = [os.path.join(REF, "gaussian_splatting"), os.path.join(REF, "main"), REF]
torch.Tensor.cuda = lambda self, *a, **k: self          # noqa: E731  (no GPU in the build container)
# `import scene` drags in the dataset readers; their third-party imports are absent here and unused by CustomCam
import types                                             # noqa: E402
for name, attrs in (("plyfile", ("PlyData", "PlyElement")), ("simple_knn", ()), ("simple_knn._C", ("distCUDA2",))):
    m = types.ModuleType(name)
    for a in attrs:
        setattr(m, a, None)
    sys.modules.setdefault(name, m)

from scene.cameras import CustomCam                      # noqa: E402  (reference)
from utils.graphics_utils import getProjectionMatrix     # noqa: E402  (reference)
from camera_utils import LookAtPoseSampler               # noqa: E402  (reference)

CASES = [  # size, fov_deg, h, v, radius
    (512, 12.0, math.pi / 2 + 0.3, math.pi / 2 - 0.2, 2.7),       # the bench camera
    (512, 5.0, math.pi / 2, math.pi / 2, 2.7),
    (2048, 17.0, math.pi / 2 - 0.9, math.pi / 2 + 0.4, 2.7),
    (128, 12.0, 0.1, 0.3, 1.0),
    (256, 18.837, math.pi, 2.9, 4.0),
]

out = {"cases": np.array(CASES, dtype=np.float64)}
for i, (size, fov_deg, h, v, radius) in enumerate(CASES):
    fov = fov_deg * math.pi / 180.0
    c2w = LookAtPoseSampler.sample(h, v, radius=radius)[0]
    cam = CustomCam(size=size, fov=fov, extr=c2w)
    out[f"c2w_{i}"] = c2w.numpy()
    out[f"view_{i}"] = cam.world_view_transform.numpy()
    out[f"proj_{i}"] = cam.projection_matrix.numpy()
    out[f"full_{i}"] = cam.full_proj_transform.numpy()
    out[f"center_{i}"] = cam.camera_center.numpy()
    out[f"P_{i}"] = getProjectionMatrix(znear=0.01, zfar=10.0, fovX=fov, fovY=fov).numpy()
np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cameras.npz"), **out)
print("wrote cameras.npz with", len(CASES), "cases")
