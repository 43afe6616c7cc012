"""Iso-surface point sets on the GPU.

Drop-in for the part of ``TargetDataloader.get_data`` between the density grid and ``result.vertices``
(main/decoder_utils/target_dataloader.py:96-118 and ``marching_cubes`` :185-190): the reference moves the
128^3 grid to the CPU, runs skimage marching cubes + trimesh, and draws one barycentric point per face until
500k points exist.  ``surface_points`` keeps the grid on the device and returns the jittered points directly::

    sigmas = ...                       # [res, res, res] float32 CUDA tensor (G.sample_mixed output, reshaped)
    result.vertices = surface_points(sigmas, level=10.0, n_points=500_000, thickness=0.1, seed=step)

Differences from the reference: tetrahedral tessellation of each cell instead of skimage's Lewiner tables, a
counter-based random stream instead of torch.rand / torch.randn (reproducible per ``seed``); the point ->
triangle order (k mod n_faces), the weights (uniforms / sum), ``v / res - 0.5`` and the radial scale are the
reference's.
"""
import ctypes as C

import torch

from .. import _lib


def surface_points(sigmas, level=10.0, n_points=500_000, thickness=0.1, seed=0, return_num_triangles=False):
    if not isinstance(sigmas, torch.Tensor) or not sigmas.is_cuda:
        raise RuntimeError("sigmas must be a CUDA tensor: the point-set producer has no CPU path")
    if sigmas.dim() != 3 or sigmas.shape[0] != sigmas.shape[1] or sigmas.shape[1] != sigmas.shape[2]:
        raise ValueError(f"sigmas must be [res, res, res], got {tuple(sigmas.shape)}")
    lib = _lib.load()
    sigmas = sigmas.to(torch.float32).contiguous()
    res, dev = int(sigmas.shape[0]), sigmas.device
    points = torch.empty(int(n_points), 3, dtype=torch.float32, device=dev)
    nbytes = lib.ggd_surface_points_scratch_bytes(res)
    scratch = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
    pinned = torch.zeros(1, dtype=torch.int32).pin_memory() if return_num_triangles else None
    with torch.cuda.device(dev):
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(lib.ggd_surface_points(s, _lib.ptr(sigmas), res, float(level), int(seed) & 0xFFFFFFFFFFFFFFFF, int(n_points),
                                          float(thickness), _lib.ptr(scratch), nbytes, _lib.ptr(points),
                                          C.c_void_p(pinned.data_ptr()) if pinned is not None else None))
        if return_num_triangles:
            torch.cuda.current_stream().synchronize()
            return points, int(pinned[0].item()) & 0xFFFFFFFF
    return points
