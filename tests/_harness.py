"""Test-side wrappers around the two checkers under oracle/:

* ``CpuOracle``  — oracle/libggd_oracle.so, our CPU restatement of the reference rasteriser;
* ``RefCuda``    — oracle/_ref/libggd_ref_cuda.so, the UNMODIFIED reference CUDA rasteriser
                   (compiled in the build container from /root/reference; runs only on a GPU).

Only tests, smoke() and bench.py's baseline legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
CPU_SO = os.path.join(ORACLE_DIR, "libggd_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libggd_ref_cuda.so")

_p, _f, _i = C.c_void_p, C.c_float, C.c_int


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(x):
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(x, dtype=np.float32)


import sys
sys.path.insert(0, ORACLE_DIR)
from raster_oracle import CpuOracle  # noqa: E402,F401


def ref_cuda_available():
    return os.path.exists(REF_SO) and torch.cuda.is_available()


class RefCuda:
    """The reference's CudaRasterizer::Rasterizer through oracle/ref_shim.cu."""

    def __init__(self):
        self.lib = C.CDLL(REF_SO)
        self.lib.ggd_ref_forward.restype = _i
        self.lib.ggd_ref_forward.argtypes = [_i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p,
                                             _f, _f, _i, _p, _p, _i, C.POINTER(_p), C.POINTER(_p), C.POINTER(_p)]
        self.lib.ggd_ref_backward.restype = _i
        self.lib.ggd_ref_backward.argtypes = [_i, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p,
                                              _f, _f, _p] + [_p] * 10 + [_i]
        for n in ("ggd_ref_geom_offsets", "ggd_ref_binning_offsets", "ggd_ref_img_offsets"):
            getattr(self.lib, n).argtypes = [C.c_size_t, C.POINTER(C.c_size_t)]
        self.lib.ggd_ref_read.restype = _i
        self.lib.ggd_ref_read.argtypes = [_i, C.c_size_t, _p, C.c_size_t]
        self.lib.ggd_ref_mark_visible.restype = None
        self.lib.ggd_ref_mark_visible.argtypes = [_i, _p, _p, _p, _p]

    def mark_visible(self, means3D, viewmatrix, projmatrix):
        """Reference CudaRasterizer::Rasterizer::markVisible (rasterizer_impl.cu:141-152) -> bool[P] (device)."""
        P = means3D.shape[0]
        present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
        m, v, pr = means3D.contiguous().float(), viewmatrix.contiguous().float(), projmatrix.contiguous().float()
        torch.cuda.synchronize()
        self.lib.ggd_ref_mark_visible(P, self._dp(m), self._dp(v), self._dp(pr), C.c_void_p(present.data_ptr()))
        torch.cuda.synchronize()
        return present

    @staticmethod
    def _dp(t):
        return None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())

    def forward(self, scene, dev):
        """scene tensors must already live on ``dev``. Returns dict of torch tensors (device)."""
        t = {k: (scene.get(k).contiguous().float() if scene.get(k) is not None else None)
             for k in ("bg", "means3D", "shs", "colors_precomp", "opacities", "scales", "rotations",
                       "cov3D_precomp", "viewmatrix", "projmatrix", "campos")}
        P = t["means3D"].shape[0]
        W, H = scene["W"], scene["H"]
        M = t["shs"].shape[1] if t["shs"] is not None else 0
        out_color = torch.zeros(3, H, W, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        g, b, im = _p(), _p(), _p()
        torch.cuda.synchronize()
        d = self._dp
        R = self.lib.ggd_ref_forward(P, int(scene.get("sh_degree", 0)), M, d(t["bg"]), W, H, d(t["means3D"]),
                                     d(t["shs"]), d(t["colors_precomp"]), d(t["opacities"]), d(t["scales"]),
                                     float(scene.get("scale_modifier", 1.0)), d(t["rotations"]),
                                     d(t["cov3D_precomp"]), d(t["viewmatrix"]), d(t["projmatrix"]),
                                     d(t["campos"]), float(scene["tanfovx"]), float(scene["tanfovy"]), 0,
                                     d(out_color), d(radii), 0, C.byref(g), C.byref(b), C.byref(im))
        torch.cuda.synchronize()
        assert R >= 0, "reference forward failed"
        self._last = dict(P=P, W=W, H=H, R=R, M=M, t=t, scene=scene, radii=radii)
        res = dict(R=R, out_color=out_color, radii=radii)

        def grab(which, off, count, dtype):
            n = count * torch.empty((), dtype=dtype).element_size()
            raw = torch.empty(n, dtype=torch.uint8, device=dev)
            if n:  # device-to-device copy out of the shim-owned buffer
                assert self.lib.ggd_ref_read(which, off, C.c_void_p(raw.data_ptr()), n) == 0
            return raw.view(dtype)

        go = (C.c_size_t * 10)(); self.lib.ggd_ref_geom_offsets(P, go)
        names = [("depths", 0, P, torch.float32), ("clamped", 1, 3 * P, torch.uint8), ("means2D", 3, 2 * P, torch.float32),
                 ("cov3D", 4, 6 * P, torch.float32), ("conic_opacity", 5, 4 * P, torch.float32),
                 ("rgb", 6, 3 * P, torch.float32), ("tiles_touched", 7, P, torch.int32),
                 ("point_offsets", 9, P, torch.int32)]
        for n, k, c, dt in names:
            res[n] = grab(0, go[k], c, dt)
        if R > 0:
            bo = (C.c_size_t * 4)(); self.lib.ggd_ref_binning_offsets(R, bo)
            res["point_list"] = grab(1, bo[0], R, torch.int32)
            res["point_list_unsorted"] = grab(1, bo[1], R, torch.int32)
            res["point_list_keys"] = grab(1, bo[2], R, torch.int64)
            res["point_list_keys_unsorted"] = grab(1, bo[3], R, torch.int64)
        io = (C.c_size_t * 3)(); self.lib.ggd_ref_img_offsets(W * H, io)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        res["accum_alpha"] = grab(2, io[0], W * H, torch.float32)
        res["n_contrib"] = grab(2, io[1], W * H, torch.int32)
        res["ranges"] = grab(2, io[2], 2 * tiles, torch.int32)
        torch.cuda.synchronize()
        return res

    def backward(self, dL_dpix, dev):
        L = self._last
        P, M, t, scene = L["P"], L["M"], L["t"], L["scene"]
        z = lambda *s: torch.zeros(*s, device=dev)
        out = dict(dL_dmean2D=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1), dL_dcolor=z(P, 3),
                   dL_dmean3D=z(P, 3), dL_dcov3D=z(P, 6), dL_dsh=z(P, max(M, 1), 3), dL_dscale=z(P, 3),
                   dL_drot=z(P, 4))
        d = self._dp
        g = dL_dpix.contiguous().float()
        torch.cuda.synchronize()
        rc = self.lib.ggd_ref_backward(P, int(scene.get("sh_degree", 0)), M, L["R"], d(t["bg"]), L["W"], L["H"],
                                       d(t["means3D"]), d(t["shs"]), d(t["colors_precomp"]), d(t["scales"]),
                                       float(scene.get("scale_modifier", 1.0)), d(t["rotations"]),
                                       d(t["cov3D_precomp"]), d(t["viewmatrix"]), d(t["projmatrix"]),
                                       d(t["campos"]), float(scene["tanfovx"]), float(scene["tanfovy"]),
                                       d(L["radii"]), d(g), d(out["dL_dmean2D"]), d(out["dL_dconic"]),
                                       d(out["dL_dopacity"]), d(out["dL_dcolor"]), d(out["dL_dmean3D"]),
                                       d(out["dL_dcov3D"]), d(out["dL_dsh"]), d(out["dL_dscale"]),
                                       d(out["dL_drot"]), 0)
        torch.cuda.synchronize()
        assert rc == 0
        return out


def run_ours(scene, dev):
    """Run the product rasteriser through the reference-facing binding and return every
    intermediate the contract names, as torch tensors on ``dev``."""
    from gaussian_gan_decoder_b200 import rasterizer as rz
    e = torch.empty(0, device=dev)
    g = lambda k: scene[k] if scene.get(k) is not None else e
    P, W, H = scene["means3D"].shape[0], scene["W"], scene["H"]
    rz.keep_reference_order_arrays = True      # the parity tests compare duplicateWithKeys' output as well
    R, color, radii, geom, binning, img = rz._C.rasterize_gaussians(
        g("bg"), g("means3D"), g("colors_precomp"), g("opacities"), g("scales"), g("rotations"),
        scene.get("scale_modifier", 1.0), g("cov3D_precomp"), g("viewmatrix"), g("projmatrix"),
        scene["tanfovx"], scene["tanfovy"], H, W, g("shs"), scene.get("sh_degree", 0), g("campos"), False, False)
    cap = getattr(binning, "ggd_capacity", 0)
    res = dict(R=R, out_color=color, radii=radii, _buffers=(geom, binning, img), capacity=cap)
    if P == 0:
        return res
    f = lambda which, name, dt: rz.field((geom, binning, img)[which], which, name, P, W, H, cap, dt)
    for n, dt in (("depths", torch.float32), ("means2D", torch.float32), ("cov3D", torch.float32),
                  ("conic_opacity", torch.float32), ("rgb", torch.float32), ("tiles_touched", torch.int32),
                  ("point_offsets", torch.int32), ("clamped", torch.uint8)):
        res[n] = f(0, n, dt)
    for n, dt in (("point_list", torch.int32), ("point_list_unsorted", torch.int32),
                  ("point_list_keys", torch.int64), ("point_list_keys_unsorted", torch.int64)):
        res[n] = f(1, n, dt)[:R]
    for n, dt in (("accum_alpha", torch.float32), ("n_contrib", torch.int32), ("ranges", torch.int32)):
        res[n] = f(2, n, dt)
    return res
