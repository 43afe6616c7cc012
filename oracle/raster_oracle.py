"""TEST INFRASTRUCTURE — ctypes wrapper of oracle/libggd_oracle.so (raster_oracle.cpp), the
CPU restatement of the reference rasteriser.  NOT the product: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
import ctypes as C
import os
import subprocess

import numpy as np
import torch

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
CPU_SO = os.path.join(ORACLE_DIR, "libggd_oracle.so")
_p, _f, _i = C.c_void_p, C.c_float, C.c_int


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(x):
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(x, dtype=np.float32)


class CpuOracle:
    FIELDS = {
        "depths": np.float32, "means2D": np.float32, "cov3D": np.float32, "conic_opacity": np.float32,
        "rgb": np.float32, "accum_alpha": np.float32, "out_color": np.float32, "radii": np.int32,
        "tiles_touched": np.uint32, "point_offsets": np.uint32, "point_list_unsorted": np.uint32,
        "point_list": np.uint32, "n_contrib": np.uint32, "ranges": np.uint32, "clamped": np.uint8,
        "point_list_keys_unsorted": np.uint64, "point_list_keys": np.uint64,
        "dL_dmean2D": np.float32, "dL_dconic": np.float32, "dL_dopacity": np.float32, "dL_dcolor": np.float32,
        "dL_dmean3D": np.float32, "dL_dcov3D": np.float32, "dL_dsh": np.float32, "dL_dscale": np.float32,
        "dL_drot": np.float32,
    }

    def __init__(self):
        if not os.path.exists(CPU_SO):
            subprocess.check_call(["make", "-f", os.path.join(ORACLE_DIR, "Makefile"), "oracle"])
        self.lib = C.CDLL(CPU_SO)
        self.lib.oracle_create.restype = _p
        self.lib.oracle_destroy.argtypes = [_p]
        self.lib.oracle_forward.restype = _i
        self.lib.oracle_forward.argtypes = [_p, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p,
                                            _p, _f, _f, _i]
        self.lib.oracle_backward.restype = _i
        self.lib.oracle_backward.argtypes = [_p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p]
        self.lib.oracle_get.restype = _i
        self.lib.oracle_get.argtypes = [_p, C.c_char_p, C.POINTER(_p), C.POINTER(C.c_size_t), C.POINTER(_i)]
        self.lib.oracle_higher_msb.restype = C.c_uint32
        self.lib.oracle_higher_msb.argtypes = [C.c_uint32]
        self.lib.oracle_num_threads.restype = _i
        self.h = C.c_void_p(self.lib.oracle_create())
        self._keep = None

    def __del__(self):
        try:
            self.lib.oracle_destroy(self.h)
        except Exception:
            pass

    def num_threads(self):
        return int(self.lib.oracle_num_threads())

    def forward(self, scene, preprocess_only=False):
        a = {k: _f32(scene.get(k)) for k in ("bg", "means3D", "shs", "colors_precomp", "opacities", "scales",
                                              "rotations", "cov3D_precomp", "viewmatrix", "projmatrix", "campos")}
        self._keep = (a, scene)
        P = a["means3D"].shape[0]
        M = a["shs"].shape[1] if a["shs"] is not None else 0
        R = self.lib.oracle_forward(self.h, P, int(scene.get("sh_degree", 0)), M, _np_ptr(a["bg"]), scene["W"],
                                    scene["H"], _np_ptr(a["means3D"]), _np_ptr(a["shs"]),
                                    _np_ptr(a["colors_precomp"]), _np_ptr(a["opacities"]), _np_ptr(a["scales"]),
                                    float(scene.get("scale_modifier", 1.0)), _np_ptr(a["rotations"]),
                                    _np_ptr(a["cov3D_precomp"]), _np_ptr(a["viewmatrix"]),
                                    _np_ptr(a["projmatrix"]), _np_ptr(a["campos"]), float(scene["tanfovx"]),
                                    float(scene["tanfovy"]), int(preprocess_only))
        return R

    def backward(self, dL_dpix):
        a, scene = self._keep
        g = _f32(dL_dpix)
        self.lib.oracle_backward(self.h, _np_ptr(a["bg"]), _np_ptr(a["means3D"]), _np_ptr(a["shs"]),
                                 _np_ptr(a["colors_precomp"]), _np_ptr(a["scales"]),
                                 float(scene.get("scale_modifier", 1.0)), _np_ptr(a["rotations"]),
                                 _np_ptr(a["cov3D_precomp"]), _np_ptr(a["viewmatrix"]), _np_ptr(a["projmatrix"]),
                                 _np_ptr(a["campos"]), float(scene["tanfovx"]), float(scene["tanfovy"]), _np_ptr(g))

    def get(self, name):
        ptr, cnt, el = _p(), C.c_size_t(), _i()
        if self.lib.oracle_get(self.h, name.encode(), C.byref(ptr), C.byref(cnt), C.byref(el)) != 0:
            raise KeyError(name)
        dt = np.dtype(self.FIELDS[name])
        assert dt.itemsize == el.value
        if cnt.value == 0:
            return np.zeros(0, dtype=dt)
        buf = (C.c_char * (cnt.value * el.value)).from_address(ptr.value)
        return np.frombuffer(buf, dtype=dt).copy()
