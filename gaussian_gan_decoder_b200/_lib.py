"""ctypes binding of libggd_b200.so (the C ABI declared in include/ggd.h).

There is no fallback: if the shared library is missing, importing the product
API raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C gaussian_gan_decoder_b200/csrc``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libggd_b200.so")

GGD_OK, GGD_E_INVALID, GGD_E_CUDA, GGD_E_CAPACITY = 0, 1, 2, 3
ABI_VERSION = 3          # ggd_abi_version() of the library these signatures describe
BINNING_RADIX, BINNING_AUTO, BINNING_TILES, BINNING_LEAN = 0, 1, 2, 4


class GGDError(RuntimeError):
    """A C-ABI call returned a non-zero status."""

    def __init__(self, code, msg):
        super().__init__(f"ggd error {code}: {msg}")
        self.code = code


_p = C.c_void_p
_f = C.c_float
_i = C.c_int
_u64 = C.c_uint64
_sz = C.c_size_t



class DecoderConfig(C.Structure):
    """ctypes mirror of ``ggd_decoder_config`` (include/ggd.h)."""
    _fields_ = [("arch", C.c_int), ("variant", C.c_int), ("plane_depth", C.c_int), ("plane_h", C.c_int),
                ("plane_w", C.c_int), ("pos_freqs", C.c_int), ("hidden", C.c_int), ("box_warp", C.c_float),
                ("scale_bias", C.c_float), ("precision", C.c_int)]


_cfg = C.POINTER(DecoderConfig)

# name -> (restype, argtypes); mirrors include/ggd.h declaration by declaration
SIGNATURES = {
    "ggd_last_error": (C.c_char_p, []),
    "ggd_abi_version": (_i, []),
    "ggd_launch_count": (_u64, []),
    "ggd_raster_workspace_sizes": (_i, [_i, _i, _i, _u64, C.POINTER(_sz)]),
    "ggd_raster_preprocess": (_i, [_p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p,
                                   _f, _f, _i, _p, _p, _p, _i]),
    "ggd_raster_render": (_i, [_p, _i, _i, _i, _p, _p, _p, _p, _p, _u64, _p, _p, _i]),
    "ggd_raster_forward": (_i, [_p, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p,
                                _f, _f, _i, _p, _p, _u64, _p, _p, _p, _p, _i]),
    "ggd_raster_backward": (_i, [_p, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p,
                                 _p, _p, _u64, _p] + [_p] * 10),
    "ggd_raster_status": (_i, [_p, _p, _i, C.POINTER(C.c_uint32)]),
    "ggd_surface_points_scratch_bytes": (C.c_size_t, [_i]),
    "ggd_surface_points": (_i, [_p, _p, _i, C.c_float, C.c_uint64, _i, C.c_float, _p, C.c_size_t, _p, _p]),
    "ggd_raster_binning": (_i, [_i]),
    "ggd_raster_binning_info": (_i, [_p, _p, _i, C.POINTER(C.c_uint32)]),
    "ggd_mark_visible": (_i, [_p, _i, _p, _p, _p, _p]),
    "ggd_raster_field": (_i, [_i, C.c_char_p, _i, _i, _i, _u64, C.POINTER(_sz), C.POINTER(_sz)]),
    "ggd_stage_timing": (_i, [_i]),
    "ggd_stage_times": (_i, [C.POINTER(C.c_float)]),
    "ggd_sort_scratch_bytes": (_sz, [_u64]),
    "ggd_sort_pairs": (_i, [_p, _p, _p, _p, _p, _p, _u64, _i, _i, _p, _p, _p, _sz]),
    "ggd_sort_pairs_static": (_i, [_p, _p, _p, _p, _p, _p, _u64, _i, _i, _p, _p]),
    "ggd_decode_param_counts": (_i, [_cfg, C.POINTER(_sz), C.POINTER(_sz)]),
    "ggd_decode_param_offset": (_i, [_cfg, _i, _i, _i, C.POINTER(_sz), C.POINTER(_i), C.POINTER(_i)]),
    "ggd_decode_pack_weights": (_i, [_p, _cfg, _p, _p]),
    "ggd_planes_channels_last": (_i, [_p, _cfg, _p, _p]),
    "ggd_decode_forward": (_i, [_p, _cfg, _p, _p, _i, _p] + [_p] * 8),
    "ggd_debug_decode_profile": (_i, [C.POINTER(C.c_uint64)]),
    "ggd_mlp_forward": (_i, [_p, _p, _p, _i, _i] + [_p] * 8 + [_i, _i, _p]),
    "ggd_activation_backward": (_i, [_p, _i, _p, _p, _p, _p, _p, _p]),
    "ggd_decode_backward_scratch_bytes": (_sz, [_cfg, _i]),
    "ggd_decode_backward": (_i, [_p, _cfg, _p, _p, _p, _i, _p, _p] + [_p] * 10 + [_p, _sz, _p, _p]),
}

_lib = None


def load():
    """Load the library once; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the sm_100a extension has not been built "
            "(run __graft_entry__.build()); there is no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        if not hasattr(lib, name):
            continue  # optional entry points are resolved lazily by their users
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.ggd_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH} has ABI version {lib.ggd_abi_version()}, this binding needs {ABI_VERSION}: "
                          "rebuild it (__graft_entry__.build())")
    _lib = lib
    return lib


def check(code):
    if code != GGD_OK:
        raise GGDError(code, load().ggd_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device (or host) address of a torch tensor; None/empty tensors map to NULL
    exactly like the reference's empty-tensor placeholders
    (DGR/diff_gaussian_rasterization/__init__.py:197-207)."""
    if t is None or t.numel() == 0:
        return None
    return C.c_void_p(t.data_ptr())
