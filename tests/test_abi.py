"""The C-ABI library loads and exports every symbol include/ggd.h declares; host-only
entry points behave (no GPU needed)."""
import ctypes as C
import os
import re

import pytest

from gaussian_gan_decoder_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "ggd.h")).read()
    return sorted(set(re.findall(r"GGD_API[^;]*?\b(ggd_\w+)\s*\(", hdr)))


def test_header_declares_something():
    syms = declared_symbols()
    assert "ggd_raster_forward" in syms and "ggd_sort_pairs" in syms and len(syms) >= 10


def test_all_declared_symbols_exported_and_bound():
    lib = _lib.load()
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in ggd.h but not exported by libggd_b200.so"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature in _lib.py"


def test_workspace_sizes_monotone_and_aligned():
    lib = _lib.load()
    prev = None
    for P, cap in [(1, 0), (1000, 5000), (500000, 8_000_000)]:
        out = (C.c_size_t * 3)()
        assert lib.ggd_raster_workspace_sizes(P, 512, 512, cap, out) == 0
        sizes = tuple(out)
        assert all(s % 256 == 0 and s > 0 for s in sizes)
        if prev:
            assert sizes[0] > prev[0] and sizes[1] > prev[1]
        prev = sizes


def test_bad_arguments_fail_loudly_without_a_gpu():
    lib = _lib.load()
    out = (C.c_size_t * 3)()
    assert lib.ggd_raster_workspace_sizes(10, 512, 512, 1 << 30, out) == _lib.GGD_E_INVALID
    assert b"2^30" in lib.ggd_last_error()
    # P <= 0 and null pointers are rejected before any CUDA call
    assert lib.ggd_raster_preprocess(None, 0, 0, 1, 64, 64, *([None] * 5), 1.0, *([None] * 5), 1.0, 1.0, 0,
                                     None, None, None, -1) == _lib.GGD_E_INVALID
    with pytest.raises(_lib.GGDError):
        _lib.check(lib.ggd_mark_visible(None, 5, None, None, None, None))


def test_field_locator():
    lib = _lib.load()
    off, cnt = C.c_size_t(), C.c_size_t()
    assert lib.ggd_raster_field(0, b"depths", 100, 64, 64, 0, C.byref(off), C.byref(cnt)) == 0
    assert off.value == 0 and cnt.value == 100
    assert lib.ggd_raster_field(2, b"ranges", 100, 64, 48, 0, C.byref(off), C.byref(cnt)) == 0
    assert cnt.value == 2 * 4 * 3
    assert lib.ggd_raster_field(1, b"nope", 100, 64, 64, 10, C.byref(off), C.byref(cnt)) == _lib.GGD_E_INVALID


def test_drop_in_package_names():
    import diff_gaussian_rasterization as d
    assert d.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    for n in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert hasattr(d._C, n)
    r = d.GaussianRasterizer(None)
    import torch
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r.forward(torch.zeros(1, 3), None, torch.zeros(1, 1))
    with pytest.raises(Exception, match="scale/rotation pair"):
        r.forward(torch.zeros(1, 3), None, torch.zeros(1, 1), shs=torch.zeros(1, 1, 3))
