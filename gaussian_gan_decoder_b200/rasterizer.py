"""Host-side mirror of ``diff_gaussian_rasterization`` over the ggd C ABI.

Same names, argument meaning and error behaviour as the reference binding
(DGR/diff_gaussian_rasterization/__init__.py:21-220 and the pybind layer
DGR/rasterize_points.cu:35-217, DGR/ext.cpp:15-19; DGR =
gaussian_splatting/submodules/diff-gaussian-rasterization), so that
``gaussian_renderer.render_simple`` runs against it unchanged.  PyTorch is used
for device memory, streams and autograd plumbing only; every kernel is reached
through ``libggd_b200.so``.

Differences that a caller can observe:
  * the three state buffers have our own (opaque) layout — they are only ever
    handed back to ``rasterize_gaussians_backward``, as in the reference;
  * the binning buffer is sized before the frame from the previous frame's
    instance count, so a frame runs without a mid-frame host sync; an overflow
    (instance count grew past the guess) re-runs the binning/blend stage once.
"""
from typing import NamedTuple
import ctypes as C
import threading

import torch
import torch.nn as nn

from . import _lib


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _f32(t, name):
    if t is None:
        return None
    if t.numel() and t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32 (got {t.dtype})")
    return t.contiguous()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# The reference-order key/value arrays ("point_list_keys_unsorted", "point_list_unsorted": duplicateWithKeys' output) are
# read by nothing but ggd_raster_field users, i.e. tests and debugging; a frame binned per tile skips writing them
# (GGD_BINNING_LEAN, 12 bytes per instance) unless this is set.  Everything downstream is unchanged either way.
keep_reference_order_arrays = False

# instance-count guess per (device, P, H, W): drives the binning-buffer capacity
_capacity_hint = {}
_tls = threading.local()


def _pinned_u32(device):
    """Pinned word the frame's instance count is copied into.  One per host thread and device: a call
    synchronises its stream before it returns, so the word is free again when the thread's next call
    starts, and two threads never share one."""
    key = device.index if device.index is not None else torch.cuda.current_device()
    words = getattr(_tls, "words", None)
    if words is None:
        words = _tls.words = {}
    if key not in words:
        words[key] = torch.zeros(4, dtype=torch.int32).pin_memory()
    return words[key]


def workspace_sizes(P, W, H, capacity):
    out = (C.c_size_t * 3)()
    _lib.check(_lib.load().ggd_raster_workspace_sizes(P, W, H, capacity, out))
    return int(out[0]), int(out[1]), int(out[2])


def field(buf, which, name, P, W, H, capacity, dtype):
    """View of a named array inside a state buffer (tests / debugging)."""
    off, cnt = C.c_size_t(), C.c_size_t()
    _lib.check(_lib.load().ggd_raster_field(which, name.encode(), P, W, H, capacity, C.byref(off), C.byref(cnt)))
    nbytes = cnt.value * torch.empty((), dtype=dtype).element_size()
    return buf[off.value: off.value + nbytes].view(dtype)


class _CNamespace:
    """Stand-in for the reference's pybind module ``diff_gaussian_rasterization._C``."""

    @staticmethod
    def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                            cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                            image_width, sh, degree, campos, prefiltered, debug):
        lib = _lib.load()
        if means3D.ndimension() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        if not means3D.is_cuda:
            raise RuntimeError("means3D must be a CUDA tensor: this rasteriser has no CPU path")
        dev = means3D.device
        P, H, W = means3D.size(0), int(image_height), int(image_width)
        out_color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        empty = torch.empty((0,), dtype=torch.uint8, device=dev)
        if P == 0:
            return 0, out_color, radii, empty, empty.clone(), empty.clone()

        M = sh.size(1) if sh.numel() != 0 else 0
        background = _f32(background, "bg"); means3D = _f32(means3D, "means3D")
        colors = _f32(colors, "colors_precomp"); opacity = _f32(opacity, "opacities")
        scales = _f32(scales, "scales"); rotations = _f32(rotations, "rotations")
        cov3D_precomp = _f32(cov3D_precomp, "cov3D_precomp"); sh = _f32(sh, "shs")
        viewmatrix = _f32(viewmatrix, "viewmatrix"); projmatrix = _f32(projmatrix, "projmatrix")
        campos = _f32(campos, "campos")
        if M == 0 and colors.numel() == 0:
            raise RuntimeError("For non-RGB, provide precomputed Gaussian colors!")

        p = _lib.ptr
        key = (dev.index, P, H, W)
        hint = _capacity_hint.get(key)
        # Stage-2 binning (ggd_raster_binning): counting instances per tile costs one atomic per (splat, tile),
        # wasted on frames whose tile lists are too long for the per-tile path.  Remember per shape whether
        # the last frame qualified; a shape that did not is rendered radix-only and probed again now and then.
        # The mode travels with the call (an argument of both stages), never through process state.
        lib_mode = lib.ggd_raster_binning(-1)
        bh = _binning_hint.setdefault(key, [min(lib_mode, 1), 0])
        mode = bh[0] if lib_mode else 0
        if mode == 0 and lib_mode != 0:
            bh[1] += 1
            if bh[1] % 64 == 0:
                mode = 1
        pinned = _pinned_u32(dev)
        host_R = C.c_void_p(pinned.data_ptr())
        with torch.cuda.device(dev):
            s = _stream()
            geom_b, _, img_b = workspace_sizes(P, W, H, 0)
            geom = torch.empty(geom_b, dtype=torch.uint8, device=dev)
            img = torch.empty(img_b, dtype=torch.uint8, device=dev)
            pre_args = (s, P, int(degree), M, W, H, p(means3D), p(sh), p(colors), p(opacity), p(scales),
                        float(scale_modifier), p(rotations), p(cov3D_precomp), p(viewmatrix), p(projmatrix),
                        p(campos), float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
                        p(geom), p(radii), host_R, mode)
            lean = 0 if keep_reference_order_arrays else _lib.BINNING_LEAN

            def render(cap):
                binning = torch.empty(workspace_sizes(P, W, H, cap)[1], dtype=torch.uint8, device=dev)
                _lib.check(lib.ggd_raster_render(s, P, W, H, p(background), p(colors), p(radii), p(geom),
                                                 p(binning), cap, p(img), p(out_color), mode | lean))
                return binning

            _lib.check(lib.ggd_raster_preprocess(*pre_args))
            if hint is None:
                torch.cuda.current_stream().synchronize()       # first frame of this shape only
                R = int(pinned[0].item()) & 0xffffffff
                cap = max(R, 1)
                binning = render(cap)
            else:
                cap = hint
                binning = render(cap)
                torch.cuda.current_stream().synchronize()       # end of frame, nothing queued behind it
                R = int(pinned[0].item()) & 0xffffffff
                if R > cap:                                     # guess too small: redo binning + blend
                    cap = R
                    binning = render(cap)
            if debug:
                torch.cuda.current_stream().synchronize()
            _capacity_hint[key] = max(int(R * 1.25) + 4096, 4096)
            if mode:
                info, st = (C.c_uint32 * 2)(), (C.c_uint32 * 2)()
                _lib.check(lib.ggd_raster_binning_info(s, p(geom), P, info))
                _lib.check(lib.ggd_raster_status(s, p(geom), P, st))
                bh[0] = 1 if info[0] else 0
                if st[1] & 8:      # per-tile sort refused a list (depths too clustered): render again with the radix sort
                    mode, bh[0] = 0, 0
                    binning = render(cap)
                    torch.cuda.current_stream().synchronize()
        # capacity travels with the buffer so backward can re-carve it
        binning.ggd_capacity = cap
        return R, out_color, radii, geom, binning, img

    @staticmethod
    def rasterize_gaussians_backward(*args):
        from . import raster_backward
        return raster_backward.rasterize_gaussians_backward(*args)

    @staticmethod
    def mark_visible(means3D, viewmatrix, projmatrix):
        lib = _lib.load()
        P = means3D.size(0)
        present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
        if P != 0:
            means3D = _f32(means3D, "means3D"); viewmatrix = _f32(viewmatrix, "viewmatrix")
            projmatrix = _f32(projmatrix, "projmatrix")
            with torch.cuda.device(means3D.device):
                _lib.check(lib.ggd_mark_visible(_stream(), P, _lib.ptr(means3D), _lib.ptr(viewmatrix),
                                                _lib.ptr(projmatrix), C.c_void_p(present.data_ptr())))
        return present


_C = _CNamespace()
_binning_hint = {}


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


def _cpu_copy(args):
    return tuple(a.cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        if rs.debug:
            snapshot = _cpu_copy(args)
            try:
                num_rendered, color, radii, geom, binning, img = _C.rasterize_gaussians(*args)
            except Exception:
                torch.save(snapshot, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            num_rendered, color, radii, geom, binning, img = _C.rasterize_gaussians(*args)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.bin_capacity = getattr(binning, "ggd_capacity", num_rendered)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning,
         img) = ctx.saved_tensors
        binning.ggd_capacity = ctx.bin_capacity
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color,
                sh, rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, rs.debug)
        if rs.debug:
            snapshot = _cpu_copy(args)
            try:
                grads = _C.rasterize_gaussians_backward(*args)
            except Exception:
                torch.save(snapshot, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            grads = _C.rasterize_gaussians_backward(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
         grad_sh, grad_scales, grad_rotations) = grads
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities,
                grad_scales, grad_rotations, grad_cov3Ds_precomp, None)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)
