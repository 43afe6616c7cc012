"""``_C.rasterize_gaussians_backward`` over the ggd C ABI (reference:
DGR/rasterize_points.cu:117-196).  Same 21 arguments, same 8 returned tensors."""
import ctypes as C

import torch

from . import _lib


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh,
                                 degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug):
    lib = _lib.load()
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
    dmeans3D, dmeans2D, dcolors, dconic = z(P, 3), z(P, 3), z(P, 3), z(P, 2, 2)
    dopacity, dcov3D, dsh, dscales, drot = z(P, 1), z(P, 6), z(P, M, 3), z(P, 3), z(P, 4)
    if P != 0:
        cap = getattr(binningBuffer, "ggd_capacity", None)
        if cap is None:
            raise RuntimeError("binning buffer without its capacity: pass the tensor returned by rasterize_gaussians")
        f = lambda t: None if t is None or t.numel() == 0 else t.contiguous().float()
        background, means3D, colors, scales, rotations = f(background), f(means3D), f(colors), f(scales), f(rotations)
        cov3D_precomp, viewmatrix, projmatrix, campos, sh = f(cov3D_precomp), f(viewmatrix), f(projmatrix), f(campos), f(sh)
        g = dL_dout_color.contiguous().float()
        p = _lib.ptr
        with torch.cuda.device(dev):
            s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(lib.ggd_raster_backward(
                s, P, int(degree), M, p(background), W, H, p(means3D), p(sh), p(colors), p(scales),
                float(scale_modifier), p(rotations), p(cov3D_precomp), p(viewmatrix), p(projmatrix), p(campos),
                float(tan_fovx), float(tan_fovy), p(radii), p(geomBuffer), p(binningBuffer), int(cap), p(imageBuffer),
                p(g), p(dmeans2D), p(dconic), p(dopacity), p(dcolors), p(dmeans3D), p(dcov3D), p(dsh), p(dscales),
                p(drot)))
            if debug:
                torch.cuda.current_stream().synchronize()
    return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot
