"""Drop-in for the reference's ``diff_gaussian_rasterization`` package
(gaussian_splatting/submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):
put this repository's root on ``sys.path`` ahead of the reference build and
``gaussian_renderer.render_simple`` runs unchanged on the sm_100a kernels."""
from gaussian_gan_decoder_b200.rasterizer import (  # noqa: F401
    GaussianRasterizationSettings,
    GaussianRasterizer,
    rasterize_gaussians,
    _RasterizeGaussians,
    _C,
)
