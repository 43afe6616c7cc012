"""B200-native (sm_100a) implementation of gaussian_gan_decoder's per-frame hot
path: tri-plane sampling + decoder MLP -> tile rasteriser, behind the
reference's own Python API.  See DESIGN.md."""
__all__ = ["rasterizer", "cameras", "synthetic"]
