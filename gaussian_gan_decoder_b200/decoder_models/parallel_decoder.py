"""``ParallelDecoder``: five independent heads on [tri-plane features | position].
Drop-in for main/decoder_models/parallel_decoder.py:12-92 of the reference."""
from ._base import FusedDecoderBase


class ParallelDecoder(FusedDecoderBase):
    VARIANT = "parallel"
    ORDER = ("xyz", "scale", "rotation", "opacity", "color")
    SCALE_BIAS = -2.0
