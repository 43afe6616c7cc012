"""``SequentialDecoderReverse`` (the reference's default ``--decoder_type``): color ->
opacity -> rotation -> scale -> xyz.  Drop-in for
main/decoder_models/sequential_decoder_reverse.py:12-99; scale offset -2.5."""
from ._base import FusedDecoderBase


class SequentialDecoderReverse(FusedDecoderBase):
    VARIANT = "sequential_reverse"
    ORDER = ("color", "opacity", "rotation", "scale", "xyz")
    SCALE_BIAS = -2.5
