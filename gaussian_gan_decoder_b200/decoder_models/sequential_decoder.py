"""``SequentialDecoder``: xyz -> scale -> rotation -> opacity -> color, each head also
fed the earlier heads' outputs.  Drop-in for main/decoder_models/sequential_decoder.py:12-96;
like the reference it passes the camera to ``G.mapping`` and zeros to ``G.synthesis``."""
from ._base import FusedDecoderBase


class SequentialDecoder(FusedDecoderBase):
    VARIANT = "sequential"
    ORDER = ("xyz", "scale", "rotation", "opacity", "color")
    SCALE_BIAS = -2.0
    MAPPING_GETS_CAMERA = True
