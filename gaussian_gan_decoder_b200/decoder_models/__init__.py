from .base_decoder import Decoder  # noqa: F401
from .parallel_decoder import ParallelDecoder  # noqa: F401
from .sequential_decoder import SequentialDecoder  # noqa: F401
from .sequential_decoder_reverse import SequentialDecoderReverse  # noqa: F401
