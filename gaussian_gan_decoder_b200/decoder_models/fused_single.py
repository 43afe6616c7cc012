"""``Decoder.forward(triplane_features, gaussian_features)`` on its own."""


def forward(decoder, triplane_features, gaussian_features):
    raise NotImplementedError(
        "stand-alone Decoder.forward is not built in this revision; use the wrapper classes "
        "(there is deliberately no PyTorch fallback)")
