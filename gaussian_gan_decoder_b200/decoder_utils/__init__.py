"""Host-side helpers around the decoders: positional embedding, on-GPU surface point sets."""
