"""Positional embedding of xyz: [x, sin(f x), cos(f x)] for 10 LINEARLY spaced
frequencies linspace(1, 512, 10) -> 63 dims.  Mirrors ``Embedder``
(main/decoder_utils/pos_encoding.py:4-50 of the reference).  Inside the decoders the
embedding is computed by the fused CUDA kernel; this class only carries the configuration
(``out_dim``, ``num_freqs``) and offers the same callable for users who want the features."""
import torch


class Embedder:
    def __init__(self, include_input=True, input_dims=3, num_freqs=10, log_sampling=False):
        if not include_input or input_dims != 3 or log_sampling:
            raise NotImplementedError("only include_input=True, input_dims=3, log_sampling=False "
                                      "(the configuration the reference decoders use) is supported")
        self.num_freqs = num_freqs
        self.input_dims = input_dims
        self.include_input = include_input
        self.out_dim = input_dims * (1 + 2 * num_freqs)
        self.freq_bands = torch.linspace(2.0 ** 0.0, 2.0 ** (num_freqs - 1), steps=num_freqs)

    def __call__(self, inputs):
        outs = [inputs]
        for f in self.freq_bands.to(inputs.device):
            outs += [torch.sin(inputs * f), torch.cos(inputs * f)]
        return torch.cat(outs, -1)
