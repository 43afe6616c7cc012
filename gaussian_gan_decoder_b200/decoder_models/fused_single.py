"""``Decoder.forward(triplane_features, gaussian_features)`` on its own (ggd_mlp_forward)."""
import ctypes as C

import torch

from .. import _lib


def forward(decoder, triplane_features, gaussian_features):
    lib = _lib.load()
    if not triplane_features.is_cuda:
        raise RuntimeError("Decoder.forward needs CUDA tensors: there is no CPU path")
    if torch.is_grad_enabled() and any(p.requires_grad for p in decoder.parameters()):
        raise NotImplementedError("stand-alone Decoder.forward is inference-only; train through the wrapper classes "
                                  "(call it under torch.no_grad())")
    if triplane_features.dim() != 3 or triplane_features.shape[0] != 3 or triplane_features.shape[2] != 32:
        raise ValueError(f"triplane_features must be [3,P,32], got {tuple(triplane_features.shape)}")
    dev = triplane_features.device
    P = triplane_features.shape[1]
    k = gaussian_features.shape[1]
    if 32 + k != decoder.n_features:
        raise ValueError(f"expected {decoder.n_features - 32} gaussian features, got {k}")
    tri = triplane_features.contiguous().float()
    gf = gaussian_features.contiguous().float()
    out = torch.empty(P, decoder.out_features, dtype=torch.float32, device=dev)
    if P == 0:
        return out
    lin = decoder.linears()
    p = _lib.ptr
    with torch.cuda.device(dev):
        _lib.check(lib.ggd_mlp_forward(C.c_void_p(torch.cuda.current_stream().cuda_stream), p(tri), p(gf), k, P,
                                       p(lin[0].weight), p(lin[0].bias), p(lin[1].weight), p(lin[1].bias),
                                       p(lin[2].weight), p(lin[2].bias), p(lin[3].weight), p(lin[3].bias),
                                       decoder.hidden_dim, decoder.out_features, p(out)))
    return out
