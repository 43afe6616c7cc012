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
    H, K = decoder.hidden_dim, 128
    if H > K:
        raise NotImplementedError(f"hidden_dim={H}: the kernel is {K} units wide")
    if H == K and all(l.weight.is_contiguous() and l.bias.is_contiguous() for l in lin):
        w = [(l.weight, l.bias) for l in lin]
    else:       # a narrower head (or one whose parameters are strided views of a padded buffer): zero-padded copies
        def pad(t, *shape):
            out = torch.zeros(*shape, dtype=torch.float32, device=dev)
            out[tuple(slice(0, n) for n in t.shape)] = t.detach()
            return out
        w = [(pad(lin[0].weight, K, decoder.n_features), pad(lin[0].bias, K)), (pad(lin[1].weight, K, K), pad(lin[1].bias, K)),
             (pad(lin[2].weight, K, K), pad(lin[2].bias, K)), (pad(lin[3].weight, decoder.out_features, K), lin[3].bias.detach().contiguous())]
    p = _lib.ptr
    with torch.cuda.device(dev):
        _lib.check(lib.ggd_mlp_forward(C.c_void_p(torch.cuda.current_stream().cuda_stream), p(tri), p(gf), k, P,
                                       p(w[0][0]), p(w[0][1]), p(w[1][0]), p(w[1][1]),
                                       p(w[2][0]), p(w[2][1]), p(w[3][0]), p(w[3][1]),
                                       K, decoder.out_features, p(out)))
    return out
