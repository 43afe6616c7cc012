"""Autograd entry of the fused decoder (backward kernels: see csrc/decode_bwd.cu)."""


def apply(owner, planes, xyz):
    raise NotImplementedError(
        "decoder backward is not built in this revision: call the decoder under torch.no_grad() "
        "(there is deliberately no PyTorch fallback)")
