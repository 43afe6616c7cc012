"""``Decoder``: the 4-layer GELU MLP head.  Same constructor, attribute names and
``state_dict`` keys (``backbone.{0,2,4,6}.{weight,bias}``) as the reference
(main/decoder_models/base_decoder.py:8-27), so reference checkpoints load unchanged.
The module only OWNS parameters; the math runs in the fused CUDA kernel driven by the
wrapper classes (see fused.py)."""
import torch
from torch import nn


class Decoder(nn.Module):
    def __init__(self, n_features, out_features=3, hidden_dim=128):
        super().__init__()
        self.n_features = n_features
        self.out_features = out_features
        self.hidden_dim = hidden_dim
        self.backbone = torch.nn.Sequential(
            torch.nn.Linear(in_features=n_features, out_features=hidden_dim),
            torch.nn.GELU(),
            torch.nn.Linear(in_features=hidden_dim, out_features=hidden_dim),
            torch.nn.GELU(),
            torch.nn.Linear(in_features=hidden_dim, out_features=hidden_dim),
            torch.nn.GELU(),
            torch.nn.Linear(in_features=hidden_dim, out_features=out_features),
        )

    def linears(self):
        return [self.backbone[i] for i in (0, 2, 4, 6)]

    def forward(self, triplane_features, gaussian_features):
        """[3,P,32], [P,k] -> [P,out] (reference signature).  Routed through the same CUDA
        kernel as the fused path; there is no PyTorch fallback."""
        from .fused import single_head_forward
        return single_head_forward(self, triplane_features, gaussian_features)
