"""Generates tests/golden/snapshot_decoder.pkl (+ .pt, + expected flat parameters) with the REFERENCE's own classes and
its own snapshot code path (main/train_pano2gaussian_decoder.py:297-307: pickle.dump of the persistent decoder object
and torch.save of its state_dict).  Run in the build container:   python tests/golden/make_decoder_pickle.py
The generator is a tiny stand-in (a real EG3D / PanoHead G is hundreds of MB and out of scope)."""
import os
import pickle
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
sys.path[:0] = [REF, os.path.join(REF, "main"), os.path.join(REF, "PanoHead")]

from main.decoder_models.sequential_decoder_reverse import SequentialDecoderReverse   # noqa: E402  (reference)
from torch_utils import persistence                                                    # noqa: E402  (reference)
from training.volumetric_rendering.renderer import generate_planes                     # noqa: E402  (reference)
import dnnlib                                                                          # noqa: E402  (reference)


class _Renderer:
    def __init__(self, axes):
        self.plane_axes = axes


@persistence.persistent_class
class TinyG(torch.nn.Module):
    def __init__(self, depth=3):
        super().__init__()
        self.rendering_kwargs = {"box_warp": 1.0, "triplane_depth": depth}
        self.renderer = _Renderer(generate_planes())
        self.w = torch.nn.Parameter(torch.zeros(3))


torch.manual_seed(11)
dec = SequentialDecoderReverse(TinyG(), hidden_dim=128, use_xyz_embedding=False, use_gen_finetune=True, device="cpu")
with torch.no_grad():
    for p in dec.parameters():
        p.copy_(p.half().float())          # fp16-representable: the expected values below store in 2 bytes each
snapshot = dict(training_set_kwargs=dnnlib.EasyDict(lr=9e-5, num_iter=10), decoder=dec, dataloader=None)
with open(os.path.join(HERE, "snapshot_decoder.pkl"), "wb") as f:
    pickle.dump(snapshot, f)
torch.save(dec.state_dict(), os.path.join(HERE, "snapshot_decoder.pt"))
flat = torch.cat([p.detach().reshape(-1) for p in dec.get_params_custom()[:40]])
np.savez_compressed(os.path.join(HERE, "snapshot_decoder_expected.npz"), flat_fp16=flat.half().numpy())
print("wrote snapshot_decoder.{pkl,pt} and expected flat params", flat.numel())
