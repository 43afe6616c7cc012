"""End-to-end training step on one GPU: loss goes down, gradients reach all 40 tensors."""
import pytest
import torch

from _stubs import StubG
from gaussian_gan_decoder_b200 import synthetic
from gaussian_gan_decoder_b200.decoder_models import SequentialDecoderReverse
from gaussian_gan_decoder_b200.gaussian_renderer import GaussianModel, render_simple
from gaussian_gan_decoder_b200.training import training_step

pytestmark = pytest.mark.gpu


def test_training_step_reduces_loss(cuda):
    g = synthetic.gen(5)
    planes = synthetic.planes("panohead", g, res=32, depth=3).to(cuda)
    xyz = synthetic.head_shell_points(20_000, g).to(cuda)
    torch.manual_seed(1)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", 3), use_xyz_embedding=False, use_gen_finetune=False, device=cuda)
    cam = synthetic.camera(128, 12.0)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).contiguous().to(cuda))
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    target = torch.full((3, 128, 128), 0.2, device=cuda)
    opt = torch.optim.Adam(dec.get_params_custom(), lr=1e-3)
    z, c = torch.zeros(1, 512, device=cuda), torch.zeros(1, 25, device=cuda)
    losses = [float(training_step(dec, opt, z, c, xyz, cam, target, bg)) for _ in range(12)]
    assert all(l == l for l in losses)
    assert losses[-1] < losses[0] * 0.9, losses
    # parameters are still views of the flat buffer after optimiser steps, packed weights follow them
    assert dec._flat.in_sync(dec)
    with torch.no_grad():
        img = render_simple(cam, GaussianModel.from_decoder(dec(z, c, xyz, 1.0), dec.activated()), bg)["render"]
    assert abs(float((img - target).abs().mean() + ((img - target) ** 2).mean()) - losses[-1]) < 0.05
