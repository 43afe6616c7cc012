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


def _setup(cuda, seed, P=20_000, res=32, size=128):
    g = synthetic.gen(seed)
    planes = synthetic.planes("panohead", g, res=res, depth=3).to(cuda)
    xyz = synthetic.head_shell_points(P, g).to(cuda)
    cam = synthetic.camera(size, 12.0)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).contiguous().to(cuda))
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    return planes, xyz, cam, bg


def _decoder(planes, cuda, seed=1, hidden=128):
    torch.manual_seed(seed)
    return SequentialDecoderReverse(StubG(planes, "panohead", 3), hidden_dim=hidden, use_xyz_embedding=False, use_gen_finetune=False,
                                    device=cuda)


@pytest.mark.parametrize("mode,hidden", [("target", 128), ("dL_dimage", 128), ("target", 48)])
def test_train_pipeline_matches_autograd_route(cuda, mode, hidden):
    """TrainPipeline (pre-allocated buffers, C ABI end to end, no autograd, no host sync inside a step) takes the same
    optimisation step as ``training_step`` (decoder classes + GaussianModel + render_simple under autograd): compared
    through the parameters after one SGD step, i.e. through the full flat gradient.  Tolerance: the backward sums are
    fp32 atomics in both routes (order differs from run to run), 1e-3 of the largest gradient entry."""
    from gaussian_gan_decoder_b200.training import TrainPipeline
    planes, xyz, cam, bg = _setup(cuda, 6)
    z, c = torch.zeros(1, 512, device=cuda), torch.zeros(1, 25, device=cuda)
    target = torch.full((3, 128, 128), 0.3, device=cuda)
    dimg = torch.randn(3, 128, 128, generator=synthetic.gen(7)).to(cuda)
    kw = dict(target=target) if mode == "target" else dict(dL_dimage=dimg)
    a, b = _decoder(planes, cuda, hidden=hidden), _decoder(planes, cuda, hidden=hidden)
    a._ensure_flat(planes), b._ensure_flat(planes)
    before = a._flat.flat.clone()
    assert torch.equal(before, b._flat.flat)
    loss_a = training_step(a, torch.optim.SGD(a.get_params_custom(), lr=1.0), z, c, xyz, cam, kw.get("target"), bg,
                           dL_dimage=kw.get("dL_dimage"))
    tp = TrainPipeline(b, torch.optim.SGD(b.get_params_custom(), lr=1.0), xyz.shape[0], 128, cuda)
    tp.step(planes, xyz, cam, bg, **kw)
    R, redo = tp.check()
    assert not redo and R > 0
    ga, gb = before - a._flat.flat, before - b._flat.flat
    scale = float(ga.abs().max())
    assert scale > 0
    assert float((ga - gb).abs().max()) <= 1e-3 * scale, (float((ga - gb).abs().max()), scale)
    assert abs(float(loss_a) - float(tp.loss)) <= 1e-4 * max(1.0, abs(float(loss_a)))
    assert b._flat.in_sync(b)
    if hidden < 128:      # the zero padding of a narrow decoder took no step
        true = sum(float(p.detach().abs().double().sum()) for p in b.get_params_custom())
        assert abs(float(b._flat.flat.abs().double().sum()) - true) <= 1e-6 * true


def test_train_pipeline_reduces_loss_and_recovers_from_overflow(cuda):
    from gaussian_gan_decoder_b200.training import TrainPipeline
    planes, xyz, cam, bg = _setup(cuda, 5)
    dec = _decoder(planes, cuda)
    target = torch.full((3, 128, 128), 0.2, device=cuda)
    tp = TrainPipeline(dec, torch.optim.Adam(dec.get_params_custom(), lr=1e-3), xyz.shape[0], 128, cuda, capacity=1000)
    tp.step(planes, xyz, cam, bg, target=target)
    R, redo = tp.check()
    assert redo and tp.capacity >= R            # the first guess was too small on purpose: buffer grown, step to be redone
    losses = []
    for _ in range(12):
        tp.step(planes, xyz, cam, bg, target=target)
        losses.append(float(tp.loss))
    assert not tp.check()[1]
    assert all(l == l for l in losses) and losses[-1] < losses[0] * 0.9, losses
