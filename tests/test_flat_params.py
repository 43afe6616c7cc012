"""Host side of the decoder's flat parameter layout (no GPU needed): the 40 nn.Linear tensors are views of one buffer in
the ggd.h layout; a decoder narrower than the kernels' 128 hidden units keeps the reference's parameter shapes and sits
zero-padded in that layout (FlatParams.view_of)."""
import pytest
import torch

from _stubs import StubG
from gaussian_gan_decoder_b200.decoder_models import ParallelDecoder, SequentialDecoderReverse, fused

HEADS = ("xyz", "scale", "rotation", "opacity", "color")


def _decoder(cls, hidden, posenc):
    planes = torch.zeros(1, 3, 32, 8, 8)
    torch.manual_seed(3)
    dec = cls(StubG(planes, "eg3d", 1), hidden_dim=hidden, use_xyz_embedding=posenc, use_gen_finetune=False, device="cpu")
    before = [p.detach().clone() for p in dec.get_params_custom()]
    dec._ensure_flat(planes)
    return dec, before


@pytest.mark.parametrize("cls,hidden,posenc", [(SequentialDecoderReverse, 128, False), (SequentialDecoderReverse, 64, False),
                                               (ParallelDecoder, 100, True), (ParallelDecoder, 1, False)])
def test_parameters_are_views_of_the_flat_buffer(cls, hidden, posenc):
    dec, before = _decoder(cls, hidden, posenc)
    flat = dec._flat
    lo, hi = flat.flat.data_ptr(), flat.flat.data_ptr() + 4 * flat.flat.numel()
    for p, b in zip(dec.get_params_custom(), before):
        assert torch.equal(p.detach(), b) and p.shape == b.shape              # values and reference shapes kept
        assert lo <= p.data_ptr() < hi and lo <= p.grad.data_ptr() - flat.flat_grad.data_ptr() + lo < hi
    # shapes are the reference's (base_decoder.py:11-27)
    lin = dec.color_decoder.linears()
    assert tuple(lin[0].weight.shape) == (hidden, dec.color_decoder.n_features)
    assert tuple(lin[1].weight.shape) == (hidden, hidden) and tuple(lin[3].weight.shape) == (3, hidden)
    # everything in the buffer that is not a parameter is zero padding
    total = sum(float(p.detach().abs().double().sum()) for p in dec.get_params_custom())
    assert abs(float(flat.flat.abs().double().sum()) - total) <= 1e-9 * max(total, 1.0)
    n_true = sum(p.numel() for p in dec.get_params_custom())
    assert (n_true == flat.flat.numel()) == (hidden == fused.KERNEL_HIDDEN)
    # an optimizer step through the views lands in the buffer and leaves the padding alone
    opt = torch.optim.SGD(dec.get_params_custom(), lr=0.5)
    flat.flat_grad.fill_(1.0)                     # as if the backward kernels had written every entry, padding included
    opt.step()
    for p, b in zip(dec.get_params_custom(), before):
        assert torch.allclose(p.detach(), b - 0.5)
    total = sum(float(p.detach().abs().double().sum()) for p in dec.get_params_custom())
    assert abs(float(flat.flat.abs().double().sum()) - total) <= 1e-9 * max(total, 1.0)
    assert flat.in_sync(dec)


def test_state_dict_of_a_narrow_decoder_round_trips():
    a, _ = _decoder(SequentialDecoderReverse, 48, False)
    b, _ = _decoder(SequentialDecoderReverse, 48, False)
    with torch.no_grad():
        for p in a.get_params_custom():
            p.add_(0.25)
    sd = {k: v.clone() for k, v in a.state_dict().items()}
    assert all(v.is_contiguous() for v in sd.values())
    res = b.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys
    for p, q in zip(a.get_params_custom(), b.get_params_custom()):
        assert torch.equal(p.detach(), q.detach())
    assert b._flat.in_sync(b) and torch.equal(a._flat.flat, b._flat.flat)


def test_wider_than_the_kernels_is_refused():
    with pytest.raises(NotImplementedError):
        ParallelDecoder(StubG(torch.zeros(1, 3, 32, 8, 8), "eg3d", 1), hidden_dim=129, device="cpu")


def test_packed_planes_host_producer_layout():
    """PackedPlanes.from_host: [1,3,32*D,H,W] fp32 -> [3][D][H][W][32] fp16, generator channel c*D + d -> (d, c)."""
    from gaussian_gan_decoder_b200.decoder_models.fused import PackedPlanes
    D, H, W = 3, 4, 5
    x = torch.arange(3 * 32 * D * H * W, dtype=torch.float32).view(1, 3, 32 * D, H, W) / 64.0
    pk = PackedPlanes.from_host(x, pin=False)
    assert pk.shape == (1, 3, 32 * D, H, W) and pk.data.dtype == torch.uint8 and pk.data.numel() == x.numel() * 2
    got = pk.data.view(torch.float16).view(3, D, H, W, 32)
    for (p, d, y, xx, c) in [(0, 0, 0, 0, 0), (2, 1, 3, 4, 31), (1, 2, 2, 1, 7)]:
        assert got[p, d, y, xx, c] == x[0, p, c * D + d, y, xx].to(torch.float16)
    with pytest.raises(ValueError):
        PackedPlanes.from_host(torch.zeros(1, 3, 30, 4, 4))
