"""Stand-ins for the upstream generator (EG3D / PanoHead are out of scope: SURVEY §1)."""
import torch


from gaussian_gan_decoder_b200.synthetic import SyntheticGenerator as StubG  # noqa: E402,F401


def load_flat_into(decoder, flat):
    """Copy a flat parameter vector (ggd.h / get_params_custom order) into a decoder's modules."""
    off = 0
    with torch.no_grad():
        for h in ("xyz", "scale", "rotation", "opacity", "color"):
            for lin in getattr(decoder, f"{h}_decoder").linears():
                for t in (lin.weight, lin.bias):
                    n = t.numel()
                    t.copy_(flat[off:off + n].view_as(t).to(t.device))
                    off += n
    assert off == flat.numel()
