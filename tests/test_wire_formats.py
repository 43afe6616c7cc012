"""Wire / disk formats either side of the path: the reference's decoder
``state_dict`` key names and shapes, and the .ply layout of decoded Gaussians."""
import glob
import json
import os

import numpy as np
import torch

from _stubs import StubG
from gaussian_gan_decoder_b200.decoder_models import ParallelDecoder, SequentialDecoder, SequentialDecoderReverse
from gaussian_gan_decoder_b200.easydict import EasyDict
from gaussian_gan_decoder_b200.gaussian_renderer import GaussianModel

HERE = os.path.dirname(os.path.abspath(__file__))
CLS = {"eg3d_parallel": (ParallelDecoder, "eg3d", 1, False), "eg3d_sequential_posenc": (SequentialDecoder, "eg3d", 1, True),
       "panohead_reverse": (SequentialDecoderReverse, "panohead", 3, False),
       "panohead_parallel_posenc": (ParallelDecoder, "panohead", 3, True),
       "panohead_sequential_boxwarp07": (SequentialDecoder, "panohead", 3, False),
       "eg3d_reverse_nonsquare_boxwarp2": (SequentialDecoderReverse, "eg3d", 1, True)}


def test_state_dict_keys_and_shapes_match_the_reference_classes():
    files = sorted(glob.glob(os.path.join(HERE, "golden", "state_dict_keys_*.json")))
    assert len(files) == len(CLS)
    for path in files:
        name = os.path.basename(path)[len("state_dict_keys_"):-len(".json")]
        cls, arch, depth, posenc = CLS[name]
        ref = json.load(open(path))
        dec = cls(StubG(torch.zeros(1, 3, 32 * depth, 4, 4), arch, depth), use_xyz_embedding=posenc, device="cpu")
        ours = {k: list(v.shape) for k, v in dec.state_dict().items() if not k.startswith("G.")}
        assert ours == ref, name                      # same keys, same order, same shapes
        # a reference checkpoint loads with strict key checking
        sd = {k: torch.randn(*s) for k, s in ref.items()}
        missing, unexpected = dec.load_state_dict(sd, strict=False)
        assert [m for m in missing if not m.startswith("G.")] == [] and unexpected == []
        assert torch.equal(dec.color_decoder.backbone[6].weight, sd["color_decoder.backbone.6.weight"])


def test_ply_round_trip_and_header(tmp_path):
    g = torch.Generator().manual_seed(0)
    P = 1234
    res = EasyDict(xyz=torch.randn(P, 3, generator=g), scale=torch.randn(P, 3, generator=g), rotation=torch.randn(P, 4, generator=g),
                   opacity=torch.randn(P, 1, generator=g), color=torch.randn(P, 3, generator=g))
    gm = GaussianModel.from_decoder(res)
    path = str(tmp_path / "out" / "checkpoint.ply")
    gm.save_ply(path)
    raw = open(path, "rb").read()
    header = raw[:raw.index(b"end_header\n") + 11].decode()
    props = [l.split()[-1] for l in header.splitlines() if l.startswith("property")]
    assert props == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2", "opacity", "scale_0", "scale_1", "scale_2",
                     "rot_0", "rot_1", "rot_2", "rot_3"]
    assert "format binary_little_endian 1.0" in header and f"element vertex {P}" in header
    assert len(raw) == len(header) + P * 17 * 4
    back = GaussianModel(0).load_ply(path)
    assert torch.equal(back._xyz, res.xyz) and torch.equal(back._scaling, res.scale)
    assert torch.equal(back._rotation, res.rotation) and torch.equal(back._opacity, res.opacity)
    assert torch.equal(back._features_dc, res.color.unsqueeze(1))
    body = np.frombuffer(raw[len(header):], dtype="<f4").reshape(P, 17)
    assert np.all(body[:, 3:6] == 0)
