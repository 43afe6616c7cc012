"""the reference's decoder snapshots open into this package's classes.  The fixtures were written by the
reference's own classes through its own snapshot code (tests/golden/make_decoder_pickle.py: pickle.dump of the
persistent decoder, torch.save of its state_dict); nothing of the reference is importable when this test runs."""
import os
import sys

import numpy as np
import pytest
import torch

from _stubs import StubG
from gaussian_gan_decoder_b200.decoder_models import ParallelDecoder, SequentialDecoderReverse
from gaussian_gan_decoder_b200.decoder_utils import checkpoint

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXPECT = torch.from_numpy(np.load(os.path.join(GOLD, "snapshot_decoder_expected.npz"))["flat_fp16"].astype(np.float32))


def _flat(dec):
    return torch.cat([p.detach().reshape(-1).cpu() for p in dec.get_params_custom()[:40]])


def _G():
    return StubG(torch.zeros(1, 3, 96, 8, 8), "panohead", 3)


def test_pickle_opens_without_the_reference_tree():
    for m in ("torch_utils", "dnnlib", "training"):
        assert m not in sys.modules or "reference" not in (getattr(sys.modules[m], "__file__", "") or "")
    dec, info = checkpoint.load_decoder_pickle(os.path.join(GOLD, "snapshot_decoder.pkl"), _G(), device="cpu")
    assert isinstance(dec, SequentialDecoderReverse) and info["class_name"] == "SequentialDecoderReverse"
    assert info["init_kwargs"] == dict(hidden_dim=128, use_xyz_embedding=False, use_gen_finetune=True)
    assert torch.equal(_flat(dec), EXPECT)
    # the pickled generator comes back as an inert record with its state, never as executed code
    assert info["generator"] is not None and info["generator"].class_name == "TinyG"
    assert "w" in info["generator"].state["_parameters"]
    assert set(info["extra"]) == {"training_set_kwargs", "dataloader"}
    # parameters are live views of the flat buffer the kernels read
    dec._ensure_flat(dec.G.planes)
    assert dec._flat.in_sync(dec) and torch.equal(dec._flat.flat.cpu(), EXPECT)


def test_state_dict_file_loads_and_mismatches_fail_loudly():
    dec = SequentialDecoderReverse(_G(), use_xyz_embedding=False, use_gen_finetune=False, device="cpu")
    rest = checkpoint.load_state_dict_file(dec, os.path.join(GOLD, "snapshot_decoder.pt"))
    assert torch.equal(_flat(dec), EXPECT)
    assert list(rest) == ["G.w"]
    other = ParallelDecoder(_G(), use_xyz_embedding=False, use_gen_finetune=False, device="cpu")
    with pytest.raises((KeyError, RuntimeError)):          # chained heads have wider first layers: shapes differ
        checkpoint.load_state_dict_file(other, os.path.join(GOLD, "snapshot_decoder.pt"))
    with pytest.raises(KeyError):
        checkpoint.load_state_dict_file(dec, {"xyz_decoder.backbone.0.weight": torch.zeros(128, 46)})
