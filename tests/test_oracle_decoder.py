"""oracle/decoder_oracle.py against outputs of the reference's own decoder classes."""
import os
import sys

import pytest
import torch

from _decoder_cases import GOLDEN, HEADS, load

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import decoder_oracle as do  # noqa: E402


def test_fixtures_present():
    assert len(GOLDEN) >= 6
    cases = [load(p) for p in GOLDEN]
    assert any(c["box_warp"] != 1.0 for c in cases) and any(c["planes"].shape[-1] != c["planes"].shape[-2] for c in cases)


@pytest.mark.parametrize("path", GOLDEN)
def test_oracle_matches_reference_classes(path):
    c = load(path)
    with torch.no_grad():
        res = do.decode(c["arch"], c["variant"], c["pos_freqs"], c["flat"], c["planes"], c["xyz"], c["box_warp"], c["depth"])
    for h in HEADS:
        err = (res[h] - c["out"][h]).abs().max().item()
        assert err <= 2e-6, (h, err)      # same ATen ops on the same machine: essentially exact


def test_flat_layout_counts():
    # 189 966 (parallel) / 193 294 (sequential_reverse) decoder parameters
    assert do.flat_layout("parallel", 0)[1] == 189966
    assert do.flat_layout("sequential_reverse", 0)[1] == 193294
    assert do.in_dims("sequential", 0) == {"xyz": 35, "scale": 38, "rotation": 41, "opacity": 45, "color": 46}
    assert do.in_dims("sequential_reverse", 0) == {"color": 35, "opacity": 38, "rotation": 39, "scale": 43, "xyz": 46}


def test_embed_order_and_linear_bands():
    x = torch.tensor([[0.1, -0.2, 0.3]])
    e = do.embed(x, 10)
    assert e.shape == (1, 63)
    f = torch.linspace(1.0, 512.0, 10)
    assert torch.allclose(e[0, 3:6], torch.sin(x[0] * f[0])) and torch.allclose(e[0, 6:9], torch.cos(x[0] * f[0]))
    assert torch.allclose(e[0, 57:60], torch.sin(x[0] * f[9]))
