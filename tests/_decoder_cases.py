"""Loader for the decoder golden fixtures (made by tests/golden/make_decoder_golden.py from
the reference's own Python classes)."""
import glob
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "decoder_*.npz")))
HEADS = ("xyz", "scale", "rotation", "opacity", "color")


def load(path):
    z = np.load(path)
    return dict(
        arch=str(z["arch"]), variant=str(z["variant"]), pos_freqs=int(z["pos_freqs"]), depth=int(z["depth"]),
        box_warp=float(z["box_warp"]) if "box_warp" in z.files else 1.0,
        planes=torch.from_numpy(z["planes"]), xyz=torch.from_numpy(z["xyz"]),
        flat=torch.from_numpy(z["flat_fp16"].astype(np.float32)),
        out={h: torch.from_numpy(z["out_" + h]) for h in HEADS},
    )
