"""Generates tests/golden/decoder_*.npz by running the reference's OWN decoder classes
(imported from /root/reference, CPU) on seeded inputs.  Run in the build container:

    python tests/golden/make_decoder_golden.py            # spawns one process per arch

A stub generator stands in for EG3D / PanoHead: mapping() returns zeros and synthesis()
returns seeded random feature_planes (the generator is upstream of the path).
Weights are rounded to fp16-representable values before the reference runs, so the
fixture can store them in 2 bytes each without changing the answer.
"""
import os
import subprocess
import zlib
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"

CASES = {
    # name: (arch, variant-class, use_xyz_embedding, P, plane_res, depth)
    "eg3d_parallel": ("eg3d", "ParallelDecoder", False, 700, 16, 1),
    "eg3d_sequential_posenc": ("eg3d", "SequentialDecoder", True, 300, 16, 1),
    "panohead_reverse": ("panohead", "SequentialDecoderReverse", False, 700, 16, 3),
    "panohead_parallel_posenc": ("panohead", "ParallelDecoder", True, 300, 12, 3),
    # plane_res may be (H, W); a seventh entry is rendering_kwargs["box_warp"] (1.0 above; the released PanoHead uses 0.7)
    "panohead_sequential_boxwarp07": ("panohead", "SequentialDecoder", False, 500, 12, 3, 0.7),
    "eg3d_reverse_nonsquare_boxwarp2": ("eg3d", "SequentialDecoderReverse", True, 300, (10, 18), 1, 2.0),
}
VARIANT = {"ParallelDecoder": "parallel", "SequentialDecoder": "sequential",
           "SequentialDecoderReverse": "sequential_reverse"}
HEADS = ("xyz", "scale", "rotation", "opacity", "color")


class _Renderer:
    def __init__(self, axes):
        self.plane_axes = axes


class StubG(torch.nn.Module):
    """Module-level (the reference's persistent_class pickles constructor arguments)."""

    def __init__(self, planes, arch, depth, axes, box_warp=1.0):
        super().__init__()
        self.planes = planes
        self.rendering_kwargs = {"box_warp": box_warp}
        if arch == "panohead":
            self.rendering_kwargs["triplane_depth"] = depth
        self.renderer = _Renderer(axes)

    def mapping(self, z, c, truncation_psi=1):
        return torch.zeros(1, 14, 512)

    def synthesis(self, ws, c, noise_mode="const"):
        return {"feature_planes": self.planes}


def run_arch(arch, only=None):
    sys.path[:0] = [REF, os.path.join(REF, "main"), os.path.join(REF, "PanoHead" if arch == "panohead" else "eg3d")]
    from main.decoder_models.parallel_decoder import ParallelDecoder
    from main.decoder_models.sequential_decoder import SequentialDecoder
    from main.decoder_models.sequential_decoder_reverse import SequentialDecoderReverse
    from training.volumetric_rendering.renderer import generate_planes
    # With use_xyz_embedding=True the reference's persistence decorator refuses to construct
    # the decoder (Embedder holds local closures that do not pickle: "Can't pickle local
    # object 'Embedder.frequency_activation.<locals>.func'").  Only that constructor-time
    # pickle check is disabled here; the forward math is untouched.
    from torch_utils import persistence
    persistence._check_pickleable = lambda obj: None
    classes = dict(ParallelDecoder=ParallelDecoder, SequentialDecoder=SequentialDecoder,
                   SequentialDecoderReverse=SequentialDecoderReverse)

    for name, (a, cls, posenc, P, res, depth, *rest) in CASES.items():
        if a != arch or (only and name not in only):
            continue
        box_warp = rest[0] if rest else 1.0
        H, W = res if isinstance(res, tuple) else (res, res)
        g = torch.Generator().manual_seed(zlib.crc32(name.encode()) % 1000 + 7)
        planes = torch.randn(1, 3, 32 * depth, H, W, generator=g)

        torch.manual_seed(303)
        dec = classes[cls](StubG(planes, arch, depth, generate_planes(), box_warp), hidden_dim=128, use_xyz_embedding=posenc, use_gen_finetune=False, device="cpu")
        if cls == "SequentialDecoderReverse":
            dec.triplane_sr = "None"      # reference quirk q1: attribute read but never set
        with torch.no_grad():
            for p in dec.parameters():
                p.copy_(p.half().float())
        xyz = (torch.rand(P, 3, generator=g) - 0.5) * 1.1 * box_warp     # a few points outside the box
        with torch.no_grad():
            out = dec(torch.zeros(1, 512), torch.zeros(1, 25), xyz, 1.0)
        flat = []
        for h in HEADS:
            sd = getattr(dec, f"{h}_decoder").state_dict()
            for l in (0, 2, 4, 6):
                flat += [sd[f"backbone.{l}.weight"].reshape(-1), sd[f"backbone.{l}.bias"].reshape(-1)]
        flat = torch.cat(flat)
        assert torch.equal(flat, flat.half().float())
        z = dict(arch=arch, variant=VARIANT[cls], pos_freqs=10 if posenc else 0, depth=depth, box_warp=box_warp,
                 planes=planes.numpy(), xyz=xyz.numpy(), flat_fp16=flat.half().numpy())
        for h in HEADS:
            z["out_" + h] = out[h].numpy()
        np.savez_compressed(os.path.join(HERE, f"decoder_{name}.npz"), **z)
        import json
        keys = {k: list(v.shape) for k, v in dec.state_dict().items() if not k.startswith("G.")}
        with open(os.path.join(HERE, f"state_dict_keys_{name}.json"), "w") as f:
            json.dump(keys, f, indent=0)
        print(name, "flat params", flat.numel(), {h: tuple(out[h].shape) for h in HEADS})


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] in ("eg3d", "panohead"):
        run_arch(sys.argv[1], sys.argv[2:])
    else:       # optional arguments: the names of the cases to (re)generate
        for arch in ("eg3d", "panohead"):
            subprocess.check_call([sys.executable, os.path.abspath(__file__), arch] + sys.argv[1:])
