"""Generates tests/golden/raster_*.npz by running the UNMODIFIED reference CUDA rasteriser
(oracle/_ref/libggd_ref_cuda.so, built from /root/reference by oracle/Makefile) on seeded
synthetic scenes.  Needs a GPU:  gpurun -- python tests/golden/make_raster_golden.py
(writes into gpurun_out/golden/, copy the files into tests/golden/)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _harness import RefCuda  # noqa: E402
from gaussian_gan_decoder_b200 import synthetic  # noqa: E402

CASES = [("small", 3000, 96, 12.0), ("mid", 12000, 160, 14.0)]


def main():
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    dev = torch.device("cuda:0")
    ref = RefCuda()
    for name, P, size, fov in CASES:
        s = synthetic.raster_scene(P, size=size, fov_deg=fov, seed_offset=P)
        sd = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in s.items()}
        r = ref.forward(sd, dev)
        g = torch.Generator(device="cpu"); g.manual_seed(P)
        dL = torch.randn(3, size, size, generator=g)
        grads = ref.backward(dL.to(dev), dev)
        z = {"in_" + k: v.numpy() for k, v in s.items() if isinstance(v, torch.Tensor)}
        z.update({k: np.asarray(v) for k, v in s.items() if not isinstance(v, torch.Tensor)})
        z["R"] = np.asarray(r["R"])
        for k in ("radii", "depths", "means2D", "conic_opacity", "rgb", "tiles_touched", "n_contrib",
                  "accum_alpha", "out_color", "ranges"):
            z[k] = r[k].cpu().numpy()
        z["keys_xor"] = np.bitwise_xor.reduce(r["point_list_keys"].cpu().numpy()) if r["R"] else np.int64(0)
        z["in_dL_dpix"] = dL.numpy()
        for k, v in grads.items():
            z[k] = v.cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, f"raster_{name}.npz"), **z)
        print(name, "R", r["R"])


if __name__ == "__main__":
    main()
