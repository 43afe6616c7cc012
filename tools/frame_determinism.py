#!/usr/bin/env python
"""Stress check: the whole frame (repack -> decode -> rasterise) must return the same bits every time.
python tools/frame_determinism.py [n] [--trained-like]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gaussian_gan_decoder_b200 import synthetic  # noqa: E402
from gaussian_gan_decoder_b200.pipeline import FramePipeline  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 300
dev = torch.device("cuda:0")
sets = [[t.to(dev) for t in s] for s in bench.make_inputs(0, dev, n_sets=3)]
dec = bench.make_decoder(sets[0][0], dev)
cam = bench.device_camera(dev)
bg = torch.tensor(synthetic.BG[bench.ARCH], device=dev)
override = None
if "--trained-like" in sys.argv:
    tl = synthetic.trained_like_attributes(bench.P_FULL, synthetic.gen(77))
    act = synthetic.activated(tl)
    override = {"scale": act["scales"].to(dev), "rotation": act["rotations"].to(dev), "opacity": act["opacities"].to(dev),
                "color": tl["color"].to(dev)}
pipe = FramePipeline(dec, bench.P_FULL, bench.SIZE, dev, capacity=16 * bench.P_FULL)
ref = {}
bad = 0
for i in range(n + 9):
    k = i % 3
    pipe.frame(*sets[k], cam, bg, attr_override=override)
    R, redo = pipe.check()
    if i < 6:
        continue                      # binning mode settles (1 -> 2) in the first frames
    cur = (R, pipe.image.clone(), pipe.out["xyz"].clone(), pipe.radii.clone())
    if k not in ref:
        ref[k] = cur
        continue
    names = ("R", "image", "xyz", "radii")
    for name, a, b in zip(names, cur, ref[k]):
        same = (a == b) if name == "R" else torch.equal(a, b)
        if not same:
            bad += 1
            print("frame", i, "set", k, name, "differs", (a, b) if name == "R" else int((a != b).sum()))
print("frames", n, "mismatches", bad)
