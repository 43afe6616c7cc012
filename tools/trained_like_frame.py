#!/usr/bin/env python
"""The benchmark's trained-like frame (500k Gaussians, 512^2, attributes overridden before the rasteriser) run a few
times — a small target for ncu / per-stage timing.   python tools/trained_like_frame.py [frames] [--random-init]"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gaussian_gan_decoder_b200 import _lib, synthetic  # noqa: E402
from gaussian_gan_decoder_b200.pipeline import FramePipeline  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
random_init = "--random-init" in sys.argv      # the headline frame: decoded attributes as they are
dev = torch.device("cuda:0")
lib = _lib.load()
planes, xyz = [t.to(dev) for t in bench.make_inputs(0, dev, n_sets=1)[0]]
dec = bench.make_decoder(planes, dev)
cam = bench.device_camera(dev)
bg = torch.tensor(synthetic.BG[bench.ARCH], device=dev)
tl = synthetic.trained_like_attributes(bench.P_FULL, synthetic.gen(77))
act = synthetic.activated(tl)
override = None if random_init else {"scale": act["scales"].to(dev), "rotation": act["rotations"].to(dev), "opacity": act["opacities"].to(dev),
            "color": tl["color"].to(dev)}
pipe = FramePipeline(dec, bench.P_FULL, bench.SIZE, dev, capacity=16 * bench.P_FULL)
for _ in range(3):
    pipe.frame(planes, xyz, cam, bg, attr_override=override)
    R, redo = pipe.check()
lib.ggd_stage_timing(1)
acc = [0.0] * 8
for _ in range(n):
    pipe.frame(planes, xyz, cam, bg, attr_override=override)
    t = (C.c_float * 8)()
    _lib.check(lib.ggd_stage_times(t))
    acc = [a + max(x, 0.0) for a, x in zip(acc, t)]
lib.ggd_stage_timing(0)
print("R", R, "mode", pipe.bin_mode, {k: round(a / n, 4) for k, a in zip(bench.STAGES, acc)})
