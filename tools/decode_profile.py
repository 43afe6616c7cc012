#!/usr/bin/env python
"""Cycle accounting of decode_tc_kernel's roles (GGD_DECODE_PROF=1 instantiation) at bench size, plus the plain
kernel's time.  Diagnostics only:  python tools/decode_profile.py"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gaussian_gan_decoder_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.load()
planes, xyz = [t.to(dev) for t in bench.make_inputs(0, dev, n_sets=1)[0]]
dec = bench.make_decoder(planes, dev)
pk = dec.pack_planes(planes)


def run(n):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.no_grad():
        for _ in range(3):
            dec.decode(pk, xyz)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            dec.decode(pk, xyz)
        e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print("decode (class API incl. output allocs), ms:", round(run(20), 4))
os.environ["GGD_DECODE_PROF"] = "1"
print("instrumented, ms:", round(run(5), 4))
out = (C.c_uint64 * 12)()
_lib.check(lib.ggd_debug_decode_profile(out))
names = ("kernel", "sampling", "wait accumulator", "hidden epilogues", "last-layer epilogues", "mma: wait operand",
         "mma: wait weights", "mma: issue", "  of which layer 0", "  of which last layer")
tot = out[0]
for n, v in zip(names, out):
    print(f"{n:24s} {v:10d} clk  {100.0 * v / max(tot, 1):5.1f} %")
tiles = -(-bench.P_FULL // 128) / 148
steps = tiles * 5
print("MMA issue clk per step: layer0 %.0f, hidden %.0f, last %.0f" % (out[8] / steps, (out[7] - out[8] - out[9]) / (2 * steps), out[9] / steps))
print("tiles per SM %.1f; clk per tile %.0f; per hidden tile-layer epilogue %.0f; per last-layer %.0f" % (
    tiles, tot / tiles, out[3] / (tiles * 15), out[4] / (tiles * 5)))
