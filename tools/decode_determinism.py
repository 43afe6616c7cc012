#!/usr/bin/env python
"""Stress check: the fp16 decoder must return the same bits on every call (same inputs).  python tools/decode_determinism.py [n]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda:0")
sets = [[t.to(dev) for t in s] for s in bench.make_inputs(0, dev, n_sets=2)]
dec = bench.make_decoder(sets[0][0], dev)
ref = []
with torch.no_grad():
    for planes, xyz in sets:
        out = dec.decode(planes, xyz)
        ref.append({k: v.clone() for k, v in out.items()})
    bad = 0
    for i in range(n):
        planes, xyz = sets[i % 2]
        out = dec.decode(planes, xyz)
        for k, v in out.items():
            if not torch.equal(v, ref[i % 2][k]):
                bad += 1
                d = (v != ref[i % 2][k]).nonzero()
                print("call", i, k, "differs at", d[:3].tolist(), "of", d.shape[0])
torch.cuda.synchronize()
print("calls", n, "mismatching tensors", bad)
