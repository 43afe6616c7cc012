#!/usr/bin/env python
"""B2 baseline of SURVEY.md section 8(d): the reference's own GPU path timed on the same B200.

    python tests/reference_cuda_timing.py [--steps 20]

* decoder: the torch restatement of the reference decoder classes (oracle/decoder_oracle.py: grid_sample,
  nn.Linear / cuBLAS SGEMM with TF32 off, erf-GELU, softplus ...) on ``cuda`` in fp32 - the same ATen ops the
  reference modules launch (main/decoder_models/*.py);
* rasteriser: the UNMODIFIED reference CUDA rasteriser (oracle/_ref/libggd_ref_cuda.so, built from
  /root/reference/diff-gaussian-rasterization by oracle/Makefile), including its per-frame D2H sync.

Same workload as bench.py (configs[1]).  This is test infrastructure (it drives the checkers under oracle/);
it prints one JSON line that DESIGN.md / BASELINE.md quote.  Not collected by pytest.
"""
import argparse
import json
import math
import os
import sys
import time

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import bench  # noqa: E402  (workload constants and input construction)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    import decoder_oracle as do
    from _harness import RefCuda, ref_cuda_available
    from gaussian_gan_decoder_b200 import synthetic
    assert ref_cuda_available(), "needs a GPU and oracle/_ref/libggd_ref_cuda.so"
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sets = [(pl.to(dev), xyz.to(dev)) for pl, xyz in bench.make_inputs(0, dev)]
    torch.manual_seed(303)
    flat = torch.cat([p.detach().reshape(-1) for p in bench.make_decoder(sets[0][0].cpu(), "cpu").get_params_custom()]).to(dev)
    cam = bench.device_camera(dev)
    bg = torch.tensor(synthetic.BG[bench.ARCH], device=dev)
    ref = RefCuda()

    def decode(i):
        planes, xyz = sets[i % len(sets)]
        with torch.no_grad():
            return do.decode(bench.ARCH, bench.VARIANT, 0, flat, planes, xyz, 1.0, bench.DEPTH)

    def scene_of(r):
        return dict(means3D=r["xyz"], scales=r["act_scale"], rotations=r["act_rotation"], opacities=r["act_opacity"],
                    shs=r["color"][:, None, :].contiguous(), viewmatrix=cam.world_view_transform,
                    projmatrix=cam.full_proj_transform, campos=cam.camera_center, bg=bg, W=bench.SIZE, H=bench.SIZE,
                    tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), sh_degree=0, scale_modifier=1.0)

    def raster_only(scene):
        """The shim call alone (RefCuda.forward also copies every workspace field out, which is not the reference's work)."""
        t = {k: (scene[k].contiguous().float()) for k in ("bg", "means3D", "shs", "opacities", "scales", "rotations",
                                                          "viewmatrix", "projmatrix", "campos")}
        P = t["means3D"].shape[0]
        out_color = torch.empty(3, bench.SIZE, bench.SIZE, device=dev)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        import ctypes as C
        g, b, im = C.c_void_p(), C.c_void_p(), C.c_void_p()
        d = ref._dp
        return lambda: ref.lib.ggd_ref_forward(P, 0, 1, d(t["bg"]), bench.SIZE, bench.SIZE, d(t["means3D"]), d(t["shs"]), None,
                                               d(t["opacities"]), d(t["scales"]), 1.0, d(t["rotations"]), None,
                                               d(t["viewmatrix"]), d(t["projmatrix"]), d(t["campos"]),
                                               float(scene["tanfovx"]), float(scene["tanfovy"]), 0, d(out_color), d(radii), 0,
                                               C.byref(g), C.byref(b), C.byref(im))

    # decoder
    for i in range(args.warmup):
        r = decode(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        r = decode(i)
    e1.record()
    torch.cuda.synchronize()
    dec_ms = e0.elapsed_time(e1) / args.steps
    # rasteriser (wall clock around the call: the reference synchronises inside for num_rendered)
    call = raster_only(scene_of(r))
    for _ in range(args.warmup):
        R = call()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        R = call()
    torch.cuda.synchronize()
    ras_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    total = dec_ms + ras_ms
    print(json.dumps({"impl": "reference-cuda (B2)", "metric": bench.METRIC, "value": round(1e3 / total, 2), "unit": "frames/s",
                      "ms_per_step": round(total, 3), "decode_ms": round(dec_ms, 3), "raster_ms": round(ras_ms, 3),
                      "num_rendered": int(R), "steps": args.steps, "warmup": args.warmup, "dtype": "f32 (TF32 off)",
                      "gpu": torch.cuda.get_device_name(0),
                      "note": "torch fp32 restatement of the reference decoder classes on cuda + unmodified reference CUDA rasteriser"}))


if __name__ == "__main__":
    main()
