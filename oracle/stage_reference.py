#!/usr/bin/env python
"""TEST INFRASTRUCTURE — stages the reference's OWN Python files for the hot path into the git-ignored
directory oracle/_ref/pyref/ (same relative layout as /root/reference), so that they can travel to the GPU
box next to oracle/_ref/libggd_ref_cuda.so.  Nothing is edited; nothing staged here is ever committed or
imported by the product package.  Run in the build container (where /root/reference exists):

    python oracle/stage_reference.py            # also run by `make -f oracle/Makefile ref`

What is staged and who uses it:
  * the decoder classes and what they import (main/decoder_models, main/decoder_utils/pos_encoding.py,
    main/dnnlib, main/torch_utils, the two volumetric_rendering/renderer.py + their math helpers)
      -> bench.py --impl reference (CPU, the reference arm's decoder half) and the `ref_cuda` leg (B2);
         tests/test_reference_callers_gpu.py (reference decoder class vs decoder_models)
  * the caller glue kept verbatim by the drop-in boundary: gaussian_renderer/__init__.py,
    scene/{gaussian_model,cameras}.py, utils/{graphics_utils,sh_utils,general_utils,system_utils}.py
      -> tests/test_reference_callers_gpu.py imports the untouched render_simple with THIS repo's
         diff_gaussian_rasterization first on sys.path
  * the reference's Python binding diff_gaussian_rasterization/__init__.py, next to which oracle/Makefile
    builds its _C extension from the reference's C++/CUDA sources (kept in its own directory, _dgr/, so it is
    only importable when a baseline asks for it)
"""
import os
import shutil
import sys

REF = os.environ.get("GGD_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref", "pyref")
DGR = "gaussian_splatting/submodules/diff-gaussian-rasterization"

FILES = [
    "main/decoder_models/base_decoder.py",
    "main/decoder_models/parallel_decoder.py",
    "main/decoder_models/sequential_decoder.py",
    "main/decoder_models/sequential_decoder_reverse.py",
    "main/decoder_utils/pos_encoding.py",
    "main/dnnlib/__init__.py",
    "main/dnnlib/util.py",
    "main/torch_utils/__init__.py",
    "main/torch_utils/persistence.py",
    "main/torch_utils/misc.py",
    "eg3d/training/volumetric_rendering/renderer.py",
    "eg3d/training/volumetric_rendering/ray_marcher.py",
    "eg3d/training/volumetric_rendering/math_utils.py",
    "PanoHead/training/volumetric_rendering/renderer.py",
    "PanoHead/training/volumetric_rendering/ray_marcher.py",
    "PanoHead/training/volumetric_rendering/math_utils.py",
    "gaussian_splatting/gaussian_renderer/__init__.py",
    "gaussian_splatting/scene/gaussian_model.py",
    "gaussian_splatting/scene/cameras.py",
    "gaussian_splatting/utils/graphics_utils.py",
    "gaussian_splatting/utils/sh_utils.py",
    "gaussian_splatting/utils/general_utils.py",
    "gaussian_splatting/utils/system_utils.py",
    "main/camera_utils.py",
]
RENAMED = {DGR + "/diff_gaussian_rasterization/__init__.py": "_dgr/diff_gaussian_rasterization/__init__.py"}


def main():
    if not os.path.isdir(REF):
        print(f"reference tree {REF} absent: keeping what is already staged in {DST}")
        return 0
    n = 0
    for rel, dst_rel in [(f, f) for f in FILES] + list(RENAMED.items()):
        src, dst = os.path.join(REF, rel), os.path.join(DST, dst_rel)
        if not os.path.exists(src):
            print("missing in the reference:", rel, file=sys.stderr)
            return 1
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        n += 1
    print(f"staged {n} reference files into {DST}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
