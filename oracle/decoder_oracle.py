"""TEST INFRASTRUCTURE — CPU restatement of the reference decoder half. NOT the product.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  It restates, with plain torch CPU ops (the same ATen
operators the reference lowers to: grid_sample, Linear, GELU(erf), Softplus):

  sample_from_planes (EG3D)      eg3d/training/volumetric_rendering/renderer.py:23-65
  sample_from_planes (PanoHead)  PanoHead/training/volumetric_rendering/renderer.py:15-58
  Embedder                       main/decoder_utils/pos_encoding.py:4-50
  Decoder                        main/decoder_models/base_decoder.py:8-27
  ParallelDecoder.forward        main/decoder_models/parallel_decoder.py:38-80
  SequentialDecoder.forward      main/decoder_models/sequential_decoder.py:38-86
  SequentialDecoderReverse.fwd   main/decoder_models/sequential_decoder_reverse.py:38-88
  GaussianModel getters          gaussian_splatting/scene/gaussian_model.py:99-121

PARITY STATUS: the reference ships no tests for this path; this restatement is pinned
against outputs of the reference's own Python classes, imported from /root/reference in
the build container by tests/golden/make_decoder_golden.py (fixtures
tests/golden/decoder_*.npz, checked by tests/test_oracle_decoder.py).
"""
import torch
import torch.nn.functional as F

HEADS = ("xyz", "scale", "rotation", "opacity", "color")      # get_params_custom() order
HEAD_OUT = {"xyz": 3, "scale": 3, "rotation": 4, "opacity": 1, "color": 3}
ORDER = {
    "parallel": ("xyz", "scale", "rotation", "opacity", "color"),
    "sequential": ("xyz", "scale", "rotation", "opacity", "color"),
    "sequential_reverse": ("color", "opacity", "rotation", "scale", "xyz"),
}
SCALE_BIAS = {"parallel": -2.0, "sequential": -2.0, "sequential_reverse": -2.5}

PLANE_AXES = {
    "eg3d": torch.tensor([[[1, 0, 0], [0, 1, 0], [0, 0, 1]], [[1, 0, 0], [0, 0, 1], [0, 1, 0]],
                          [[0, 0, 1], [1, 0, 0], [0, 1, 0]]], dtype=torch.float32),
    "panohead": torch.tensor([[[1, 0, 0], [0, 1, 0], [0, 0, 1]], [[1, 0, 0], [0, 0, 1], [0, 1, 0]],
                              [[0, 1, 0], [0, 0, 1], [1, 0, 0]]], dtype=torch.float32),
}


def sample_from_planes(arch, planes, xyz, box_warp, depth=1):
    """planes [1,3,32*depth,H,W], xyz [P,3] -> [3,P,32]"""
    N, n_planes, CD, H, W = planes.shape
    coords = (2 / box_warp) * xyz[None]                                   # [1,P,3]
    inv = torch.linalg.inv(PLANE_AXES[arch].to(xyz.device))
    proj = torch.bmm(coords.expand(n_planes, -1, -1), inv)                # [3,P,3]
    if arch == "eg3d":
        grid = proj[..., :2].unsqueeze(1)                                 # [3,1,P,2]
        out = F.grid_sample(planes.view(n_planes, CD, H, W), grid.float(), mode="bilinear",
                            padding_mode="zeros", align_corners=False)   # [3,C,1,P]
        return out.permute(0, 3, 2, 1).reshape(n_planes, -1, CD)
    C = CD // depth
    grid = proj.unsqueeze(1).unsqueeze(2)                                 # [3,1,1,P,3]
    out = F.grid_sample(planes.view(n_planes, C, depth, H, W), grid.float(), mode="bilinear",
                        padding_mode="zeros", align_corners=False)       # [3,C,1,1,P]
    return out.permute(0, 4, 3, 2, 1).reshape(n_planes, -1, C)


def embed(xyz, num_freqs):
    if not num_freqs:
        return xyz
    outs = [xyz]
    for f in torch.linspace(2.0 ** 0.0, 2.0 ** (num_freqs - 1), steps=num_freqs):
        outs.append(torch.sin(xyz * f))
        outs.append(torch.cos(xyz * f))
    return torch.cat(outs, -1)


def in_dims(variant, pos_freqs):
    pos = 3 + 6 * pos_freqs if pos_freqs else 3
    dims, chained = {}, 0
    for h in ORDER[variant]:
        dims[h] = 32 + pos + (0 if variant == "parallel" else chained)
        chained += HEAD_OUT[h]
    return dims


def flat_layout(variant, pos_freqs, hidden=128):
    """[(head, layer, 'w'|'b', offset, shape)] of the flat parameter buffer (ggd.h layout)."""
    dims, off, out = in_dims(variant, pos_freqs), 0, []
    for h in HEADS:
        for l in range(4):
            i = dims[h] if l == 0 else hidden
            o = HEAD_OUT[h] if l == 3 else hidden
            out.append((h, l, "w", off, (o, i))); off += o * i
            out.append((h, l, "b", off, (o,))); off += o
    return out, off


def mlp(flat, layout, head, x):
    p = {(h, l, k): flat[o:o + int(torch.tensor(s).prod())].view(*s) for h, l, k, o, s in layout}
    for l in range(4):
        x = F.linear(x, p[(head, l, "w")], p[(head, l, "b")])
        if l < 3:
            x = F.gelu(x)
    return x


def decode(arch, variant, pos_freqs, flat, planes, xyz, box_warp=1.0, depth=1, hidden=128):
    """Returns the reference's result dict (raw attributes) plus the activated tensors the
    rasteriser receives.  ``hidden``: the reference's --hidden_dim (flat holds the true, unpadded shapes)."""
    layout, _ = flat_layout(variant, pos_freqs, hidden)
    feat = sample_from_planes(arch, planes, xyz, box_warp, depth).mean(0)   # [P,32]
    info = embed(xyz, pos_freqs)
    res = {}
    for h in ORDER[variant]:
        y = mlp(flat, layout, h, torch.cat([feat, info], -1))
        if h == "xyz":
            y = y * 0.01 + xyz
        elif h == "scale":
            y = -F.softplus(y + 5) + SCALE_BIAS[variant]
        res[h] = y
        if variant != "parallel":
            info = torch.cat([info, y], -1)
    res["act_scale"] = torch.exp(res["scale"])
    res["act_rotation"] = F.normalize(res["rotation"])
    res["act_opacity"] = torch.sigmoid(res["opacity"])
    return res
