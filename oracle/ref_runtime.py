"""TEST INFRASTRUCTURE — run the reference's OWN Python code for the path from oracle/_ref/pyref/.

oracle/_ref/pyref/ is staged from /root/reference by oracle/stage_reference.py (git-ignored; it travels to the
GPU box like the prebuilt .so files).  Only tests/, bench.py's baseline legs (`--impl reference`, `ref_cuda`)
and __graft_entry__ may import this module; the product package never does.

The reference's files are imported UNCHANGED.  What this module adds around them is what any harness for that
code needs (SURVEY 8b "import-time baggage"):
  * sys.path entries in the reference's own layout: <pyref>, <pyref>/main, <pyref>/gaussian_splatting and ONE of
    <pyref>/eg3d | <pyref>/PanoHead (main/train_pano2gaussian_decoder.py:127,146,159-162 picks it the same way);
  * stub modules for third-party imports that are absent from this image and unused on the path:
    ``plyfile`` (gaussian_model.py:20) and ``simple_knn._C`` (gaussian_model.py:22);
  * a stand-in generator G (mapping -> zeros, synthesis -> given feature planes): EG3D / PanoHead themselves are
    upstream of the path and out of scope;
  * ``SequentialDecoderReverse.triplane_sr``: the reference reads this attribute in forward() but never sets it
    (quirk q1); its training script sets nothing either, the released checkpoints carry it.  Set to "None".
"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
PYREF = os.path.join(HERE, "_ref", "pyref")
DGR_DIR = os.path.join(PYREF, "_dgr")


def available(with_rasterizer=False):
    ok = os.path.exists(os.path.join(PYREF, "main", "decoder_models", "parallel_decoder.py"))
    if with_rasterizer:
        ok = ok and os.path.exists(os.path.join(DGR_DIR, "diff_gaussian_rasterization", "_C.so"))
    return ok


def _stub(name, **attrs):
    if name not in sys.modules:
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
    return sys.modules[name]


def install_stubs():
    """Absent third-party modules the kept callers import at module load and never use on the path."""
    _stub("plyfile", PlyData=None, PlyElement=None)
    pkg = _stub("simple_knn")
    pkg._C = _stub("simple_knn._C", distCUDA2=None)


def add_paths(arch, reference_rasterizer=False, first=()):
    """Put the staged reference tree on sys.path.  ``first``: directories that must win (the drop-in test puts
    THIS repo's root there so ``import diff_gaussian_rasterization`` resolves to the product);
    ``reference_rasterizer``: make the reference's own binding importable instead (baselines only)."""
    gen = "PanoHead" if arch == "panohead" else "eg3d"
    paths = list(first)
    if reference_rasterizer:
        paths.append(DGR_DIR)
    paths += [PYREF, os.path.join(PYREF, "main"), os.path.join(PYREF, "gaussian_splatting"), os.path.join(PYREF, gen)]
    for p in reversed(paths):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    install_stubs()


class _Renderer:
    def __init__(self, axes):
        self.plane_axes = axes


class StubG(torch.nn.Module):
    """Module-level on purpose: the reference's persistent_class pickles constructor arguments."""

    def __init__(self, planes, arch, depth, axes):
        super().__init__()
        self.planes = planes
        self.rendering_kwargs = {"box_warp": 1.0}
        if arch == "panohead":
            self.rendering_kwargs["triplane_depth"] = depth
        self.renderer = _Renderer(axes)

    def mapping(self, z, c, truncation_psi=1):
        return torch.zeros(1, 14, 512, device=z.device)

    def synthesis(self, ws, c, noise_mode="const"):
        return {"feature_planes": self.planes}


VARIANT_CLASS = {"parallel": "ParallelDecoder", "sequential": "SequentialDecoder",
                 "sequential_reverse": "SequentialDecoderReverse"}


def make_decoder(variant, arch, depth, planes, device, use_xyz_embedding=False, seed=303, hidden_dim=128):
    """An instance of the reference's decoder class (call ``add_paths(arch, ...)`` first).  Default nn.Linear
    initialisation under ``seed``; NOTE the reference constructs its five heads in its own order, so the random
    stream differs from the product classes — copy parameters with ``load_flat`` for like-for-like runs."""
    import importlib
    mod = importlib.import_module({"parallel": "main.decoder_models.parallel_decoder",
                                   "sequential": "main.decoder_models.sequential_decoder",
                                   "sequential_reverse": "main.decoder_models.sequential_decoder_reverse"}[variant])
    from training.volumetric_rendering.renderer import generate_planes
    if use_xyz_embedding:
        # the reference's persistence decorator refuses to construct a decoder whose Embedder holds local closures
        # ("Can't pickle local object"); only that constructor-time check is switched off
        from torch_utils import persistence
        persistence._check_pickleable = lambda obj: None
    torch.manual_seed(seed)
    G = StubG(planes, arch, depth, generate_planes().to(device))
    dec = getattr(mod, VARIANT_CLASS[variant])(G, hidden_dim=hidden_dim, use_xyz_embedding=use_xyz_embedding,
                                              use_gen_finetune=False, device=device)
    dec.triplane_sr = "None"
    return dec


HEADS = ("xyz", "scale", "rotation", "opacity", "color")      # get_params_custom() order


def load_flat(dec, flat):
    """Copy a flat parameter vector (get_params_custom order: per head W0,b0,..,W3,b3) into a reference decoder."""
    off = 0
    with torch.no_grad():
        for h in HEADS:
            for p in getattr(dec, f"{h}_decoder").parameters():
                n = p.numel()
                p.copy_(flat[off:off + n].view_as(p).to(p.device))
                off += n
    assert off == flat.numel(), (off, flat.numel())
    return dec


def gaussian_model_from(decoded):
    """The assignment block of the reference's loops (main/train_pano2gaussian_decoder.py:223-227,
    main/load_decoder.py:14-19) with the reference's own GaussianModel."""
    from scene.gaussian_model import GaussianModel
    g = GaussianModel(0)
    g._xyz = decoded.xyz
    g._scaling = decoded.scale
    g._rotation = decoded.rotation
    g._opacity = decoded.opacity
    g._features_dc = decoded.color.unsqueeze(1)
    # get_features concatenates _features_dc with _features_rest, which the constructor leaves as an empty CPU tensor;
    # keep it empty but on the attributes' device so the cat does not depend on this torch version's device checks
    g._features_rest = torch.empty(0, device=decoded.color.device)
    return g
