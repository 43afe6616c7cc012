"""Host side of the fused sample+decode kernel: flat parameter storage, config
marshalling and the autograd boundary.  Everything numeric happens in
``libggd_b200.so`` (ggd_decode_* in include/ggd.h).
"""
import ctypes as C
import os

import torch

from .. import _lib

HEADS = ("xyz", "scale", "rotation", "opacity", "color")     # get_params_custom() order
VARIANTS = {"parallel": 0, "sequential": 1, "sequential_reverse": 2}
ARCHS = {"eg3d": 0, "panohead": 1}


DecoderConfig = _lib.DecoderConfig


PRECISIONS = {"fp32": 0, "fp16": 1}
# "fp16": tcgen05 tensor-core kernel (fp16 operands, fp32 accumulation) — the production path.
# "fp32": exact CUDA-core kernel — the accuracy anchor.  Overridable per decoder (``.precision``).
DEFAULT_PRECISION = os.environ.get("GGD_DECODE_PRECISION", "fp16")
# Width the kernels are specialised for (tile N = 128, TMEM column budget).  A narrower decoder (--hidden_dim < 128 in the
# reference's training script) runs on the same kernels with its weights zero-padded into the 128-wide layout: the padded
# hidden units compute gelu(0) = 0 — exactly, in the erf, tanh and polynomial forms alike — and feed zero weights, so the
# function and every gradient of a real parameter are those of the narrow network; gradients of padded entries are 0.
KERNEL_HIDDEN = 128


def make_config(arch, variant, depth, h, w, pos_freqs, hidden, box_warp, scale_bias, precision="fp32"):
    return DecoderConfig(ARCHS[arch], VARIANTS[variant], int(depth), int(h), int(w), int(pos_freqs), int(hidden),
                         float(box_warp), float(scale_bias), PRECISIONS[precision])


def param_counts(cfg):
    lib = _lib.load()
    a, b = C.c_size_t(), C.c_size_t()
    _lib.check(lib.ggd_decode_param_counts(C.byref(cfg), C.byref(a), C.byref(b)))
    return a.value, b.value


def param_offset(cfg, head, layer, is_bias):
    lib = _lib.load()
    off, fi, fo = C.c_size_t(), C.c_int(), C.c_int()
    _lib.check(lib.ggd_decode_param_offset(C.byref(cfg), head, layer, int(is_bias), C.byref(off), C.byref(fi), C.byref(fo)))
    return off.value, fi.value, fo.value


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class FlatParams:
    """One contiguous fp32 buffer holding all 40 decoder tensors in the ggd.h layout; the
    nn.Linear parameters are re-pointed to views of it, so the optimizer updates the buffer
    in place and a data-parallel job all-reduces ONE gradient buffer."""

    def __init__(self, owner, cfg):
        self.cfg = cfg
        self.count, self.packed_count = param_counts(cfg)
        self.flat = None
        self.flat_grad = None
        self.packed = None
        self._packed_version = None
        self.adopt(owner)

    def set_config(self, cfg):
        """Switch kernel flavour (precision): same flat parameters, different packed layout."""
        self.cfg = cfg
        _, self.packed_count = param_counts(cfg)
        self.packed = None
        self._packed_version = None

    def _tensors(self, owner):
        for hi, h in enumerate(HEADS):
            for li, lin in enumerate(getattr(owner, f"{h}_decoder").linears()):
                yield hi, li, 0, lin.weight
                yield hi, li, 1, lin.bias

    def view_of(self, buf, hi, li, is_b, t):
        """The part of ``buf`` (a flat buffer in the ggd.h layout) that belongs to parameter ``t``: the whole block for a
        128-wide decoder, its leading rows / columns — a strided view — for a narrower one (see KERNEL_HIDDEN)."""
        off, fi, fo = param_offset(self.cfg, hi, li, is_b)
        if is_b:
            return buf[off:off + fo][:t.shape[0]]
        return buf[off:off + fo * fi].view(fo, fi)[:t.shape[0], :t.shape[1]]

    def adopt(self, owner):
        first = next(self._tensors(owner))[3]
        dev = first.device
        hidden = int(getattr(owner, "hidden_dim", KERNEL_HIDDEN))
        flat = torch.zeros(self.count, dtype=torch.float32, device=dev)      # zeros: the padding of a narrow decoder
        grad = torch.zeros(self.count, dtype=torch.float32, device=dev)
        narrow = lambda d: hidden if d == KERNEL_HIDDEN else d
        with torch.no_grad():
            for hi, li, is_b, t in self._tensors(owner):
                off, fi, fo = param_offset(self.cfg, hi, li, is_b)
                shape = (narrow(fo),) if is_b else (narrow(fo), narrow(fi))
                if tuple(t.shape) != shape:
                    raise ValueError(f"{HEADS[hi]}_decoder layer {li}: expected {shape}, got {tuple(t.shape)}")
                view = self.view_of(flat, hi, li, is_b, t)
                view.copy_(t.detach().to(torch.float32))
                t.data = view
                t.grad = self.view_of(grad, hi, li, is_b, t)
        self.flat, self.flat_grad = flat, grad
        self._params = [t for _, _, _, t in self._tensors(owner)]
        self._packed_version = None

    def invalidate(self):
        """Force the packed kernel weights to be rebuilt on the next decode.  The cache follows the parameters' autograd
        version counters, which every in-place op on the parameter bumps (``optimizer.step()``, ``copy_``,
        ``load_state_dict``); writes that bypass them — ``p.data.add_(...)``, ``flat.flat[...] = ...``, a raw CUDA copy —
        do not, so call this after such a write (``decoder.invalidate_packed_weights()``)."""
        self._packed_version = None
        self._alt_packed = None

    def in_sync(self, owner):
        first = next(self._tensors(owner))[3]
        return self.flat is not None and first.data_ptr() == self.flat.data_ptr() and first.device == self.flat.device

    def packed_weights_for(self, cfg):
        """Packed copy for another kernel flavour (the backward always uses the fp32 pack)."""
        if cfg.precision == self.cfg.precision:
            return self.packed_weights()
        ver = sum(t._version for t in self._params)
        cache = getattr(self, "_alt_packed", None)
        if cache is None or cache[0] != (cfg.precision, ver) or cache[1].device != self.flat.device:
            lib = _lib.load()
            _, n = param_counts(cfg)
            buf = torch.empty(n, dtype=torch.float32, device=self.flat.device)
            with torch.cuda.device(self.flat.device):
                _lib.check(lib.ggd_decode_pack_weights(_stream(), C.byref(cfg), _lib.ptr(self.flat), _lib.ptr(buf)))
            self._alt_packed = ((cfg.precision, ver), buf)
        return self._alt_packed[1]

    def packed_weights(self):
        """Kernel-layout copy of the parameters, rebuilt only when the flat buffer changed."""
        # parameters are views of the flat buffer but carry their own version counters
        ver = sum(t._version for t in self._params)
        if self.packed is None or self.packed.device != self.flat.device or self._packed_version != ver:
            lib = _lib.load()
            if self.packed is None or self.packed.device != self.flat.device:
                self.packed = torch.empty(self.packed_count, dtype=torch.float32, device=self.flat.device)
            with torch.cuda.device(self.flat.device):
                _lib.check(lib.ggd_decode_pack_weights(_stream(), C.byref(self.cfg), _lib.ptr(self.flat), _lib.ptr(self.packed)))
            self._packed_version = ver
        return self.packed


class PackedPlanes:
    """Feature planes already in the layout the tensor-core decoder reads: fp16, channels-last
    ``[3][D][H][W][32]`` (one texel = one 64-byte half-line).  ``shape`` is the generator's NCHW shape
    ``(1, 3, 32*D, H, W)`` so the decoder classes can treat it like the tensor it stands for.

    This is the plane-producer handshake of SURVEY 8(f) N3: a generator (or a cache of per-identity
    planes in host memory) that emits this layout saves the per-latent repack kernel and half of the
    upload.  Inference only - the backward needs the fp32 planes."""
    requires_grad = False

    def __init__(self, data, shape):
        self.data, self.shape = data, tuple(shape)

    def dim(self):
        return len(self.shape)

    @classmethod
    def from_host(cls, planes_nchw, pin=True):
        """The producer for planes that live on the HOST (a cache of per-identity planes, a data loader): fp32
        ``[1, 3, 32*D, H, W]`` in the generator's layout -> the kernel's layout, built with torch CPU ops, same bytes as
        ``ggd_planes_channels_last`` writes on the device (round-to-nearest-even fp16, channel index c*D + d as in
        PanoHead/training/volumetric_rendering/renderer.py:51-52).  Upload it with ``.to(device, non_blocking=True)``:
        half the bytes of the fp32 planes and no repack kernel per frame."""
        if planes_nchw.dim() != 5 or planes_nchw.shape[0] != 1 or planes_nchw.shape[1] != 3 or planes_nchw.shape[2] % 32:
            raise ValueError(f"feature_planes must be [1,3,32*D,H,W], got {tuple(planes_nchw.shape)}")
        _, _, cd, h, w = planes_nchw.shape
        d = cd // 32
        x = planes_nchw.detach().to("cpu", torch.float32).view(3, 32, d, h, w)          # channel = c * D + d
        packed = x.permute(0, 2, 3, 4, 1).contiguous().to(torch.float16)                    # [3][D][H][W][32]
        data = packed.view(torch.uint8).reshape(-1)
        if pin and torch.cuda.is_available():
            data = data.pin_memory()
        return cls(data, planes_nchw.shape)

    def to(self, device, non_blocking=False):
        return PackedPlanes(self.data.to(device, non_blocking=non_blocking), self.shape)

    def pin_memory(self):
        return PackedPlanes(self.data.pin_memory(), self.shape)


def pack_planes(cfg, planes_nchw):
    """fp32 NCHW planes (device) -> PackedPlanes; cfg must be an fp16-precision config."""
    lib = _lib.load()
    if cfg.precision != 1:
        raise ValueError("packed planes are the fp16 tensor-core path's layout")
    planes_nchw = planes_nchw.to(torch.float32).contiguous()
    data = torch.empty(planes_nchw.numel() * 2, dtype=torch.uint8, device=planes_nchw.device)
    with torch.cuda.device(planes_nchw.device):
        _lib.check(lib.ggd_planes_channels_last(_stream(), C.byref(cfg), _lib.ptr(planes_nchw), _lib.ptr(data)))
    return PackedPlanes(data, planes_nchw.shape)


def decode_forward(cfg, planes_nchw, xyz, packed, with_activations=True):
    """planes [1,3,32*D,H,W] fp32 (generator layout) or PackedPlanes, xyz [P,3] -> dict of raw attributes
    (+ activated scale/rotation/opacity for the rasteriser)."""
    lib = _lib.load()
    if not xyz.is_cuda:
        raise RuntimeError("init_position must be a CUDA tensor: the decoder has no CPU path")
    dev = xyz.device
    P = xyz.shape[0]
    if planes_nchw.dim() != 5 or planes_nchw.shape[0] != 1 or planes_nchw.shape[1] != 3:
        raise ValueError(f"feature_planes must be [1,3,C*D,H,W], got {tuple(planes_nchw.shape)}")
    prepacked = isinstance(planes_nchw, PackedPlanes)
    if prepacked:
        if cfg.precision != 1:
            raise ValueError("PackedPlanes need the fp16 decoder (decoder.precision = 'fp16')")
        if not planes_nchw.data.is_cuda:
            raise RuntimeError("PackedPlanes must be on the GPU: the decoder has no CPU path")
    else:
        planes_nchw = planes_nchw.to(device=dev, dtype=torch.float32).contiguous()
    xyz = xyz.to(torch.float32).contiguous()
    new = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
    out = dict(xyz=new(P, 3), scale=new(P, 3), rotation=new(P, 4), opacity=new(P, 1), color=new(P, 3))
    act = dict(act_scale=new(P, 3), act_rotation=new(P, 4), act_opacity=new(P, 1)) if with_activations else {}
    if P == 0:
        return {**out, **act}
    p = _lib.ptr
    with torch.cuda.device(dev):
        s = _stream()
        if prepacked:
            planes_cl = planes_nchw.data
        else:
            planes_cl = torch.empty(planes_nchw.numel() * (2 if cfg.precision == 1 else 4), dtype=torch.uint8, device=dev)
            _lib.check(lib.ggd_planes_channels_last(s, C.byref(cfg), p(planes_nchw), p(planes_cl)))
        _lib.check(lib.ggd_decode_forward(s, C.byref(cfg), p(planes_cl), p(xyz), P, p(packed),
                                          p(out["xyz"]), p(out["scale"]), p(out["rotation"]), p(out["opacity"]),
                                          p(out["color"]), p(act.get("act_scale")), p(act.get("act_rotation")),
                                          p(act.get("act_opacity"))))
    return {**out, **act}


def single_head_forward(decoder, triplane_features, gaussian_features):
    from . import fused_single
    return fused_single.forward(decoder, triplane_features, gaussian_features)
