"""Autograd boundary of the fused decoder.

``apply(owner, planes, xyz)`` runs the fused forward kernel and registers a backward that
calls ``ggd_decode_backward`` (csrc/decode_bwd.cu + decode_bwd_dw.cu).  The 40 ``nn.Linear``
parameters are the differentiable inputs (plus the feature planes when the generator is being
fine-tuned); their gradients come back as views of ONE flat buffer in ``get_params_custom()``
order — the message a data-parallel job all-reduces (see ``parallel.py``).
"""
import ctypes as C

import torch

from .. import _lib
from . import fused


class _FusedDecode(torch.autograd.Function):
    @staticmethod
    def forward(ctx, owner, planes, xyz, *params):
        cfg = owner._cfg_for(planes)
        res = fused.decode_forward(cfg, planes, xyz, owner._flat.packed_weights())
        owner._last_activated = {k: res[k] for k in ("act_scale", "act_rotation", "act_opacity")}
        outs = tuple(res[k] for k in ("xyz", "scale", "rotation", "opacity", "color"))
        ctx.owner = owner
        ctx.save_for_backward(planes, xyz, *outs)
        ctx.planes_need_grad = planes.requires_grad
        return outs

    @staticmethod
    def backward(ctx, d_xyz, d_scale, d_rot, d_opacity, d_color):
        owner = ctx.owner
        planes, xyz, *outs = ctx.saved_tensors
        dflat, dplanes = decode_backward(owner, planes, xyz, outs, (d_xyz, d_scale, d_rot, d_opacity, d_color),
                                         ctx.planes_need_grad)
        owner._flat.last_dflat = dflat
        grads = []
        for hi, li, is_b, t in owner._flat._tensors(owner):
            grads.append(owner._flat.view_of(dflat, hi, li, is_b, t))
        return (None, dplanes, None, *grads)


def decode_backward(owner, planes, xyz, outs, douts, want_dplanes):
    """Returns (dflat [flat_count], dplanes or None)."""
    lib = _lib.load()
    dev = xyz.device
    P = xyz.shape[0]
    flat = owner._flat
    cfg32 = owner._cfg_for(planes, precision="fp32")
    p = _lib.ptr
    dflat = torch.empty(flat.count, dtype=torch.float32, device=dev)
    dplanes = torch.empty_like(planes, dtype=torch.float32) if want_dplanes else None
    if P == 0:
        dflat.zero_()
        if dplanes is not None:
            dplanes.zero_()
        return dflat, dplanes
    planes = planes.to(device=dev, dtype=torch.float32).contiguous()
    xyz = xyz.to(torch.float32).contiguous()
    f = lambda t: None if t is None else t.contiguous().float()
    outs = [f(t) for t in outs]
    douts = [f(t) for t in douts]
    with torch.cuda.device(dev):
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        planes_cl = torch.empty(planes.numel() * 4, dtype=torch.uint8, device=dev)
        _lib.check(lib.ggd_planes_channels_last(s, C.byref(cfg32), p(planes), p(planes_cl)))
        packed32 = flat.packed_weights_for(cfg32)
        cfg = owner._cfg_for(planes)            # its precision picks the backward flavour
        planes_cl16 = None
        if cfg.precision == 1:                  # the tensor-core chain recomputes from the planes the fp16 forward sampled
            planes_cl16 = torch.empty(planes.numel() * 2, dtype=torch.uint8, device=dev)
            _lib.check(lib.ggd_planes_channels_last(s, C.byref(cfg), p(planes), p(planes_cl16)))
        nbytes = lib.ggd_decode_backward_scratch_bytes(C.byref(cfg), P)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.ggd_decode_backward(s, C.byref(cfg), p(planes_cl), p(planes_cl16), p(xyz), P, p(flat.flat), p(packed32),
                                           p(outs[0]), p(outs[1]), p(outs[2]), p(outs[3]), p(outs[4]),
                                           p(douts[0]), p(douts[1]), p(douts[2]), p(douts[3]), p(douts[4]),
                                           p(scratch), nbytes, p(dflat), p(dplanes)))
    return dflat, dplanes


def apply(owner, planes, xyz):
    params = [t for _, _, _, t in owner._flat._tensors(owner)]
    return _FusedDecode.apply(owner, planes, xyz, *params)
