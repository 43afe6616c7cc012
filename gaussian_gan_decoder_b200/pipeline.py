"""Sync-free decode -> rasterise frame loop over pre-allocated buffers.

This is the "N1" caller shape of SURVEY §8(f): what the reference's demo / viewer loop
(main/load_decoder.py:12-25: decoder(...) -> GaussianModel -> CustomCam -> render_simple)
does per frame, but with every buffer allocated once and no host synchronisation inside a
frame.  It talks to ``libggd_b200.so`` directly; results are identical to going through the
decoder classes + ``render_simple`` (tests/test_pipeline_gpu.py).
"""
import ctypes as C

import torch

from . import _lib, rasterizer
from .decoder_models import fused


class FramePipeline:
    def __init__(self, decoder, P, size, device, capacity=None):
        self.lib = _lib.load()
        self.dec = decoder
        self.P, self.W, self.H = int(P), int(size), int(size)
        self.dev = torch.device(device)
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.out = {k: torch.empty(P, n, **f32) for k, n in
                    (("xyz", 3), ("scale", 3), ("rotation", 4), ("opacity", 1), ("color", 3))}
        self.act = {k: torch.empty(P, n, **f32) for k, n in (("scale", 3), ("rotation", 4), ("opacity", 1))}
        self.image = torch.empty(3, self.H, self.W, **f32)
        self.radii = torch.empty(P, dtype=torch.int32, device=self.dev)
        self.planes_cl = None
        self.capacity = None
        self.geom = self.binning = self.img = None
        self.pinned = torch.zeros(4, dtype=torch.int32).pin_memory()
        # stage-2 binning mode of this loop, passed with every call (this pipeline's own: another pipeline or thread
        # cannot change it): starts at the library default; check() moves it to 2 (tiles only: no radix launches)
        # once a frame qualified, and back to 1 if one did not.
        self.bin_mode = self.lib.ggd_raster_binning(-1)
        # a frame loop has no use for the reference-order key/value arrays (ggd.h: GGD_BINNING_LEAN); set False to keep them
        self.lean = True
        self._alloc(capacity or 4 * P)

    def _alloc(self, capacity):
        g, b, i = rasterizer.workspace_sizes(self.P, self.W, self.H, int(capacity))
        u8 = dict(dtype=torch.uint8, device=self.dev)
        if self.geom is None:
            self.geom, self.img = torch.empty(g, **u8), torch.empty(i, **u8)
        self.binning = torch.empty(b, **u8)
        self.capacity = int(capacity)

    def frame(self, planes_nchw, xyz, cam, bg, attr_override=None):
        """Enqueue one frame on the current stream; returns the image tensor (device).  No sync.
        Call ``check()`` afterwards (e.g. once per batch of frames) to validate the capacity.
        ``attr_override``: optional dict of device tensors copied over the decoded attributes before the
        rasteriser runs — keys ``scale`` / ``rotation`` / ``opacity`` (the ACTIVATED values the rasteriser reads),
        ``xyz``, ``color`` — the benchmark's "trained-like splats" leg."""
        lib, p, P = self.lib, _lib.ptr, self.P
        flat = self.dec._ensure_flat(planes_nchw)
        cfg = self.dec._cfg_for(planes_nchw)
        packed = flat.packed_weights()
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        if isinstance(planes_nchw, fused.PackedPlanes):      # producer already emits the kernel's layout
            planes_cl = planes_nchw.data
        else:
            nbytes = planes_nchw.numel() * (2 if cfg.precision == 1 else 4)
            if self.planes_cl is None or self.planes_cl.numel() != nbytes:
                self.planes_cl = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
            planes_cl = self.planes_cl
            _lib.check(lib.ggd_planes_channels_last(s, C.byref(cfg), p(planes_nchw), p(planes_cl)))
        _lib.check(lib.ggd_decode_forward(s, C.byref(cfg), p(planes_cl), p(xyz), P, p(packed),
                                          p(self.out["xyz"]), p(self.out["scale"]), p(self.out["rotation"]),
                                          p(self.out["opacity"]), p(self.out["color"]),
                                          p(self.act["scale"]), p(self.act["rotation"]), p(self.act["opacity"])))
        if attr_override:
            for k, v in attr_override.items():
                (self.act if k in self.act else self.out)[k].copy_(v, non_blocking=True)
        self._raster(s, cam, bg)
        return self.image

    def _raster(self, s, cam, bg):
        import math
        lib, p, P = self.lib, _lib.ptr, self.P
        _lib.check(lib.ggd_raster_forward(
            s, P, 0, 1, p(bg), self.W, self.H, p(self.out["xyz"]), p(self.out["color"]), None,
            p(self.act["opacity"]), p(self.act["scale"]), 1.0, p(self.act["rotation"]), None,
            p(cam.world_view_transform), p(cam.full_proj_transform), p(cam.camera_center),
            math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), 0,
            p(self.geom), p(self.binning), self.capacity, p(self.img), p(self.image), p(self.radii),
            C.c_void_p(self.pinned.data_ptr()), self.bin_mode | (_lib.BINNING_LEAN if self.lean else 0)))

    def capture(self, planes_nchw, xyz, cam, bg):
        """Record one frame (planes -> decode -> rasterise) into a CUDA graph.  The graph reads the given
        tensors in place: refill them (``copy_``) and call ``replay()``.  Run ``frame()`` + ``check()``
        once before capturing so the binning capacity is settled and the weights are packed."""
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self.frame(planes_nchw, xyz, cam, bg)
        return self._graph

    def replay(self):
        self._graph.replay()
        return self.image

    def render_only(self, cam, bg):
        """Another view of the attributes decoded by the last ``frame()`` (decoded attributes depend
        on the latent only): rasterise again with a different camera."""
        self._raster(C.c_void_p(torch.cuda.current_stream().cuda_stream), cam, bg)
        return self.image

    def check(self):
        """Synchronise and verify the last frame: it must fit the binning buffer (grown otherwise) and, in
        binning mode 2, must have qualified for per-tile binning (mode falls back to 1 otherwise).
        Returns (num_rendered, redo): ``redo`` means the last image is invalid, enqueue the frame again."""
        torch.cuda.current_stream().synchronize()
        R = int(self.pinned[0].item()) & 0xffffffff
        over = R > self.capacity
        if over:
            self._alloc(int(R * 1.25) + 4096)
        if self.bin_mode:
            st, info = (C.c_uint32 * 2)(), (C.c_uint32 * 2)()
            s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            _lib.check(self.lib.ggd_raster_status(s, _lib.ptr(self.geom), self.P, st))
            _lib.check(self.lib.ggd_raster_binning_info(s, _lib.ptr(self.geom), self.P, info))
            if st[1] & 8:        # depths too clustered for the chunked long-list sort: this scene needs the radix sort
                self.bin_mode, over = 0, True
            elif st[1] & 2:
                self.bin_mode, over = 1, True
            elif self.bin_mode == 1 and info[0] and not over:
                self.bin_mode = 2
            elif self.bin_mode == 1 and not info[0] and not over:
                self.bin_mode = 0       # tile lists too long for the per-tile path: stop counting per tile
        return R, over
