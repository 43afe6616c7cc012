"""One decoder training step on the fused path (BASELINE configs[3]).

The shape of the reference's loop body (main/train_pano2gaussian_decoder.py:219-265): decode,
copy the five results into a GaussianModel, render, loss, backward, optimiser step — with the
gradient all-reduce of ``data_parallel.py`` between backward and step when several GPUs train
data-parallel.  The reference's extra losses (LPIPS / ID / Sobel) need pretrained networks and
are out of scope; L1 + L2 on the image are used (gaussian_splatting/utils/loss_utils.py:17-24).
"""
import torch

from .data_parallel import allreduce_flat, allreduce_gradients
from .gaussian_renderer import GaussianModel, render_simple


def l1_loss(a, b):
    return torch.abs(a - b).mean()


def l2_loss(a, b):
    return ((a - b) ** 2).mean()


def training_step(decoder, optimizer, z, gan_camera_params, init_position, camera, target, bg,
                  truncation_psi=1.0, group=None, dL_dimage=None):
    """Returns the loss tensor (device).  ``dL_dimage``: if given, backpropagate that image
    gradient instead of the L1+L2 loss (benchmark mode of SURVEY §8d, config 4)."""
    result = decoder(z, gan_camera_params, init_position, truncation_psi)
    gaussians = GaussianModel.from_decoder(result)        # activations through autograd
    image = render_simple(camera, gaussians, bg)["render"]
    if dL_dimage is not None:
        loss = (image * dL_dimage).sum()
    else:
        loss = l1_loss(image, target) + l2_loss(image, target)
    loss.backward()
    extra = list(decoder.G.parameters()) if decoder.use_gen_finetune else ()
    allreduce_gradients(decoder, extra_params=extra, group=group)
    optimizer.step()
    # Dropping the .grad tensors costs nothing; the next backward hands out fresh views of ONE flat gradient
    # buffer (fused_backward.py), so neither 40 fill kernels nor 40 accumulate kernels run per step.
    optimizer.zero_grad(set_to_none=True)
    return loss.detach()


class TrainPipeline:
    """The training-loop body of the reference (main/train_pano2gaussian_decoder.py:217-265) as a sync-free caller
    of the C ABI: every forward and backward buffer is allocated once, the binning capacity comes
    from earlier steps, nothing is read back inside a step, the decoder gradient lands directly in the flat buffer
    the ``.grad`` tensors are views of, and the data-parallel all-reduce of that buffer runs on a side stream that
    the main stream joins right before ``optimizer.step()``.

        tp = TrainPipeline(decoder, torch.optim.Adam(decoder.get_params_custom(), lr=...), P, 512, device)
        for ...:
            tp.step(feature_planes, init_position, camera, bg, target=image)     # enqueue only
            if it % 50 == 0: tp.check()                                            # capacity / loss read-back

    Same numbers as the autograd route ``training_step`` (tests/test_training_gpu.py); autograd is not involved:
    decode fwd -> raster fwd -> dL/dimage -> raster bwd -> activation bwd -> decode bwd -> all-reduce -> step.
    """

    def __init__(self, decoder, optimizer, P, size, device, capacity=None, group=None):
        import ctypes as C
        from . import _lib, rasterizer
        self.C, self._lib, self.lib = C, _lib, _lib.load()
        self.dec, self.opt, self.group = decoder, optimizer, group
        self.P, self.W, self.H = int(P), int(size), int(size)
        self.dev = torch.device(device)
        f32 = dict(dtype=torch.float32, device=self.dev)
        P = self.P
        self.out = {k: torch.empty(P, n, **f32) for k, n in (("xyz", 3), ("scale", 3), ("rotation", 4), ("opacity", 1), ("color", 3))}
        self.act = {k: torch.empty(P, n, **f32) for k, n in (("scale", 3), ("rotation", 4), ("opacity", 1))}
        self.image = torch.empty(3, self.H, self.W, **f32)
        self.dimage = torch.empty(3, self.H, self.W, **f32)
        self.radii = torch.empty(P, dtype=torch.int32, device=self.dev)
        # one block for every rasteriser gradient (ggd_raster_backward wants them zero-filled: one memset per step)
        names = (("mean2D", 3), ("conic", 4), ("opacity", 1), ("color", 3), ("mean3D", 3), ("cov3D", 6), ("sh", 3),
                 ("scale", 3), ("rot", 4))
        self._gblock = torch.empty(P * sum(n for _, n in names), **f32)
        self.g, off = {}, 0
        for k, n in names:
            self.g[k] = self._gblock[off:off + P * n].view(P, n)
            off += P * n
        self.loss = torch.zeros((), **f32)
        self.pinned = torch.zeros(4, dtype=torch.int32).pin_memory()
        self.bin_mode = self.lib.ggd_raster_binning(-1)
        self.geom = self.binning = self.img = None
        self._ws = rasterizer.workspace_sizes
        self._alloc(capacity or 4 * P)
        self.planes16 = self.planes32 = self.scratch = None
        self.side = torch.cuda.Stream(device=self.dev)
        self._ar_events = None
        self._ar_bytes = 0

    def _alloc(self, capacity):
        g, b, i = self._ws(self.P, self.W, self.H, int(capacity))
        u8 = dict(dtype=torch.uint8, device=self.dev)
        if self.geom is None:
            self.geom, self.img = torch.empty(g, **u8), torch.empty(i, **u8)
        self.binning = torch.empty(b, **u8)
        self.capacity = int(capacity)

    def step(self, planes_nchw, xyz, cam, bg, target=None, dL_dimage=None):
        """Enqueue one optimisation step on the current stream.  Loss: L1 + L2 against ``target`` (the reference's two
        pixel losses, gaussian_splatting/utils/loss_utils.py:17-24) or, with ``dL_dimage``, that image gradient is
        back-propagated as is (benchmark mode).  Returns the image tensor; ``self.loss`` holds the loss (device)."""
        import math
        C, lib, p, P, dec = self.C, self.lib, self._lib.ptr, self.P, self.dec
        check = self._lib.check
        flat = dec._ensure_flat(planes_nchw)
        cfg, cfg32 = dec._cfg_for(planes_nchw), dec._cfg_for(planes_nchw, precision="fp32")
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        n = planes_nchw.numel()
        if self.planes32 is None or self.planes32.numel() != 4 * n:
            self.planes32 = torch.empty(4 * n, dtype=torch.uint8, device=self.dev)
            self.planes16 = torch.empty(2 * n, dtype=torch.uint8, device=self.dev) if cfg.precision == 1 else self.planes32
            nbytes = lib.ggd_decode_backward_scratch_bytes(C.byref(cfg), P)
            self.scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
        planes_nchw = planes_nchw.contiguous()
        # the tensor-core backward chain samples the fp16 planes of the forward: no fp32 copy needed then
        need_f32 = not (cfg.precision == 1 and cfg.pos_freqs == 0)
        if need_f32:
            check(lib.ggd_planes_channels_last(s, C.byref(cfg32), p(planes_nchw), p(self.planes32)))
        if cfg.precision == 1:
            check(lib.ggd_planes_channels_last(s, C.byref(cfg), p(planes_nchw), p(self.planes16)))
        o, a, g = self.out, self.act, self.g
        check(lib.ggd_decode_forward(s, C.byref(cfg), p(self.planes16), p(xyz), P, p(flat.packed_weights()),
                                     p(o["xyz"]), p(o["scale"]), p(o["rotation"]), p(o["opacity"]), p(o["color"]),
                                     p(a["scale"]), p(a["rotation"]), p(a["opacity"])))
        tfx, tfy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
        view, proj, campos = cam.world_view_transform, cam.full_proj_transform, cam.camera_center
        check(lib.ggd_raster_forward(s, P, 0, 1, p(bg), self.W, self.H, p(o["xyz"]), p(o["color"]), None, p(a["opacity"]),
                                     p(a["scale"]), 1.0, p(a["rotation"]), None, p(view), p(proj), p(campos), tfx, tfy, 0,
                                     p(self.geom), p(self.binning), self.capacity, p(self.img), p(self.image), p(self.radii),
                                     C.c_void_p(self.pinned.data_ptr()), self.bin_mode | self._lib.BINNING_LEAN))
        if dL_dimage is not None:
            dimg = dL_dimage
            self.loss.copy_((self.image * dL_dimage).sum())
        else:
            diff = self.image - target
            self.loss.copy_(diff.abs().mean() + (diff * diff).mean())
            torch.add(torch.sign(diff), diff, alpha=2.0, out=self.dimage)
            self.dimage.div_(diff.numel())
            dimg = self.dimage
        self._gblock.zero_()
        check(lib.ggd_raster_backward(s, P, 0, 1, p(bg), self.W, self.H, p(o["xyz"]), p(o["color"]), None, p(a["scale"]), 1.0,
                                      p(a["rotation"]), None, p(view), p(proj), p(campos), tfx, tfy, p(self.radii),
                                      p(self.geom), p(self.binning), self.capacity, p(self.img), p(dimg),
                                      p(g["mean2D"]), p(g["conic"]), p(g["opacity"]), p(g["color"]), p(g["mean3D"]),
                                      p(g["cov3D"]), p(g["sh"]), p(g["scale"]), p(g["rot"])))
        # the rasteriser saw exp / normalize / sigmoid of the decoder outputs: chain through them in place
        check(lib.ggd_activation_backward(s, P, p(a["scale"]), p(o["rotation"]), p(a["opacity"]),
                                          p(g["scale"]), p(g["rot"]), p(g["opacity"])))
        # decoder backward straight into the flat gradient buffer the parameters' .grad tensors view
        check(lib.ggd_decode_backward(s, C.byref(cfg), p(self.planes32) if need_f32 else None, p(self.planes16) if cfg.precision == 1 else None, p(xyz), P, p(flat.flat), p(flat.packed_weights_for(cfg32)),
                                      p(o["xyz"]), p(o["scale"]), p(o["rotation"]), p(o["opacity"]), p(o["color"]),
                                      p(g["mean3D"]), p(g["scale"]), p(g["rot"]), p(g["opacity"]), p(g["sh"]),
                                      p(self.scratch), self.scratch.numel(), p(flat.flat_grad), None))
        self._point_grads_at(flat)
        self._allreduce(flat.flat_grad)
        self.opt.step()
        return self.image

    def _point_grads_at(self, flat):
        """``.grad`` of the 40 parameters = views of the flat gradient buffer (set once; optimizer.step reads them)."""
        if getattr(self, "_grads_of", None) is flat.flat_grad:
            return
        for hi, li, is_b, t in flat._tensors(self.dec):
            t.grad = flat.view_of(flat.flat_grad, hi, li, is_b, t)
        self._grads_of = flat.flat_grad

    def _allreduce(self, buf):
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return
        main = torch.cuda.current_stream()
        if self._ar_events is None:
            self._ar_events = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        self._ar_bytes = buf.numel() * 4
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            self._ar_events[0].record()
            allreduce_flat(buf, self.group)
            self._ar_events[1].record()
        # anything enqueued on the main stream between here and this wait (a generator's own backward when
        # use_gen_finetune is on, logging) overlaps the collective
        main.wait_stream(self.side)

    def allreduce_bytes(self):
        return self._ar_bytes

    def allreduce_ms(self):
        """Device time of the last step's collective (+ scale), None on a single GPU."""
        if self._ar_events is None:
            return None
        self._ar_events[1].synchronize()
        return self._ar_events[0].elapsed_time(self._ar_events[1])

    def check(self):
        """Synchronise; returns (num_rendered, redo).  ``redo``: the last step ran with too small a binning buffer (now
        grown) — its gradient was computed from an incomplete image; run the step again."""
        torch.cuda.current_stream().synchronize()
        R = int(self.pinned[0].item()) & 0xffffffff
        over = R > self.capacity
        if over:
            self._alloc(int(R * 1.25) + 4096)
        if self.bin_mode:       # as FramePipeline.check: once a step was binned per tile, stop enqueueing the radix kernels
            C, check, ptr = self.C, self._lib.check, self._lib.ptr
            st, info = (C.c_uint32 * 2)(), (C.c_uint32 * 2)()
            s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            check(self.lib.ggd_raster_status(s, ptr(self.geom), self.P, st))
            check(self.lib.ggd_raster_binning_info(s, ptr(self.geom), self.P, info))
            if st[1] & 8:        # depths too clustered for the per-tile sort: this scene needs the radix sort
                self.bin_mode, over = 0, True
            elif st[1] & 2:      # mode 2 met a frame that needed the radix kernels
                self.bin_mode, over = 1, True
            elif self.bin_mode == 1 and info[0] and not over:
                self.bin_mode = 2
            elif self.bin_mode == 1 and not info[0] and not over:
                self.bin_mode = 0
        return R, over
