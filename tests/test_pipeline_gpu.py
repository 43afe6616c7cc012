"""FramePipeline (pre-allocated, sync-free frame loop) gives the same image as the class API, eagerly
and replayed from a CUDA graph."""
import pytest
import torch

from _stubs import StubG
from gaussian_gan_decoder_b200 import synthetic
from gaussian_gan_decoder_b200.decoder_models import SequentialDecoderReverse
from gaussian_gan_decoder_b200.gaussian_renderer import GaussianModel, render_simple
from gaussian_gan_decoder_b200.pipeline import FramePipeline

pytestmark = pytest.mark.gpu


def test_pipeline_matches_class_api_and_graph_replay(cuda):
    g = synthetic.gen(8)
    planes = synthetic.planes("panohead", g, res=64, depth=3).to(cuda)
    xyz = synthetic.head_shell_points(30_000, g).to(cuda)
    torch.manual_seed(2)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", 3), use_xyz_embedding=False, use_gen_finetune=False, device=cuda)
    cam = synthetic.camera(256, 12.0)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).contiguous().to(cuda))
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    with torch.no_grad():
        res = dec(torch.zeros(1, 512, device=cuda), torch.zeros(1, 25, device=cuda), xyz, 1.0)
        ref = render_simple(cam, GaussianModel.from_decoder(res, dec.activated()), bg)["render"]
    pipe = FramePipeline(dec, 30_000, 256, cuda, capacity=1000)      # too small on purpose
    pipe.frame(planes, xyz, cam, bg)
    R, over = pipe.check()
    assert over and pipe.capacity >= R                                # grown after the overflow
    img = pipe.frame(planes, xyz, cam, bg).clone()
    R2, over = pipe.check()
    assert not over and R2 == R
    assert torch.equal(img, ref)
    pl2, xyz2 = planes.clone(), xyz.clone()
    pipe.capture(pl2, xyz2, cam, bg)
    xyz2.mul_(0.5)                                                    # new inputs, same graph
    out = pipe.replay().clone()
    torch.cuda.synchronize()
    eager = pipe.frame(pl2, xyz2, cam, bg)
    torch.cuda.synchronize()
    assert torch.equal(out, eager) and not torch.equal(out, ref)
    other = pipe.render_only(synthetic_cam(cuda, 0.6), bg)
    assert other.shape == (3, 256, 256) and bool(torch.isfinite(other).all())


def synthetic_cam(dev, h_off):
    import math
    cam = synthetic.camera(256, 12.0, h=math.pi / 2 + h_off)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).contiguous().to(dev))
    return cam


def test_packed_planes_handshake(cuda):
    """A plane producer may hand over fp16 channels-last planes (PackedPlanes, also via pinned host memory):
    same attributes and image as the fp32 NCHW route, without the repack kernel."""
    g = synthetic.gen(9)
    planes = synthetic.planes("panohead", g, res=64, depth=3).to(cuda)
    xyz = synthetic.head_shell_points(20_000, g).to(cuda)
    torch.manual_seed(3)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", 3), use_xyz_embedding=False, use_gen_finetune=False, device=cuda)
    cam = synthetic_cam(cuda, 0.0)
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    z, c = torch.zeros(1, 512, device=cuda), torch.zeros(1, 25, device=cuda)
    with torch.no_grad():
        ref = dec(z, c, xyz, 1.0)
    packed = dec.pack_planes(planes)
    assert packed.shape == tuple(planes.shape) and packed.data.numel() == planes.numel() * 2
    host = packed.to("cpu").pin_memory()
    dec.G.planes = host.to(cuda, non_blocking=True)            # the generator now "emits" packed planes
    with torch.no_grad():
        res = dec(z, c, xyz, 1.0)
    for k in ("xyz", "scale", "rotation", "opacity", "color"):
        assert torch.equal(res[k], ref[k]), k
    pipe = FramePipeline(dec, 20_000, 256, cuda)
    a = pipe.frame(planes, xyz, cam, bg).clone()
    b = pipe.frame(dec.G.planes, xyz, cam, bg).clone()
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    xyz_g = xyz.clone().requires_grad_(False)
    with pytest.raises(RuntimeError):
        dec(z, c, xyz_g, 1.0)                                    # parameters require grad and grad mode is on
    # the host-side producer writes the same bytes as the device kernel
    from gaussian_gan_decoder_b200.decoder_models.fused import PackedPlanes
    made_on_host = PackedPlanes.from_host(planes.cpu())
    assert made_on_host.shape == packed.shape and torch.equal(made_on_host.data, packed.data.cpu())


def test_pipeline_recovers_when_a_frame_stops_qualifying(cuda):
    """The pipeline promotes itself to tiles-only binning after a qualifying frame.  When a later frame needs more
    (here: splats blown up so that tile lists and the instance count explode) check() reports `redo`, falls back, and
    the re-run gives the image of the reference-facing path."""
    from gaussian_gan_decoder_b200 import rasterizer as rz
    g = synthetic.gen(10)
    planes = synthetic.planes("panohead", g, res=64, depth=3).to(cuda)
    P = 120_000
    xyz = synthetic.head_shell_points(P, g).to(cuda)
    torch.manual_seed(4)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", 3), use_xyz_embedding=False, use_gen_finetune=False, device=cuda)
    cam = synthetic_cam(cuda, 0.2)
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    pipe = FramePipeline(dec, P, 64, cuda)            # 16 tiles only
    pipe.frame(planes, xyz, cam, bg)
    _, redo = pipe.check()
    assert not redo and pipe.bin_mode == 2
    pipe.act["scale"].fill_(0.2)                       # every splat now covers the whole image: 16 lists of 120k (> 98304)
    for _ in range(4):
        img = pipe.render_only(cam, bg).clone()
        R, redo = pipe.check()
        if not redo:
            break
    assert not redo and pipe.bin_mode == 0 and R > 98304 * 16
    s = rz.GaussianRasterizationSettings(64, 64, __import__("math").tan(cam.FoVx * 0.5), __import__("math").tan(cam.FoVy * 0.5), bg, 1.0,
                                         cam.world_view_transform, cam.full_proj_transform, 0, cam.camera_center, False, False)
    ref, _ = rz.GaussianRasterizer(s)(means3D=pipe.out["xyz"], means2D=torch.zeros_like(pipe.out["xyz"]), opacities=pipe.act["opacity"],
                                      shs=pipe.out["color"][:, None, :].contiguous(), scales=pipe.act["scale"], rotations=pipe.act["rotation"])
    assert torch.equal(img, ref)


def test_two_host_threads_two_pipelines_different_binning_modes(cuda):
    """Two host threads drive two FramePipelines on two streams, one pinned to the radix sort and one to
    tiles-only binning, frames interleaving freely.  Each image must be bit-identical to the serial run:
    nothing on the hot path is process-global any more (binning mode, pinned instance-count word, SM count)."""
    import threading
    g = synthetic.gen(12)
    planes = synthetic.planes("panohead", g, res=64, depth=3).to(cuda)
    P = 60_000
    xyzs = [synthetic.head_shell_points(P, g).to(cuda) for _ in range(2)]
    torch.manual_seed(6)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", 3), use_xyz_embedding=False, use_gen_finetune=False, device=cuda)
    cams = [synthetic_cam(cuda, 0.0), synthetic_cam(cuda, 0.4)]
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    pipes = [FramePipeline(dec, P, 256, cuda) for _ in range(2)]
    modes = (0, 2)
    serial = []
    for k in range(2):
        pipes[k].bin_mode = modes[k]
        pipes[k].frame(planes, xyzs[k], cams[k], bg)
        R, redo = pipes[k].check()
        assert not redo and R > 0 and pipes[k].bin_mode == modes[k]
        serial.append(pipes[k].image.clone())
    assert not torch.equal(serial[0], serial[1])
    torch.cuda.synchronize()
    errors = []

    def worker(k):
        try:
            torch.cuda.set_device(cuda)
            st = torch.cuda.Stream(device=cuda)
            with torch.cuda.stream(st):
                for _ in range(40):
                    img = pipes[k].frame(planes, xyzs[k], cams[k], bg)
                    R, redo = pipes[k].check()
                    assert not redo and pipes[k].bin_mode == modes[k]
                    if not torch.equal(img, serial[k]):
                        errors.append((k, "image differs from the serial run"))
                        return
        except Exception as e:   # surfaced in the main thread
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_frames_are_bit_reproducible(cuda):
    """The same inputs give the same bits, frame after frame (decoder outputs, radii, instance count, image): three
    tiles in flight per CTA, persistent queues and atomically claimed tile segments must not leak their scheduling
    into the results.  (tools/frame_determinism.py runs the same check for hundreds of full-size frames.)"""
    g = synthetic.gen(14)
    planes = synthetic.planes("panohead", g, res=128, depth=3).to(cuda)
    P = 150_000
    xyz = synthetic.head_shell_points(P, g).to(cuda)
    torch.manual_seed(7)
    dec = SequentialDecoderReverse(StubG(planes, "panohead", 3), use_xyz_embedding=False, use_gen_finetune=False, device=cuda)
    cam = synthetic_cam(cuda, 0.1)
    bg = torch.tensor(synthetic.BG["panohead"], device=cuda)
    tl = synthetic.trained_like_attributes(P, synthetic.gen(15))
    act = synthetic.activated(tl)
    override = {"scale": act["scales"].to(cuda), "rotation": act["rotations"].to(cuda), "opacity": act["opacities"].to(cuda)}
    for ov in (None, override):
        pipe = FramePipeline(dec, P, 256, cuda, capacity=32 * P)
        ref = None
        for i in range(24):
            pipe.frame(planes, xyz, cam, bg, attr_override=ov)
            R, redo = pipe.check()
            assert not redo or i < 3
            if i < 4:
                continue                       # the binning mode settles in the first frames
            cur = (R, pipe.image.clone(), pipe.out["xyz"].clone(), pipe.out["rotation"].clone(), pipe.radii.clone())
            if ref is None:
                ref = cur
                continue
            assert cur[0] == ref[0]
            for a, b in zip(cur[1:], ref[1:]):
                assert torch.equal(a, b)
