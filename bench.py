#!/usr/bin/env python
"""Benchmark of the decode -> rasterise hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one frame: tri-grid sampling + 5-head
decoder for 500k Gaussians, then projection / binning / sort / blend of one 512x512 view
(BASELINE configs[1]).  ``value`` is frames/s with inputs resident in HBM and two independent frames in flight
(two FramePipelines on two CUDA streams; ``single_stream`` holds the one-after-the-other
figure, whose ms_per_step is the frame latency); ``e2e`` is the
same frame through the reference-facing Python API with HOST inputs (pinned H2D of planes
and positions, D2H of the image inside the timed region).  Multi-GPU: one process per GPU
(torchrun), frames are independent -> no collective on the data path, weak scaling.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

P_FULL, SIZE, FOV = 500_000, 512, 12.0
ARCH, VARIANT, DEPTH = "panohead", "sequential_reverse", 3
FLOPS_PER_POINT = {"parallel": 376_064, "sequential_reverse": 382_720, "sequential": 383_744}
METRIC = "decoded+rendered frames/sec @512^2, 500k Gaussians"
STAGES = ("planes_channels_last", "decode", "preprocess", "scan", "duplicate", "sort", "ranges", "blend")


def peaks():
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return dict(hbm=m["hbm_gbs"], tf=m["bf16_tflops_sustained"], tf_burst=m["bf16_tflops"], src="measured")
    except Exception:
        return dict(hbm=6650.0, tf=1400.0, tf_burst=1590.0, src="fallback")


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled from a thread every few
    milliseconds (the region is tens of milliseconds; an nvidia-smi child would deliver its first line too
    late), falling back to ``nvidia-smi -lms`` when NVML cannot be loaded."""

    NAMES = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index, self.nvml, self.h = [], None, index, None, None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample(self):
        n = self.nvml
        sm = n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM)
        try:
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        bits = (getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8), getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4))
        self.rows.append([str(sm), str(mx)] + ["Active" if r & b else "Not Active" for b in bits])

    def sample_now(self):
        """One sample from the calling thread (call it while the timed work is still queued on the GPU)."""
        if self.nvml is not None:
            try:
                self._sample()
            except Exception:
                pass

    def _poll(self):
        while not self._stop.is_set():
            try:
                self._sample()
            except Exception:
                return
            self._stop.wait(0.002)

    def __enter__(self):
        if self.nvml is not None:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return self
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.nvml is not None:
            self._stop.set()
            self.t.join(timeout=2)
        elif self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        reasons = [n for i, n in enumerate(self.NAMES) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def make_inputs(seed_offset, device, n_sets=3):
    """n_sets rotating input sets (planes + positions): 3 x 81 MB > the 126 MB L2, so no timed
    step finds its inputs cached from the previous one."""
    from gaussian_gan_decoder_b200 import synthetic
    sets = []
    for i in range(n_sets):
        g = synthetic.gen(seed_offset * 16 + i)
        sets.append((synthetic.planes(ARCH, g, res=256, depth=DEPTH), synthetic.head_shell_points(P_FULL, g)))
    return sets


def make_decoder(planes, device):
    from gaussian_gan_decoder_b200.synthetic import SyntheticGenerator as StubG
    from gaussian_gan_decoder_b200.decoder_models import SequentialDecoderReverse
    torch.manual_seed(303)
    return SequentialDecoderReverse(StubG(planes, ARCH, DEPTH), hidden_dim=128, use_xyz_embedding=False,
                                    use_gen_finetune=False, device=device)


def device_camera(device):
    from gaussian_gan_decoder_b200 import synthetic
    cam = synthetic.camera(SIZE, FOV)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).contiguous().to(device))
    return cam


def run_ours(args, rank, world, dev):
    import ctypes as C
    from gaussian_gan_decoder_b200 import _lib, synthetic
    from gaussian_gan_decoder_b200.gaussian_renderer import GaussianModel, render_simple
    from gaussian_gan_decoder_b200.pipeline import FramePipeline
    lib = _lib.load()
    torch.cuda.set_device(dev)
    host_sets = make_inputs(rank, dev)
    dev_sets = [(pl.to(dev), xyz.to(dev)) for pl, xyz in host_sets]
    dec = make_decoder(dev_sets[0][0], dev)
    cam = device_camera(dev)
    bg = torch.tensor(synthetic.BG[ARCH], dtype=torch.float32, device=dev)
    pipe = FramePipeline(dec, P_FULL, SIZE, dev)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value") ----
    for i in range(max(args.warmup, 3)):
        pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg)
        R, over = pipe.check()
    assert not over
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    R, over = pipe.check()
    assert not over, "binning capacity overflow inside the timed region"
    info = (C.c_uint32 * 2)()
    _lib.check(lib.ggd_raster_binning_info(None, _lib.ptr(pipe.geom), P_FULL, info))
    binning = {"mode": pipe.bin_mode, "per_tile": bool(info[0]), "longest_tile_list": int(info[1])}

    # ---- same frames, two in flight: frames are independent, so a second FramePipeline on a second stream lets the
    # rasteriser kernels of one frame share the SMs with the decoder kernel of the next (the persistent decoder
    # leaves ~18 KB of shared memory, 9k registers and 40 % of the issue slots of every SM unused)
    n_fl = int(os.environ.get("GGD_BENCH_IN_FLIGHT", "2"))
    pipes = [pipe] + [FramePipeline(dec, P_FULL, SIZE, dev) for _ in range(n_fl - 1)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(n_fl)]
    main = torch.cuda.current_stream()

    def run_two(n):
        for st in streams:
            st.wait_stream(main)
        for i in range(n):
            with torch.cuda.stream(streams[i % n_fl]):
                pipes[i % n_fl].frame(*dev_sets[i % len(dev_sets)], cam, bg)
        for st in streams:
            main.wait_stream(st)

    for k in range(1, n_fl):
        for i in range(3):
            with torch.cuda.stream(streams[k]):
                pipes[k].frame(*dev_sets[i % len(dev_sets)], cam, bg)
                assert not pipes[k].check()[1]
    run_two(6)
    barrier()
    with ClockSampler(dev.index or 0) as clk:
        e0.record()
        run_two(args.steps)
        e1.record()
        clk.sample_now()      # the steps are still executing: at least one sample under load even for a short region
        barrier()
    ms_two = e0.elapsed_time(e1)
    for pp, st in zip(pipes, streams):
        with torch.cuda.stream(st):
            assert not pp.check()[1]

    # ---- per-stage times (separate pass, stage events on) ----
    lib.ggd_stage_timing(1)
    acc = [0.0] * 8
    reps = 5
    for i in range(reps):
        pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg)
        t = (C.c_float * 8)()
        _lib.check(lib.ggd_stage_times(t))
        acc = [a + max(x, 0.0) for a, x in zip(acc, t)]
    lib.ggd_stage_timing(0)
    stage_ms = {n: a / reps for n, a in zip(STAGES, acc)}

    # ---- end to end through the public API with host inputs ("e2e") ----
    # Every step uploads ITS planes + positions from pinned host memory and downloads ITS image; the
    # upload of step i+1 runs on a copy stream while step i computes (two device buffers), as any
    # streaming caller would do.  Timed from the first upload to the last download.
    pinned = [(pl.pin_memory(), xyz.pin_memory()) for pl, xyz in host_sets]
    d_in = [(torch.empty_like(dev_sets[0][0]), torch.empty_like(dev_sets[0][1])) for _ in range(2)]
    img_host = [torch.empty(3, SIZE, SIZE).pin_memory() for _ in range(2)]
    z, c = torch.zeros(1, 512, device=dev), torch.zeros(1, 25, device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream()

    # variant: the plane producer keeps fp16 channels-last planes (PackedPlanes) in host memory
    from gaussian_gan_decoder_b200.decoder_models.fused import PackedPlanes
    pk_host = [dec.pack_planes(pl).to("cpu").pin_memory() for pl, _ in dev_sets]
    pk_dev = [PackedPlanes(torch.empty_like(pk_host[0].data, device=dev), pk_host[0].shape) for _ in range(2)]

    def e2e_run(n, packed=False):
        uploaded = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]

        def upload(i):
            b = i % 2
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(consumed[b])
                pl, xyz = pinned[i % len(pinned)]
                if packed:
                    pk_dev[b].data.copy_(pk_host[i % len(pk_host)].data, non_blocking=True)
                else:
                    d_in[b][0].copy_(pl, non_blocking=True)
                d_in[b][1].copy_(xyz, non_blocking=True)
                uploaded[b].record(copy_stream)

        upload(0)
        for i in range(n):
            b = i % 2
            if i + 1 < n:
                upload(i + 1)
            main_stream.wait_event(uploaded[b])
            if packed:      # streaming caller: sync-free FramePipeline over the C ABI, planes already in kernel layout
                image = pipe.frame(pk_dev[b], d_in[b][1], cam, bg)
            else:           # reference-facing classes: decoder.forward -> GaussianModel -> render_simple
                dec.G.planes = d_in[b][0]
                with torch.no_grad():
                    res = dec(z, c, d_in[b][1], 1.0)
                    gm = GaussianModel.from_decoder(res, dec.activated())
                    image = render_simple(cam, gm, bg)["render"]
            consumed[b].record(main_stream)
            img_host[b].copy_(image, non_blocking=True)
        torch.cuda.synchronize()
        if packed:
            assert not pipe.check()[1], "frame needs to be re-run (capacity / binning mode)"

    e2e_run(4)
    barrier()
    e0.record()
    e2e_run(args.steps)
    e1.record()
    barrier()
    e2e_s = e0.elapsed_time(e1) * 1e-3
    e2e_run(4, packed=True)
    barrier()
    e0.record()
    e2e_run(args.steps, packed=True)
    e1.record()
    barrier()
    e2e_pk_s = e0.elapsed_time(e1) * 1e-3
    dec.G.planes = dev_sets[0][0]

    if world > 1:
        t = torch.tensor([ms_total, e2e_s, e2e_pk_s, ms_two], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms_total, e2e_s, e2e_pk_s, ms_two = t.tolist()
    if rank != 0:
        return None
    pk = peaks()
    ms_step = ms_two / args.steps
    flops = FLOPS_PER_POINT[VARIANT] * P_FULL
    dec_ms = stage_ms["decode"]
    h2d = host_sets[0][0].numel() * 4 + host_sets[0][1].numel() * 4
    hbm = lambda nbytes, ms: round(nbytes / (ms * 1e-3) / 1e9, 1) if ms > 0 else None
    stage_bytes = {"planes_channels_last": 2 * host_sets[0][0].numel() * 4, "preprocess": 131 * P_FULL,
                   "scan": 8 * P_FULL, "duplicate": 20 * P_FULL + 12 * R, "sort": 152 * R,
                   "ranges": 8 * R + 8 * 1024, "blend": 40 * R + 20 * SIZE * SIZE}
    line = {
        "metric": METRIC, "value": round(world * args.steps / (ms_two * 1e-3), 2), "unit": "frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 x f16 -> f32 (decoder MMA), f32 (rasteriser)", "data": "synthetic",
        "config": {"workload": "configs[1]: PanoHead-arch (tri-grid D=3, 256^2x32ch) SequentialDecoderReverse + "
                               "rasterise, 500k Gaussians, one 512x512 view, random-init decoder weights",
                   "gaussians": P_FULL, "image": SIZE, "num_rendered": R, "fov_deg": FOV,
                   "l2_policy": "3 rotating input sets (3 x 81.5 MB > 126 MB L2)", "parallelism": f"replicas x{world}",
                   "binning": binning, "frames_in_flight": n_fl,
                   "schedule": "K independent frames over %d FramePipelines / CUDA streams (frame i on stream i mod %d); "
                               "one after the other: see single_stream" % (n_fl, n_fl)},
        "e2e": {"value": round(world * args.steps / e2e_s, 2), "unit": "frames/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 3 * SIZE * SIZE * 4, "ms_per_step": round(1e3 * e2e_s / args.steps, 3)},
        # planes repack, decode, preprocess, scan, blend + binning: tile scan / scatter / tile sort (mode 2), plus
        # duplicate, 6 digit passes and the range kernel when the radix kernels are enqueued as well (modes 0, 1)
        # same upload/compute/download loop for a streaming caller: planes kept on the host as fp16 channels-last
        # (PackedPlanes: half the upload, no repack) and frames enqueued through the sync-free FramePipeline
        "e2e_packed_planes": {"via": "FramePipeline.frame(PackedPlanes, ...)", "value": round(world * args.steps / e2e_pk_s, 2), "unit": "frames/s",
                              "h2d_bytes_per_step": pk_host[0].data.numel() + host_sets[0][1].numel() * 4,
                              "d2h_bytes_per_step": 3 * SIZE * SIZE * 4, "ms_per_step": round(1e3 * e2e_pk_s / args.steps, 3)},
        # the same K frames strictly one after the other on one stream: ms_per_step here is the latency of a frame
        "single_stream": {"value": round(world * args.steps / (ms_total * 1e-3), 2), "unit": "frames/s",
                          "ms_per_step": round(ms_total / args.steps, 4)},
        "gpu_launches": args.steps * {0: 13, 1: 16, 2: 8}[pipe.bin_mode],
        "clocks": clk.summary(),
        "roofline": {"kernel": "decode_tc_kernel", "bound": "tensor", "achieved": round(flops / (dec_ms * 1e-3) / 1e12, 2),
                     "peak": pk["tf"], "unit": "TFLOP/s", "frac": round(flops / (dec_ms * 1e-3) / 1e12 / pk["tf"], 4),
                     # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full (profiles/r1_d_summary.md):
                     # the fp16 planes (37.7 MB) are still in L2 from the repack kernel and the 48 MB of results
                     # are not written back before the kernel ends, so DRAM sees far less than the 92 MB moved
                     "traffic": 18.75e6, "peak_source": pk["src"] + " (sustained bf16)",
                     "note": "tcgen05 fp16 decode incl. tri-grid sampling; algorithmic FLOPs = 382720/pt x 500k"},
        "stages_ms": {k: round(v, 4) for k, v in stage_ms.items()},
        "stages_hbm_gbs": {k: hbm(b, stage_ms[k]) for k, b in stage_bytes.items()},
        "hbm_peak_gbs": pk["hbm"],
    }
    return line, dev_sets, dec, cam, bg


# --------------------------------------------------------------------------------------
# CPU side: the oracle (a port — the reference has no CPU rasteriser and its decoder is
# PyTorch code that cannot travel to the GPU box), timed on the host cores.
# --------------------------------------------------------------------------------------
def cpu_frame_fn(P_s):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import decoder_oracle as do
    from raster_oracle import CpuOracle
    from gaussian_gan_decoder_b200 import synthetic
    g = synthetic.gen(0)
    planes = synthetic.planes(ARCH, g, res=256, depth=DEPTH)
    xyz = synthetic.head_shell_points(P_FULL, g)[:P_s].contiguous()
    torch.manual_seed(303)
    flat = torch.cat([p.detach().reshape(-1) for p in make_decoder(planes, "cpu").get_params_custom()])
    cam = synthetic.camera(SIZE, FOV)
    oracle = CpuOracle()
    torch.set_num_threads(os.cpu_count() or 1)

    def frame():
        with torch.no_grad():
            r = do.decode(ARCH, VARIANT, 0, flat, planes, xyz, 1.0, DEPTH)
        scene = dict(means3D=r["xyz"], scales=r["act_scale"], rotations=r["act_rotation"], opacities=r["act_opacity"],
                     shs=r["color"][:, None, :], viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
                     campos=cam.camera_center, bg=torch.tensor(synthetic.BG[ARCH]), W=SIZE, H=SIZE,
                     tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), sh_degree=0, scale_modifier=1.0)
        return oracle.forward(scene)

    return frame, oracle.num_threads()


def time_cpu(P_s, steps, warmup):
    frame, threads = cpu_frame_fn(P_s)
    for _ in range(warmup):
        frame()
    t0 = time.perf_counter()
    for _ in range(steps):
        frame()
    dt = (time.perf_counter() - t0) / steps
    # a sample of P_s Gaussians stands for P_s / P_FULL of a frame (decode and binning scale with P)
    fps = 1.0 / (dt * P_FULL / P_s)
    return fps, dt, threads


def run_reference(args):
    # bounded sample: ~1.2 s of CPU work per step at 50k Gaussians on 16 cores
    P_s = 50_000 if args.steps * max(args.warmup, 1) > 6 else 100_000
    fps, dt, threads = time_cpu(P_s, args.steps, max(args.warmup, 1))
    sample = f"{P_s} of {P_FULL} Gaussians per step (torch-CPU decoder oracle + OpenMP raster oracle), scaled x{P_FULL // P_s}"
    return {
        "impl": "reference", "metric": METRIC, "value": round(fps, 4), "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": round(1e3 / fps, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: PanoHead-arch (tri-grid D=3, 256^2x32ch) SequentialDecoderReverse + "
                               "rasterise, 500k Gaussians, one 512x512 view, random-init decoder weights",
                   "gaussians": P_FULL, "image": SIZE, "fov_deg": FOV},
        "cpu_baseline": {"value": round(fps, 4), "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(fps, 4), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def run_extra(args, rank, world, dev):
    """Secondary workloads (reported in DESIGN.md / profiles/, not the driver's headline):
    train  = BASELINE configs[3]: decoder training step (decode + raster, fwd + bwd, Adam, grad all-reduce)
    views  = BASELINE configs[2]: 64 latents x 8 views at 512^2, latent-major sharding over the ranks
    hires  = BASELINE configs[4]: 2048^2 raster of 2M trained-like Gaussians (sort / blend bound)"""
    from gaussian_gan_decoder_b200 import synthetic
    from gaussian_gan_decoder_b200.data_parallel import shard_frames
    from gaussian_gan_decoder_b200.pipeline import FramePipeline
    from gaussian_gan_decoder_b200.training import training_step
    torch.cuda.set_device(dev)
    bg = torch.tensor(synthetic.BG[ARCH], dtype=torch.float32, device=dev)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    extra = {}
    if args.workload == "train":
        sets = [(pl.to(dev), xyz.to(dev)) for pl, xyz in make_inputs(rank, dev)]
        dec = make_decoder(sets[0][0], dev)
        cam = device_camera(dev)
        opt = torch.optim.Adam(dec.get_params_custom(), lr=9e-5)
        z, c = torch.zeros(1, 512, device=dev), torch.zeros(1, 25, device=dev)
        g = synthetic.gen(99 + rank)
        dimg = torch.randn(3, SIZE, SIZE, generator=g).to(dev)

        def step(i):
            dec.G.planes = sets[i % len(sets)][0]
            training_step(dec, opt, z, c, sets[i % len(sets)][1], cam, None, bg, dL_dimage=dimg)

        units, name = 1, "decoder training steps/s (decode+raster fwd+bwd, Adam, grad all-reduce), 500k Gaussians @512^2"
    elif args.workload == "views":
        pairs = shard_frames(64, 8, rank, world, "latent")
        lat = sorted({l for l, _ in pairs})
        planes = {l: synthetic.planes("eg3d", synthetic.gen(1000 + l), res=256).to(dev) for l in lat[:4]}   # 4 distinct plane sets, reused
        xyz = synthetic.head_shell_points(P_FULL, synthetic.gen(7)).to(dev)
        from gaussian_gan_decoder_b200.decoder_models import ParallelDecoder
        torch.manual_seed(303)
        dec = ParallelDecoder(synthetic.SyntheticGenerator(planes[lat[0]], "eg3d", 1), use_xyz_embedding=False, use_gen_finetune=False, device=dev)
        cams = []
        for v in range(8):
            cam = synthetic.camera(SIZE, FOV, h=math.pi / 2 + 0.25 * (v - 3.5))
            for k in ("world_view_transform", "full_proj_transform", "camera_center"):
                setattr(cam, k, getattr(cam, k).contiguous().to(dev))
            cams.append(cam)
        pipe = FramePipeline(dec, P_FULL, SIZE, dev)

        def step(i):   # one step = this rank's whole share: decode each latent once, render its 8 views
            for l in lat:
                pl = planes[lat[(lat.index(l)) % len(planes)] if l not in planes else l]
                pipe.frame(pl, xyz, cams[0], bg)
                for v in range(1, 8):
                    pipe.render_only(cams[v], bg)
            pipe.check()

        units, name = len(pairs), "frames/s over 64 latents x 8 views @512^2 (EG3D planes, ParallelDecoder), latent-major shards"
    else:   # hires
        P5, S5 = 2_000_000, 2048
        sc = synthetic.raster_scene(P5, size=S5, fov_deg=FOV, seed_offset=50 + rank, device=dev)
        from gaussian_gan_decoder_b200 import rasterizer as rz
        e = torch.empty(0, device=dev)

        def step(i):
            return rz._C.rasterize_gaussians(sc["bg"], sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                            sc["viewmatrix"], sc["projmatrix"], sc["tanfovx"], sc["tanfovy"], S5, S5, sc["shs"], 0,
                                            sc["campos"], False, False)[0]

        extra["num_rendered"] = step(0)
        units, name = 1, "rasterised frames/s @2048^2, 2M trained-like Gaussians (raster only)"
    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t[0])
        total_units = torch.tensor([units], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(total_units)
        units_all = float(total_units[0])
    else:
        units_all = units
    if rank == 0:
        print(json.dumps({"metric": name, "value": round(units_all * args.steps / (ms * 1e-3), 3), "unit": "1/s", "n_gpus": world,
                          "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 4),
                          "higher_is_better": True, "scaling": "weak" if args.workload != "views" else "strong",
                          "data": "synthetic", "config": {"workload": args.workload, **extra}}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="frame", choices=["frame", "train", "views", "hires"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args)), flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: the product has no CPU path")
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=dev)
    if args.workload != "frame":
        run_extra(args, rank, world, dev)
        if world > 1:
            torch.distributed.barrier()
            torch.distributed.destroy_process_group()
        return
    res = run_ours(args, rank, world, dev)
    if rank == 0:
        line = res[0]
        if world == 1 and not args.no_cpu_baseline:
            fps, dt, threads = time_cpu(100_000, 3, 1)
            line["cpu_baseline"] = {"value": round(fps, 4), "unit": "frames/s", "cores": threads, "kind": "port",
                                    "sample": "3 steps of 100000 of 500000 Gaussians (torch-CPU decoder oracle + "
                                              "OpenMP raster oracle), scaled x5"}
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
