#!/usr/bin/env python
"""Benchmark of the decode -> rasterise hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one frame: tri-grid sampling + 5-head decoder for 500k
Gaussians, then projection / binning / sort / blend of one 512x512 view (BASELINE configs[1]).

  value            frames/s, inputs resident in HBM, two independent frames in flight (two FramePipelines on two
                   CUDA streams); ``single_stream`` is the one-after-the-other figure = the frame latency
  e2e              the same frame through the reference-facing Python API with HOST inputs: pinned H2D of that
                   step's planes + positions and D2H of its image inside the timed region
  e2e_views        configs[2]'s shape through the same classes: planes uploaded once per latent, 8 views rendered
  roofline         decode_tc_kernel: algorithmic FLOPs / its measured launch duration, against the measured peak
  stages_*         per-stage device times of the path the frame actually took, with that path's own
                   algorithmic bytes; ``radix_path``: one pass with the global radix sort forced (mode 0), so
                   duplicate / onesweep / ranges have their own honest GB/s
  trained_like     the same frame with trained-like splat attributes overriding the decoded ones
  train            configs[3]: decoder training step (fwd + bwd + Adam + gradient all-reduce), scales with --gpus
  ref_cuda         B2: the UNMODIFIED reference (its decoder classes, GaussianModel, render_simple and its own
                   diff_gaussian_rasterization extension, staged/built under oracle/_ref) on the same GPU
  cpu_baseline     B1: the reference's decoder classes on the host cores + the CPU port of its rasteriser

Multi-GPU: one process per GPU (torchrun), frames are independent -> no collective on the inference data
path, weak scaling; the training leg all-reduces the flat decoder gradient over NCCL.
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
WORKLOAD = ("configs[1]: PanoHead-arch (tri-grid D=3, 256^2x32ch) SequentialDecoderReverse + rasterise, "
            "500k Gaussians, one 512x512 view, random-init decoder weights")
STAGES = ("planes_channels_last", "decode", "preprocess", "scan", "duplicate", "sort", "ranges", "blend")
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "roofline_traffic.json")


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


def bind_to_gpu_numa(index):
    """Pin this process to the CPUs NVML reports as local to GPU ``index`` BEFORE any pinned host buffer is
    allocated, so the pages land on the GPU's own NUMA node (first touch).  Eight ranks uploading 54 GB/s each
    from whichever node the allocator picked is what limited the 8-GPU e2e figure of round 1.  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(index).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


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


def device_camera(device, **kw):
    from gaussian_gan_decoder_b200 import synthetic
    cam = synthetic.camera(SIZE, FOV, **kw)
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(cam, k, getattr(cam, k).contiguous().to(device))
    return cam


def stage_bytes_for(per_tile, R, planes_numel, tiles=1024):
    """ALGORITHMIC bytes of each stage for the binning path the frame actually took.
    per tile : scan+scatter reads 20 B/Gaussian (means2D, depth, offset, radius), writes the reference-order key/value
               pair (12 B) and the tile-segment word (8 B) per instance; the tile sort reads that word and writes the
               sorted key/value pair (8 + 12 B); there is no separate ranges pass (the scan writes them).
    radix    : duplicate 20 B/Gaussian + 12 B/instance; sort 8 B (histogram read) + 6 passes x 24 B per instance;
               ranges 8 B/instance + 8 B/tile."""
    b = {"planes_channels_last": planes_numel * 4 + planes_numel * 2, "preprocess": 131 * P_FULL, "scan": 8 * P_FULL,
         "blend": 40 * R + 20 * SIZE * SIZE}
    if per_tile:
        b.update({"duplicate": 20 * P_FULL + 20 * R, "sort": 20 * R, "ranges": None})
    else:
        b.update({"duplicate": 20 * P_FULL + 12 * R, "sort": 152 * R, "ranges": 8 * R + 8 * tiles})
    return b


def hbm_gbs(nbytes, ms):
    return round(nbytes / (ms * 1e-3) / 1e9, 1) if (nbytes and ms and ms > 0) else None


def run_ours(args, rank, world, dev):
    import ctypes as C
    from gaussian_gan_decoder_b200 import _lib, synthetic
    from gaussian_gan_decoder_b200.gaussian_renderer import GaussianModel, render_simple
    from gaussian_gan_decoder_b200.pipeline import FramePipeline
    lib = _lib.load()
    torch.cuda.set_device(dev)
    numa_cpus = bind_to_gpu_numa(dev.index or 0)
    host_sets = make_inputs(rank, dev)
    dev_sets = [(pl.to(dev), xyz.to(dev)) for pl, xyz in host_sets]
    dec = make_decoder(dev_sets[0][0], dev)
    cam = device_camera(dev)
    bg = torch.tensor(synthetic.BG[ARCH], dtype=torch.float32, device=dev)
    pipe = FramePipeline(dec, P_FULL, SIZE, dev)
    warm = max(args.warmup, 3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def binning_info(pp):
        info = (C.c_uint32 * 2)()
        _lib.check(lib.ggd_raster_binning_info(None, _lib.ptr(pp.geom), P_FULL, info))
        return {"mode": pp.bin_mode, "per_tile": bool(info[0]), "longest_tile_list": int(info[1])}

    def stage_pass(pp, frame_fn, reps):
        """Per-stage device times: CUDA events the library records around each stage on the launching stream."""
        lib.ggd_stage_timing(1)
        acc, cnt = [0.0] * 8, [0] * 8
        for i in range(reps):
            frame_fn(i)
            t = (C.c_float * 8)()
            _lib.check(lib.ggd_stage_times(t))
            for k in range(8):
                if t[k] >= 0:
                    acc[k] += t[k]
                    cnt[k] += 1
        lib.ggd_stage_timing(0)
        return {n: (acc[k] / cnt[k] if cnt[k] else None) for k, n in enumerate(STAGES)}

    # ---- device-resident, one frame after the other ("single_stream": its ms_per_step is the frame latency) ----
    for i in range(warm):
        pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg)
        R, over = pipe.check()
    assert not over
    barrier()
    e0.record()
    for i in range(args.steps):
        pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    R, over = pipe.check()
    assert not over, "binning capacity overflow inside the timed region"
    binning = binning_info(pipe)

    # ---- "value": the same frames, two in flight.  Frames are independent, so a second FramePipeline on a second
    # stream lets the rasteriser kernels of one frame share the SMs with the persistent decoder kernel of the next.
    n_fl = int(os.environ.get("GGD_BENCH_IN_FLIGHT", "2"))
    pipes = [pipe] + [FramePipeline(dec, P_FULL, SIZE, dev) for _ in range(n_fl - 1)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(n_fl)]
    main = torch.cuda.current_stream()

    def run_in_flight(n):
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
    run_in_flight(6)
    barrier()
    with ClockSampler(dev.index or 0) as clk:
        launches0 = lib.ggd_launch_count()
        e0.record()
        run_in_flight(args.steps)
        e1.record()
        launches = lib.ggd_launch_count() - launches0     # counted by the library, one per kernel launch
        clk.sample_now()      # the steps are still executing: at least one sample under load even for a short region
        barrier()
    ms_two = e0.elapsed_time(e1)
    for pp, st in zip(pipes, streams):
        with torch.cuda.stream(st):
            assert not pp.check()[1]

    # ---- per-stage times of the path taken, then one pass with the radix sort forced ----
    reps = max(5, min(args.steps, 20))
    stage_ms = stage_pass(pipe, lambda i: pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg), reps)
    radix_pipe = FramePipeline(dec, P_FULL, SIZE, dev, capacity=pipe.capacity)
    radix_pipe.bin_mode = 0
    for i in range(3):
        radix_pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg)
    assert not radix_pipe.check()[1] and radix_pipe.bin_mode == 0
    radix_ms = stage_pass(radix_pipe, lambda i: radix_pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg), reps)
    del radix_pipe

    # ---- trained-like splats: same decode, attributes overridden before the rasteriser ----
    tl = synthetic.trained_like_attributes(P_FULL, synthetic.gen(77 + rank))
    act = synthetic.activated(tl)
    override = {"scale": act["scales"].to(dev), "rotation": act["rotations"].to(dev),
                "opacity": act["opacities"].to(dev), "color": tl["color"].to(dev)}
    tl_pipe = FramePipeline(dec, P_FULL, SIZE, dev, capacity=16 * P_FULL)
    tl_frame = lambda i: tl_pipe.frame(*dev_sets[i % len(dev_sets)], cam, bg, attr_override=override)
    for i in range(4):
        tl_frame(i)
        tl_R, tl_over = tl_pipe.check()
    assert not tl_over
    barrier()
    e0.record()
    for i in range(args.steps):
        tl_frame(i)
    e1.record()
    barrier()
    tl_ms = e0.elapsed_time(e1) / args.steps
    assert not tl_pipe.check()[1]
    tl_binning = binning_info(tl_pipe)
    tl_stage = stage_pass(tl_pipe, tl_frame, reps)
    tl_bytes = stage_bytes_for(tl_binning["per_tile"], tl_R, host_sets[0][0].numel())
    del tl_pipe

    # ---- end to end through the public API with host inputs ("e2e") ----
    # Every step uploads ITS planes + positions from pinned host memory and downloads ITS image; the upload of step
    # i+1 runs on a copy stream while step i computes (two device buffers), as any streaming caller would do.
    pinned = [(pl.pin_memory(), xyz.pin_memory()) for pl, xyz in host_sets]
    d_in = [(torch.empty_like(dev_sets[0][0]), torch.empty_like(dev_sets[0][1])) for _ in range(2)]
    img_host = [torch.empty(8, 3, SIZE, SIZE).pin_memory() for _ in range(2)]
    z, c = torch.zeros(1, 512, device=dev), torch.zeros(1, 25, device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream()
    view_cams = [device_camera(dev, h=math.pi / 2 + 0.25 * (v - 3.5)) for v in range(8)]

    # measured H2D rate of this process's pinned memory: the floor under every e2e figure below
    barrier()
    e0.record()
    for _ in range(4):
        d_in[0][0].copy_(pinned[0][0], non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    h2d_gbs = 4 * pinned[0][0].numel() * 4 / (e0.elapsed_time(e1) * 1e-3) / 1e9

    from gaussian_gan_decoder_b200.decoder_models.fused import PackedPlanes
    pk_host = [dec.pack_planes(pl).to("cpu").pin_memory() for pl, _ in dev_sets]
    pk_dev = [PackedPlanes(torch.empty_like(pk_host[0].data, device=dev), pk_host[0].shape) for _ in range(2)]

    def e2e_run(n, mode="frame"):
        """mode "frame": one view per upload through the decoder classes + render_simple; "views": 8 views per
        upload (configs[2]: attributes depend on the latent only); "packed": FramePipeline over host PackedPlanes."""
        uploaded = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]

        def upload(i):
            b = i % 2
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(consumed[b])
                pl, xyz = pinned[i % len(pinned)]
                if mode == "packed":
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
            if mode == "packed":
                image = pipe.frame(pk_dev[b], d_in[b][1], cam, bg)
                img_host[b][0].copy_(image, non_blocking=True)
            else:
                dec.G.planes = d_in[b][0]
                with torch.no_grad():
                    res = dec(z, c, d_in[b][1], 1.0)
                    gm = GaussianModel.from_decoder(res, dec.activated())
                    for v in range(8 if mode == "views" else 1):
                        image = render_simple(view_cams[v] if mode == "views" else cam, gm, bg)["render"]
                        img_host[b][v].copy_(image, non_blocking=True)
            consumed[b].record(main_stream)
        torch.cuda.synchronize()
        if mode == "packed":
            assert not pipe.check()[1], "frame needs to be re-run (capacity / binning mode)"

    e2e_s = {}
    for mode, n in (("frame", args.steps), ("views", max(4, args.steps // 2)), ("packed", args.steps)):
        e2e_run(4, mode)
        barrier()
        e0.record()
        e2e_run(n, mode)
        e1.record()
        barrier()
        e2e_s[mode] = (e0.elapsed_time(e1) * 1e-3, n)
    dec.G.planes = dev_sets[0][0]

    # ---- decoder training step (configs[3]) ----
    train = run_train_leg(args, rank, world, dev, dev_sets, cam, bg)

    red = [ms_total, ms_two, tl_ms] + [e2e_s[m][0] for m in ("frame", "views", "packed")]
    if world > 1:
        t = torch.tensor(red, device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        red = t.tolist()
        t = torch.tensor([h2d_gbs], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.SUM)
        h2d_total = float(t[0])
    else:
        h2d_total = h2d_gbs
    ms_total, ms_two, tl_ms, e2e_frame_s, e2e_views_s, e2e_pk_s = red
    if rank != 0:
        return None
    pk = peaks()
    flops = FLOPS_PER_POINT[VARIANT] * P_FULL
    dec_ms = stage_ms["decode"]
    tflops = flops / (dec_ms * 1e-3) / 1e12
    planes_numel = host_sets[0][0].numel()
    h2d = planes_numel * 4 + host_sets[0][1].numel() * 4
    sb = stage_bytes_for(binning["per_tile"], R, planes_numel)
    rb = stage_bytes_for(False, R, planes_numel)
    rnd = lambda d: {k: (round(v, 4) if v is not None else None) for k, v in d.items()}
    traffic, traffic_src = None, "no ncu capture of this build on record (profiles/roofline_traffic.json absent)"
    try:
        tj = json.load(open(TRAFFIC_FILE))
        traffic, traffic_src = tj["dram_bytes_per_launch"], tj.get("source", TRAFFIC_FILE)
    except Exception:
        pass
    n_views = e2e_s["views"][1]
    line = {
        "metric": METRIC, "value": round(world * args.steps / (ms_two * 1e-3), 2), "unit": "frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": warm, "ms_per_step": round(ms_two / args.steps, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16 x f16 -> f32 (decoder MMA), f32 (rasteriser)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "gaussians": P_FULL, "image": SIZE, "num_rendered": R, "fov_deg": FOV,
                   "l2_policy": "3 rotating input sets (3 x 81.5 MB > 126 MB L2)", "parallelism": f"replicas x{world}",
                   "binning": binning, "frames_in_flight": n_fl,
                   "schedule": "K independent frames over %d FramePipelines / CUDA streams (frame i on stream i mod %d); "
                               "one after the other: see single_stream" % (n_fl, n_fl)},
        "e2e": {"value": round(world * args.steps / e2e_frame_s, 2), "unit": "frames/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 3 * SIZE * SIZE * 4, "ms_per_step": round(1e3 * e2e_frame_s / args.steps, 3),
                "via": "SequentialDecoderReverse.forward -> GaussianModel -> render_simple (reference-facing classes)",
                "h2d_gbs_per_gpu": round(h2d_gbs, 1), "h2d_gbs_all_gpus": round(h2d_total, 1),
                "pcie_floor_ms": round(h2d / (h2d_gbs * 1e9) * 1e3, 3), "numa_bound_cpus": numa_cpus,
                "note": "bounded below by the upload: pcie_floor_ms = h2d bytes / this rank's measured pinned H2D rate"},
        # configs[2]'s shape: planes + positions uploaded once per latent, decoded once, 8 views rendered and downloaded
        "e2e_views": {"value": round(world * 8 * n_views / e2e_views_s, 2), "unit": "frames/s", "views_per_upload": 8,
                      "h2d_bytes_per_frame": h2d // 8, "d2h_bytes_per_frame": 3 * SIZE * SIZE * 4,
                      "ms_per_frame": round(1e3 * e2e_views_s / (8 * n_views), 3), "via": "same classes as e2e"},
        # same upload/compute/download loop for a streaming caller that keeps fp16 channels-last planes on the host
        "e2e_packed_planes": {"via": "FramePipeline.frame(PackedPlanes, ...)", "value": round(world * args.steps / e2e_pk_s, 2),
                              "unit": "frames/s", "h2d_bytes_per_step": pk_host[0].data.numel() + host_sets[0][1].numel() * 4,
                              "d2h_bytes_per_step": 3 * SIZE * SIZE * 4, "ms_per_step": round(1e3 * e2e_pk_s / args.steps, 3)},
        "single_stream": {"value": round(world * args.steps / (ms_total * 1e-3), 2), "unit": "frames/s",
                          "ms_per_step": round(ms_total / args.steps, 4)},
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        "roofline": {"kernel": "decode_tc_kernel", "bound": "tensor", "achieved": round(tflops, 2),
                     "peak": pk["tf_burst"], "unit": "TFLOP/s", "frac": round(tflops / pk["tf_burst"], 4),
                     "frac_of_sustained_peak": round(tflops / pk["tf"], 4), "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": pk["src"] + " (burst bf16: the kernel is timed alone, between events, in a %d-frame pass)" % reps,
                     "launch_ms": round(dec_ms, 4),
                     "note": "tcgen05 fp16 decode incl. tri-grid sampling; algorithmic FLOPs = 382720/pt x 500k"},
        "stages_ms": rnd(stage_ms),
        "stages_bytes": sb,
        "stages_hbm_gbs": {k: hbm_gbs(b, stage_ms.get(k)) for k, b in sb.items()},
        "stages_note": ("bytes are algorithmic, for the binning path taken (%s); stages served from L2 / shared memory can "
                        "exceed what DRAM delivers" % ("per tile" if binning["per_tile"] else "radix")),
        "radix_path": {"stages_ms": rnd({k: radix_ms[k] for k in ("duplicate", "sort", "ranges")}),
                       "stages_hbm_gbs": {k: hbm_gbs(rb[k], radix_ms[k]) for k in ("duplicate", "sort", "ranges")},
                       "note": "same frames with the global radix sort forced (binning mode 0): onesweep's own figure"},
        "hbm_peak_gbs": pk["hbm"],
        "trained_like": {"ms_per_frame": round(tl_ms, 4), "frames_per_s": round(world * 1e3 / tl_ms, 1), "num_rendered": tl_R,
                         "binning": tl_binning, "stages_ms": rnd(tl_stage),
                         "stages_hbm_gbs": {k: hbm_gbs(b, tl_stage.get(k)) for k, b in tl_bytes.items()},
                         "attributes": "log-scale N(-5.5,0.7), rotation randn, opacity logit N(2,2), colour U(-1,1) "
                                       "copied over the decoded ones before the rasteriser (decode still runs)"},
        "train": train,
    }
    return line


def run_train_leg(args, rank, world, dev, dev_sets, cam, bg):
    """BASELINE configs[3]: decode + rasterise forward and backward, Adam, one all-reduce of the flat decoder
    gradient per step (the idiom of eg3d/training/training_loop.py:288-297).  Reported inside the main line so the
    driver's 1/2/4/8-GPU run records the training curve as well."""
    from gaussian_gan_decoder_b200 import synthetic
    from gaussian_gan_decoder_b200.training import TrainPipeline
    dec = make_decoder(dev_sets[0][0], dev)
    # the reference's optimiser (Adam, lr 9e-5: main/train_pano2gaussian_decoder.py); torch's fused implementation: one kernel
    # over the 40 tensors instead of eight element-wise ones
    tp = TrainPipeline(dec, torch.optim.Adam(dec.get_params_custom(), lr=9e-5, fused=True), P_FULL, SIZE, dev)
    dimg = torch.randn(3, SIZE, SIZE, generator=synthetic.gen(99 + rank)).to(dev)
    steps = max(5, min(args.steps, 20))
    for i in range(3):
        tp.step(*dev_sets[i % len(dev_sets)], cam, bg, dL_dimage=dimg)
    tp.check()
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        tp.step(*dev_sets[i % len(dev_sets)], cam, bg, dL_dimage=dimg)
    e1.record()
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    tp.check()
    ar_ms = tp.allreduce_ms()
    if world > 1:
        t = torch.tensor([ms, ar_ms or 0.0], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms, ar_ms = t.tolist()
    return {"workload": "configs[3]: decoder training step (decode+raster fwd+bwd, Adam, gradient all-reduce), 500k Gaussians @512^2",
            "steps_per_s": round(world * 1e3 / ms, 2), "ms_per_step": round(ms, 4), "steps": steps, "n_gpus": world,
            "allreduce_bytes": tp.allreduce_bytes() if world > 1 else 0,
            "allreduce_ms": round(ar_ms, 4) if (world > 1 and ar_ms) else None,
            "allreduce": "one NCCL all_reduce of the flat fp32 gradient per step, issued on a side stream, waited "
                         "before optimizer.step()" if world > 1 else "single GPU: none"}


# --------------------------------------------------------------------------------------
# Baselines.  B1 (CPU): the reference's OWN decoder classes (staged by oracle/stage_reference.py)
# on the host cores + the CPU port of its rasteriser (the reference has no CPU rasteriser).
# B2 (GPU): the unmodified reference end to end on the same B200.
# --------------------------------------------------------------------------------------
def cpu_frame_fn():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_runtime
    from raster_oracle import CpuOracle
    from gaussian_gan_decoder_b200 import synthetic
    g = synthetic.gen(0)
    planes = synthetic.planes(ARCH, g, res=256, depth=DEPTH)
    xyz = synthetic.head_shell_points(P_FULL, g)
    cam = synthetic.camera(SIZE, FOV)
    oracle = CpuOracle()
    torch.set_num_threads(os.cpu_count() or 1)
    z, c = torch.zeros(1, 512), torch.zeros(1, 25)
    if ref_runtime.available():
        ref_runtime.add_paths(ARCH)
        dec = ref_runtime.make_decoder(VARIANT, ARCH, DEPTH, planes, "cpu")
        kind = "reference"
        parts = ("decoder: the reference's own main/decoder_models classes (oracle/_ref/pyref, unmodified) on torch CPU; "
                 "rasteriser: CPU port oracle/raster_oracle.cpp (OpenMP) — the reference has no CPU rasteriser")

        def decode():
            r = dec(z, c, xyz, 1.0)
            return dict(xyz=r.xyz, scale=torch.exp(r.scale), rot=torch.nn.functional.normalize(r.rotation),
                        op=torch.sigmoid(r.opacity), color=r.color)
    else:
        import decoder_oracle as do
        flat = torch.cat([p.detach().reshape(-1) for p in make_decoder(planes, "cpu").get_params_custom()])
        kind = "port"
        parts = "decoder: oracle/decoder_oracle.py (torch-CPU port; oracle/_ref/pyref not staged); rasteriser: oracle/raster_oracle.cpp"

        def decode():
            r = do.decode(ARCH, VARIANT, 0, flat, planes, xyz, 1.0, DEPTH)
            return dict(xyz=r["xyz"], scale=r["act_scale"], rot=r["act_rotation"], op=r["act_opacity"], color=r["color"])

    def frame():
        with torch.no_grad():
            r = decode()
        scene = dict(means3D=r["xyz"], scales=r["scale"], rotations=r["rot"], opacities=r["op"],
                     shs=r["color"][:, None, :], viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
                     campos=cam.camera_center, bg=torch.tensor(synthetic.BG[ARCH]), W=SIZE, H=SIZE,
                     tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), sh_degree=0, scale_modifier=1.0)
        return oracle.forward(scene)

    return frame, max(oracle.num_threads(), torch.get_num_threads()), kind, parts


def time_cpu(steps, warmup):
    frame, threads, kind, parts = cpu_frame_fn()
    for _ in range(warmup):
        frame()
    t0 = time.perf_counter()
    for _ in range(steps):
        frame()
    dt = (time.perf_counter() - t0) / steps
    return 1.0 / dt, dt, threads, kind, parts


def run_reference(args):
    """The reference arm: every step is one FULL frame (500k Gaussians, 512^2) on the host cores."""
    fps, dt, threads, kind, parts = time_cpu(args.steps, max(args.warmup, 1))
    sample = f"{args.steps} full frames of {P_FULL} Gaussians @ {SIZE}^2 after {max(args.warmup, 1)} warm-up (no scaling)"
    return {
        "impl": "reference", "metric": METRIC, "value": round(fps, 4), "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": round(1e3 * dt, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "gaussians": P_FULL, "image": SIZE, "fov_deg": FOV},
        "cpu_baseline": {"value": round(fps, 4), "unit": "frames/s", "cores": threads, "kind": kind, "parts": parts,
                         "sample": sample, "cpu_count": os.cpu_count()},
        "e2e": {"value": round(fps, 4), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def run_ref_cuda(args):
    """B2: the reference's own GPU path on this GPU — its decoder classes (torch fp32, TF32 off as shipped), its
    GaussianModel getters, its render_simple and its diff_gaussian_rasterization extension, all unmodified, staged
    and built under oracle/_ref by oracle/Makefile.  Runs in its own process (``--impl ref_cuda``) so that
    ``diff_gaussian_rasterization`` resolves to the reference's package there and to nothing of ours."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_runtime
    if not (ref_runtime.available(with_rasterizer=True) and torch.cuda.is_available()):
        return {"impl": "ref_cuda", "unavailable": "oracle/_ref/pyref (staged reference + its built extension) or a GPU is missing"}
    from gaussian_gan_decoder_b200 import synthetic
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    ref_runtime.add_paths(ARCH, reference_rasterizer=True)
    import diff_gaussian_rasterization as dgr
    assert os.path.realpath(dgr.__file__).startswith(os.path.realpath(ref_runtime.DGR_DIR)), dgr.__file__
    from gaussian_renderer import render_simple             # reference
    from scene.cameras import CustomCam                     # reference
    sets = [(pl.to(dev), xyz.to(dev)) for pl, xyz in make_inputs(0, dev)]
    dec = ref_runtime.make_decoder(VARIANT, ARCH, DEPTH, sets[0][0], dev)
    host_cam = synthetic.camera(SIZE, FOV)
    cam = CustomCam(size=SIZE, fov=host_cam.FoVx, extr=synthetic.look_at_cam2world(math.pi / 2 + 0.3, math.pi / 2 - 0.2, 2.7).to(dev))
    bg = torch.tensor(synthetic.BG[ARCH], dtype=torch.float32, device=dev)
    z, c = torch.zeros(1, 512, device=dev), torch.zeros(1, 25, device=dev)

    def decode(i):
        dec.G.planes = sets[i % len(sets)][0]
        return dec(z, c, sets[i % len(sets)][1], 1.0)

    def frame(i):
        with torch.no_grad():
            return render_simple(cam, ref_runtime.gaussian_model_from(decode(i)), bg_color=bg)["render"]

    warm, steps = 20, max(args.steps, 20)
    for i in range(warm):
        img = frame(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        img = frame(i)
    e1.record()
    torch.cuda.synchronize()
    total = e0.elapsed_time(e1) / steps
    e0.record()
    with torch.no_grad():
        for i in range(steps):
            r = decode(i)
    e1.record()
    torch.cuda.synchronize()
    dec_ms = e0.elapsed_time(e1) / steps
    gm = ref_runtime.gaussian_model_from(r)
    t0 = time.perf_counter()          # the reference's rasteriser synchronises inside (num_rendered D2H): wall clock
    with torch.no_grad():
        for i in range(steps):
            out = render_simple(cam, gm, bg_color=bg)
    torch.cuda.synchronize()
    ras_ms = (time.perf_counter() - t0) * 1e3 / steps
    return {"impl": "ref_cuda", "what": "unmodified reference: decoder classes (fp32, TF32 off) + GaussianModel + render_simple + "
                                        "diff_gaussian_rasterization/_C built from its sources for sm_100a",
            "value": round(1e3 / total, 2), "unit": "frames/s", "ms_per_step": round(total, 3), "decode_ms": round(dec_ms, 3),
            "raster_ms": round(ras_ms, 3), "visible": int((out["radii"] > 0).sum()), "steps": steps, "warmup": warm,
            "gpu": torch.cuda.get_device_name(0)}


def ref_cuda_subprocess(args):
    try:
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "ref_cuda", "--steps", str(args.steps)],
                             capture_output=True, text=True, timeout=600, env={**os.environ, "CUDA_VISIBLE_DEVICES": os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0]})
        for line in reversed(out.stdout.strip().splitlines()):
            if line.startswith("{"):
                return json.loads(line)
        return {"impl": "ref_cuda", "unavailable": "no JSON from the child: " + (out.stderr.strip().splitlines() or ["?"])[-1][:300]}
    except Exception as e:
        return {"impl": "ref_cuda", "unavailable": repr(e)[:300]}


def run_extra(args, rank, world, dev):
    """Secondary workloads (reported in DESIGN.md / profiles/, not the driver's headline):
    train  = BASELINE configs[3] alone (also a sub-object of the main line)
    views  = BASELINE configs[2]: 64 latents x 8 views at 512^2, latent-major sharding over the ranks
    hires  = BASELINE configs[4]: 2048^2 raster of 2M trained-like Gaussians (sort / blend bound)
    config0 = BASELINE configs[0]: 100k-point decoder forward on EG3D tri-planes, GPU beside the reference class on the host cores"""
    from gaussian_gan_decoder_b200 import synthetic
    from gaussian_gan_decoder_b200.data_parallel import shard_frames
    from gaussian_gan_decoder_b200.pipeline import FramePipeline
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
        if rank == 0:
            print(json.dumps(run_train_leg(args, rank, world, dev, sets, device_camera(dev), bg)), flush=True)
        else:
            run_train_leg(args, rank, world, dev, sets, device_camera(dev), bg)
        return
    if args.workload == "config0":
        # BASELINE configs[0]: EG3D tri-planes (3 x 32 x 256^2), ParallelDecoder as the reference constructs it (positional
        # embedding on), 100k points, decode only.  The generator is out of scope: seeded synthetic planes stand in for
        # G.synthesis, on both sides.  Timed here: this package on the GPU (tensor-core path, then the exact fp32 path) and
        # the reference's own class on the host cores (oracle/_ref/pyref; rank 0).
        from gaussian_gan_decoder_b200.decoder_models import ParallelDecoder
        P0 = 100_000
        planes = synthetic.planes("eg3d", synthetic.gen(11), res=256)
        xyz = synthetic.head_shell_points(P0, synthetic.gen(12))
        torch.manual_seed(303)
        dec = ParallelDecoder(synthetic.SyntheticGenerator(planes.to(dev), "eg3d", 1), use_xyz_embedding=True, use_gen_finetune=False, device=dev)
        z, c, xyz_d = torch.zeros(1, 512, device=dev), torch.zeros(1, 25, device=dev), xyz.to(dev)

        def gpu_ms(precision):
            dec.precision = precision
            with torch.no_grad():
                for _ in range(max(args.warmup, 3)):
                    dec(z, c, xyz_d, 1.0)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(args.steps):
                    dec(z, c, xyz_d, 1.0)
                e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / args.steps
        ms16, ms32 = gpu_ms("fp16"), gpu_ms("fp32")
        cpu = None
        if rank == 0:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import ref_runtime
            if ref_runtime.available():
                ref_runtime.add_paths("eg3d")
                torch.set_num_threads(os.cpu_count() or 1)
                rdec = ref_runtime.make_decoder("parallel", "eg3d", 1, planes, "cpu", use_xyz_embedding=True)
                zc, cc = torch.zeros(1, 512), torch.zeros(1, 25)
                with torch.no_grad():
                    rdec(zc, cc, xyz, 1.0)
                    t0 = time.perf_counter()
                    n = max(3, min(args.steps, 10))
                    for _ in range(n):
                        rdec(zc, cc, xyz, 1.0)
                    cpu = {"ms": round(1e3 * (time.perf_counter() - t0) / n, 2), "cores": torch.get_num_threads(), "runs": n,
                           "kind": "reference", "what": "main/decoder_models/parallel_decoder.py (oracle/_ref/pyref, unmodified), torch CPU"}
            print(json.dumps({"metric": "decoder forwards/s, 100k points, EG3D tri-planes, ParallelDecoder + positional embedding (configs[0], "
                                        "without G.synthesis: generator out of scope, synthetic planes)",
                              "value": round(1e3 / ms16, 2), "unit": "1/s", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
                              "ms_per_step": round(ms16, 4), "higher_is_better": True, "scaling": "weak", "data": "synthetic", "dtype": "fp16",
                              "config": {"workload": "config0", "points": P0, "fp32_path_ms": round(ms32, 4), "cpu_reference": cpu}}), flush=True)
        return
    if args.workload == "views":
        pairs = shard_frames(64, 8, rank, world, "latent")
        lat = sorted({l for l, _ in pairs})
        planes = {l: synthetic.planes("eg3d", synthetic.gen(1000 + l), res=256).to(dev) for l in lat[:4]}   # 4 distinct plane sets, reused
        xyz = synthetic.head_shell_points(P_FULL, synthetic.gen(7)).to(dev)
        from gaussian_gan_decoder_b200.decoder_models import ParallelDecoder
        torch.manual_seed(303)
        dec = ParallelDecoder(synthetic.SyntheticGenerator(planes[lat[0]], "eg3d", 1), use_xyz_embedding=False, use_gen_finetune=False, device=dev)
        cams = [device_camera(dev, h=math.pi / 2 + 0.25 * (v - 3.5)) for v in range(8)]
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
        import ctypes as C
        from gaussian_gan_decoder_b200 import _lib, rasterizer as rz
        lib = _lib.load()
        P5, S5 = 2_000_000, 2048
        sc = synthetic.raster_scene(P5, size=S5, fov_deg=FOV, seed_offset=50 + rank, device=dev)
        e = torch.empty(0, device=dev)

        def step(i):
            return rz._C.rasterize_gaussians(sc["bg"], sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                            sc["viewmatrix"], sc["projmatrix"], sc["tanfovx"], sc["tanfovy"], S5, S5, sc["shs"], 0,
                                            sc["campos"], False, False)

        first = step(0)
        R5 = extra["num_rendered"] = first[0]
        step(1)
        lib.ggd_stage_timing(1)
        out = step(2)
        t = (C.c_float * 8)()
        _lib.check(lib.ggd_stage_times(t))
        lib.ggd_stage_timing(0)
        info = (C.c_uint32 * 2)()
        _lib.check(lib.ggd_raster_binning_info(None, _lib.ptr(out[3]), P5, info))
        st = {n: (round(t[k], 4) if t[k] >= 0 else None) for k, n in enumerate(STAGES)}
        tiles5 = (S5 // 16) ** 2
        if info[0]:    # binned per tile (class API: reference-order arrays kept): scatter 20 B/Gaussian + 20 B/instance, tile sort 20 B/instance
            by = {"preprocess": 131 * P5, "duplicate": 20 * P5 + 20 * R5, "sort": 20 * R5, "ranges": None, "blend": 40 * R5 + 20 * S5 * S5}
            note = "per-tile path: algorithmic bytes scatter 20 P + 20 R, tile sort 20 R (lists mostly stream through L2 / shared memory)"
        else:
            by = {"preprocess": 131 * P5, "duplicate": 20 * P5 + 12 * R5, "sort": 152 * R5, "ranges": 8 * R5 + 8 * tiles5,
                  "blend": 40 * R5 + 20 * S5 * S5}
            note = "radix-path algorithmic bytes (sort = 8 + 6 x 24 B per instance)"
        extra.update(stages_ms=st, binning={"per_tile": bool(info[0]), "longest_tile_list": int(info[1])},
                     stages_hbm_gbs={k: hbm_gbs(b, st.get(k)) for k, b in by.items()}, hbm_peak_gbs=peaks()["hbm"],
                     stages_note=note)
        del first, out
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
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ref_cuda"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--workload", default="frame", choices=["frame", "train", "views", "hires", "config0"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args)), flush=True)
        return
    if args.impl == "ref_cuda":
        if rank == 0:
            print(json.dumps(run_ref_cuda(args)), flush=True)
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
    line = run_ours(args, rank, world, dev)
    if rank == 0:
        if world == 1 and not args.no_ref_cuda:
            torch.cuda.synchronize()
            line["ref_cuda"] = ref_cuda_subprocess(args)
        if world == 1 and not args.no_cpu_baseline:
            fps, dt, threads, kind, parts = time_cpu(3, 1)
            line["cpu_baseline"] = {"value": round(fps, 4), "unit": "frames/s", "cores": threads, "kind": kind, "parts": parts,
                                    "sample": "3 full frames (500000 Gaussians @ 512^2) after 1 warm-up, no scaling",
                                    "ms_per_step": round(1e3 * dt, 1), "cpu_count": os.cpu_count()}
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
