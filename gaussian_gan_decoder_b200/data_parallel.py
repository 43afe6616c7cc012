"""Data-parallel decoder training: one process per GPU, one collective per step.

Frames ((latent, view) samples) are independent, so the only exchange on the path is the
gradient all-reduce.  The reference has no collective on the decoder path (its training loop is
single GPU, main/train_pano2gaussian_decoder.py:118); the idiom mirrored here is the vendored
GAN trainer's flat-gradient all-reduce (eg3d/training/training_loop.py:288-297): concatenate,
``all_reduce``, divide by the world size, ``nan_to_num``, split back.  Because the decoder keeps
its 40 parameter tensors as views of one flat buffer (decoder_models/fused.py::FlatParams) and
the backward kernels write one flat gradient buffer, the concatenate / split steps disappear:
ONE ``all_reduce`` over 0.76 MB (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def flat_gradient(decoder):
    """The contiguous gradient buffer behind the decoder's 40 parameter tensors, or None if the
    ``.grad`` tensors are not (all) views of one buffer."""
    flat = getattr(decoder, "_flat", None)
    if flat is None:
        return None
    params = [t for _, _, _, t in flat._tensors(decoder)]
    if any(p.grad is None for p in params):
        return None
    for cand in (flat.flat_grad, getattr(flat, "last_dflat", None)):
        if cand is None:
            continue
        base, ok, off = cand.data_ptr(), True, 0
        for p in params:
            if p.grad.data_ptr() != base + 4 * off or not p.grad.is_contiguous():
                ok = False
                break
            off += p.numel()
        if ok and off == cand.numel():
            return cand
    return None


def allreduce_gradients(decoder, extra_params=(), group=None, async_op=False):
    """Average the decoder gradients (and ``extra_params``' — e.g. the generator's when
    use_gen_finetune is on) over the process group.  Returns the work handle(s) when async."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return []
    world = dist.get_world_size(group)
    handles = []
    flat = flat_gradient(decoder)
    if flat is not None:
        bufs = [(flat, None)]
    else:   # gradients live in separate tensors: the reference's cat / split idiom
        params = [p for p in decoder.get_params_custom()[:40] if p.grad is not None]
        bufs = [(torch.cat([p.grad.reshape(-1) for p in params]), params)]
    extra = [p for p in extra_params if p.grad is not None]
    if extra:
        bufs.append((torch.cat([p.grad.reshape(-1) for p in extra]), extra))
    for buf, params in bufs:
        h = dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group, async_op=True)
        handles.append((h, buf, params))

    def finish():
        for h, buf, params in handles:
            h.wait()
            buf.div_(world)
            torch.nan_to_num(buf, nan=0, posinf=1e5, neginf=-1e5, out=buf)
            if params is not None:
                off = 0
                for p in params:
                    p.grad.copy_(buf[off:off + p.numel()].view_as(p.grad))
                    off += p.numel()

    if async_op:
        return finish
    finish()
    return []


def allreduce_flat(buf, group=None):
    """In place: ``buf`` <- mean over the group's ranks, NaN/Inf cleaned like the vendored trainer does
    (eg3d/training/training_loop.py:288-297, minus its cat / split: the buffer IS the flat gradient).  Enqueue-only
    on CUDA tensors (NCCL); TrainPipeline calls it on a side stream."""
    world = dist.get_world_size(group)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    buf.div_(world)
    torch.nan_to_num(buf, nan=0, posinf=1e5, neginf=-1e5, out=buf)
    return buf


def shard_frames(num_latents, num_views, rank, world, by="latent"):
    """(latent, view) pairs owned by ``rank``.  "latent": latent-major (each GPU decodes its
    latents once and renders all their views — the recommended layout); "view": BASELINE
    configs[2] as written (GPU g renders view g of every latent, so every GPU decodes every
    latent)."""
    if by == "latent":
        return [(l, v) for l in range(rank, num_latents, world) for v in range(num_views)]
    if by == "view":
        return [(l, v) for v in range(rank, num_views, world) for l in range(num_latents)]
    raise ValueError(by)
