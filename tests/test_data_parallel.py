"""Host logic of the multi-GPU path on CPU: world_size-2 gloo processes."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from _stubs import StubG
from gaussian_gan_decoder_b200 import data_parallel as dp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, flat_mode, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gaussian_gan_decoder_b200.decoder_models import SequentialDecoderReverse
        torch.manual_seed(0)
        planes = torch.zeros(1, 3, 32, 8, 8)
        dec = SequentialDecoderReverse(StubG(planes), use_xyz_embedding=False, use_gen_finetune=False, device="cpu")
        dec._ensure_flat(planes)                          # host-only: builds the flat parameter/gradient buffers
        params = dec.get_params_custom()
        if flat_mode:
            dec._flat.flat_grad.copy_(torch.arange(dec._flat.count, dtype=torch.float32) * (rank + 1))
            assert dp.flat_gradient(dec) is dec._flat.flat_grad
        else:
            off = 0
            for p in params:                              # separate grad tensors: the cat/split idiom
                p.grad = (torch.arange(off, off + p.numel(), dtype=torch.float32) * (rank + 1)).view_as(p)
                off += p.numel()
            assert dp.flat_gradient(dec) is None
        if rank == 0:
            params[3].grad.view(-1)[0] = float("nan")     # nan_to_num like the reference idiom
        if flat_mode == "pipeline":       # what TrainPipeline issues (on a side stream when on a GPU)
            assert dp.allreduce_flat(dec._flat.flat_grad) is dec._flat.flat_grad
        else:
            dp.allreduce_gradients(dec)
        got = torch.cat([p.grad.reshape(-1) for p in params])
        expect = torch.arange(dec._flat.count, dtype=torch.float32) * (sum(range(1, world + 1)) / world)
        nan_at = sum(p.numel() for p in params[:3])
        mask = torch.ones_like(expect, dtype=torch.bool); mask[nan_at] = False
        ok = torch.allclose(got[mask], expect[mask]) and float(got[nan_at]) == 0.0
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("flat_mode", [True, False, "pipeline"])
def test_allreduce_gradients_world2_gloo(flat_mode):
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, flat_mode, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def test_shard_frames_partition():
    for by in ("latent", "view"):
        all_pairs = []
        for r in range(8):
            all_pairs += dp.shard_frames(64, 8, r, 8, by)
        assert sorted(all_pairs) == [(l, v) for l in range(64) for v in range(8)]
    assert len(dp.shard_frames(64, 8, 3, 8, "latent")) == 64
    assert {l for l, _ in dp.shard_frames(64, 8, 3, 8, "latent")} == set(range(3, 64, 8))      # 8 latents, all views
    assert {v for _, v in dp.shard_frames(64, 8, 3, 8, "view")} == {3}                          # one view, all latents
