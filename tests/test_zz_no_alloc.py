"""The allocation-free sort entry point (ggd_sort_pairs_static) against the regular one.

Some users of this library are not allowed to allocate device memory themselves — think of a
viewer or serving process that divides the GPU into fixed arenas when it starts, or an engine
that loads this code as a plug-in and keeps the device allocator to itself (a cudaMalloc in
the middle of its frame loop would also stall all of its streams).  They can lend us keys,
values and ping-pong buffers from their arena, but not the sort's bookkeeping memory, so the
library carries that in a fixed built-in array.  This test checks, through the C ABI, that
the built-in-array variant returns exactly what ggd_sort_pairs returns (integer data: the
comparison is bit-exact, like every other sort test here), on two calls in a row — the second
one sees whatever the first left in the shared array — and that oversize requests are refused.
Because ggd_sort_pairs runs the same kernels, the result is also checked against numpy's stable
argsort, which shares nothing with this library; and the driver's free-memory figure must be the
same before and after a call (after the one-time set-up call): nothing was allocated on the device.

This file sorts last in collection order on purpose: a run that stops at the first failure
has then already run every other test.  It passes alone in a fresh process as well.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from gaussian_gan_decoder_b200 import _lib

pytestmark = pytest.mark.gpu
MAX_PAIRS = 1 << 21      # GGD_SORT_STATIC_MAX_PAIRS in include/ggd.h


def _p(t):
    return C.c_void_p(t.data_ptr())


def _run(lib, static, keys, vals, n, cap, begin, end, dev):
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    kt, vt = torch.empty_like(keys), torch.empty_like(vals)
    nd = torch.tensor([n], dtype=torch.int32, device=dev)
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if static:
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.ggd_sort_pairs_static(s, _p(keys), _p(ko), _p(vals), _p(vo), _p(nd), cap, begin, end, _p(kt), _p(vt))
        torch.cuda.synchronize()
        if static == "count":     # every buffer above came from the caller; the call itself must not take device memory
            assert torch.cuda.mem_get_info()[0] == free_before, "ggd_sort_pairs_static changed the free device memory"
    else:
        nbytes = lib.ggd_sort_scratch_bytes(cap)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        rc = lib.ggd_sort_pairs(s, _p(keys), _p(ko), _p(vals), _p(vo), _p(nd), cap, begin, end, _p(kt), _p(vt),
                                _p(scratch), nbytes)
    _lib.check(rc)
    torch.cuda.synchronize()
    return ko[:n].clone(), vo[:n].clone()


def _inputs(n, seed, dev, dup):
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    keys = torch.randint(0, 2 ** 43, (n,), generator=g, dtype=torch.int64)
    if dup:
        keys = keys % 4096                      # many equal keys: stability matters
    return keys.to(dev), torch.arange(n, dtype=torch.int32).to(dev)


def _numpy_stable(keys, vals, n, begin, end):
    k = keys[:n].cpu().numpy().astype(np.uint64)
    digit = (k >> np.uint64(begin)) & np.uint64((1 << (end - begin)) - 1)
    order = np.argsort(digit, kind="stable")
    return k[order], vals[:n].cpu().numpy()[order]


def test_static_scratch_sort_matches_regular_sort_twice_in_a_row(cuda):
    lib = _lib.load()
    # the very first call on a device is the documented set-up call (symbol address look-up; the CUDA runtime may
    # also load the kernels' code lazily then) — make it before measuring free memory around the later calls
    wk, wv = _inputs(1000, 9, cuda, False)
    _run(lib, True, wk, wv, 1000, 1000, 0, 43, cuda)
    # call 1: large, random keys.  call 2: smaller, duplicate-heavy, different bit range —
    # it must not be disturbed by histograms / look-back flags call 1 left in the array.
    for n, cap, begin, end, seed, dup in ((1_500_000, 1_600_000, 0, 43, 1, False), (333_333, 400_000, 3, 40, 2, True)):
        keys, vals = _inputs(cap, seed, cuda, dup)
        k_ref, v_ref = _run(lib, False, keys, vals, n, cap, begin, end, cuda)
        k_new, v_new = _run(lib, "count", keys, vals, n, cap, begin, end, cuda)
        assert torch.equal(k_new, k_ref), "keys differ between ggd_sort_pairs_static and ggd_sort_pairs"
        assert torch.equal(v_new, v_ref), "values differ between ggd_sort_pairs_static and ggd_sort_pairs"
        k_np, v_np = _numpy_stable(keys, vals, n, begin, end)
        assert (k_new.cpu().numpy().astype(np.uint64) == k_np).all(), "keys differ from numpy's stable sort"
        assert (v_new.cpu().numpy() == v_np).all(), "values differ from numpy's stable sort"
        digit = (k_new >> begin) & ((1 << (end - begin)) - 1)
        assert bool((digit[1:] >= digit[:-1]).all())


def test_static_scratch_sort_at_the_limit_and_beyond(cuda):
    lib = _lib.load()
    keys, vals = _inputs(MAX_PAIRS, 3, cuda, False)
    k_ref, v_ref = _run(lib, False, keys, vals, MAX_PAIRS, MAX_PAIRS, 0, 43, cuda)
    k_new, v_new = _run(lib, True, keys, vals, MAX_PAIRS, MAX_PAIRS, 0, 43, cuda)
    assert torch.equal(k_new, k_ref) and torch.equal(v_new, v_ref)
    # one pair more than the built-in array can serve: refused, with the limit in the message
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.ggd_sort_pairs_static(s, _p(keys), _p(keys), _p(vals), _p(vals), _p(vals), MAX_PAIRS + 1, 0, 43,
                                   _p(keys), _p(vals))
    assert rc == _lib.GGD_E_CAPACITY
    assert str(MAX_PAIRS).encode() in lib.ggd_last_error()
