"""adlm_step_counts_noalloc (no counterpart in the reference): the loss
normalisers of a batch shard computed with the kernel's scratch in a static device array instead of a caller workspace.
Called through the C ABI and compared with adlm_step_counts on the same inputs — integers, so exactly — on two calls
in a row (the array keeps the first call's contents), then on a smaller and a larger shape, and at its size limits.
Collected last (test_zz_*)."""
import ctypes

import numpy as np
import pytest
import torch

from adlm_inherently_interpretable_segmentation_b200 import _lib, head, synth

pytestmark = pytest.mark.gpu


def _counts_noalloc(target, pc, N, C, ignore_void=True):
    lib = _lib.load()
    B, HW = target.shape[0], target[0].numel()
    out = torch.full((2,), -7, device="cuda", dtype=torch.int32)
    rc = lib.adlm_step_counts_noalloc(_lib.ptr(target), B, HW, _lib.ptr(pc), N, C, ctypes.c_int(1 if ignore_void else 0),
                                      _lib.ptr(out), _lib.stream_ptr())
    return rc, out


def _case(B, h, w, N, C, seed):
    target = torch.from_numpy(synth.block_labels(B, h, w, C, block=3, seed=seed).astype(np.int32)).cuda()
    pc = (torch.arange(N, dtype=torch.int32, device="cuda") // max(N // C, 1)).clamp_(max=C - 1)
    return target, pc


def test_no_alloc_counts_equal_the_workspace_entry_point_twice_in_a_row():
    for (B, h, w, N, C, seed) in [(16, 65, 65, 190, 19, 1), (16, 65, 65, 190, 19, 2), (3, 9, 11, 30, 3, 3),
                                  (64, 33, 29, 1000, 100, 4)]:
        target, pc = _case(B, h, w, N, C, seed)
        for ignore_void in (True, False):
            want = head.step_counts(target, pc, C, ignore_void=ignore_void)
            for rep in range(2):                                   # two calls in a row share the static array
                rc, got = _counts_noalloc(target, pc, N, C, ignore_void)
                assert rc == 0, _lib.last_error()
                torch.cuda.synchronize()
                assert got.cpu().tolist() == want.cpu().tolist(), (B, N, C, ignore_void, rep)
    # the numbers themselves: CE patches = non-void labels, checked against numpy
    target, pc = _case(4, 13, 17, 190, 19, 9)
    rc, got = _counts_noalloc(target, pc, 190, 19, True)
    assert rc == 0 and int(got[0]) == int((target.cpu().numpy() >= 1).sum())


def test_no_alloc_rejects_sizes_beyond_the_static_scratch():
    lib = _lib.load()
    target, pc = _case(2, 5, 5, 20, 4, 5)
    out = torch.zeros(2, device="cuda", dtype=torch.int32)
    big_pc = torch.zeros(2000, dtype=torch.int32, device="cuda")
    rc = lib.adlm_step_counts_noalloc(_lib.ptr(target), 2, 25, _lib.ptr(big_pc), 2000, 4, 1, _lib.ptr(out), _lib.stream_ptr())
    assert rc == _lib.E_UNSUPPORTED and "exceed the static scratch" in _lib.last_error()
    many = torch.ones((300, 4), dtype=torch.int32, device="cuda")
    rc = lib.adlm_step_counts_noalloc(_lib.ptr(many), 300, 4, _lib.ptr(pc), 20, 4, 1, _lib.ptr(out), _lib.stream_ptr())
    assert rc == _lib.E_UNSUPPORTED
    rc = lib.adlm_step_counts_noalloc(None, 2, 25, _lib.ptr(pc), 20, 4, 1, _lib.ptr(out), _lib.stream_ptr())
    assert rc == _lib.E_INVALID
    # and the entry point still works after the refusals
    rc, got = _counts_noalloc(target, pc, 20, 4)
    assert rc == 0 and got.cpu().tolist() == head.step_counts(target, pc, 4).cpu().tolist()
