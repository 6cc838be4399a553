"""world_size-2 checks of the multi-GPU host logic on CPU (gloo): the collectives in parallel.py and the
data-parallel split of the step (global normalisers + additive shard gradients), using the oracle as the
arithmetic.  No CUDA involved."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import head_oracle as orc
from adlm_inherently_interpretable_segmentation_b200 import _lib, parallel, push, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(rank, size, port, fn, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        ret[rank] = fn(rank, size)
    finally:
        dist.destroy_process_group()


def spawn(fn, size=2):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_run, args=(size, _free_port(), fn, ret), nprocs=size, join=True)
    return [ret[r] for r in range(size)]


def _push_case():
    n_img, K, h, w, N, C, H, W = 4, 16, 5, 6, 20, 4, 40, 48
    z, P, _ = synth.head_inputs(n_img, K, h, w, N, C, seed=11)
    labels = synth.block_labels(n_img, H, W, C, block=8, seed=12)
    return z, P, labels, C


def _push_worker(rank, size):
    """each rank searches its contiguous image range with the oracle, then the packed words are merged"""
    z, P, labels, C = _push_case()
    n_img, K, h, w = z.shape
    N = P.shape[0]
    lo, hi = parallel.image_range(n_img, rank, size)
    proto_class = (np.arange(N) // (N // C)).astype(np.int64)
    packed = np.full(N, _lib.PUSH_NONE, dtype=np.int64)
    rows = np.zeros((N, K), dtype=np.float32)
    for i in range(lo, hi):
        mask = orc.patch_class_mask(labels[i], h, w)
        dmin, amin = orc.push_search_image(z[i], P, mask, proto_class)
        for j in range(N):
            if amin[j] >= 0:
                word = (int(np.float32(dmin[j]).view(np.uint32)) << 32) | (i * h * w + int(amin[j]))
                packed[j] = min(packed[j], word)
    t = torch.from_numpy(packed)
    parallel.merge_packed_min(t)
    dmin, ids = push.unpack(t.numpy())
    for j in range(N):
        if ids[j] >= 0 and lo <= ids[j] // (h * w) < hi:
            img, p = divmod(int(ids[j]), h * w)
            rows[j] = z[img].reshape(K, h * w)[:, p]
    r = torch.from_numpy(rows)
    parallel.merge_owned_rows(r)
    return t.numpy().copy(), r.numpy().copy()


def test_push_shard_merge_equals_single_pass():
    z, P, labels, C = _push_case()
    gmin, gpatch, gidx, _ = orc.push_search(z, P, labels, C)
    out = spawn(_push_worker)
    for packed, rows in out:                      # identical on both ranks
        d, ids = push.unpack(packed)
        assert np.array_equal(ids, gidx)
        fin = gidx >= 0
        assert np.allclose(d[fin], gmin[fin], rtol=1e-6) and np.all(np.isinf(d[~fin]))
        assert np.array_equal(rows, gpatch.astype(np.float32))


def _dp_case():
    B, K, h, w, N, C = 4, 16, 6, 7, 20, 4
    z, P, W = synth.head_inputs(B, K, h, w, N, C, w_noise=0.1, seed=21)
    target = synth.block_labels(B, h, w, C, block=2, seed=22).astype(np.int64)
    return z, P, W, target, (1.0, 0.25, 1e-4)


def _dp_worker(rank, size):
    z, P, W, target, wts = _dp_case()
    lo, hi = parallel.image_range(z.shape[0], rank, size)
    local = orc.step_loss(z[lo:hi], P, W, target[lo:hi], weights=wts, want_grads=False)
    counts = torch.tensor([local["n_patches"], local["n_kld_terms"]], dtype=torch.int32)
    parallel.allreduce_counts(counts)
    norm = (int(counts[0]), int(counts[1]))
    r = orc.step_loss(z[lo:hi], P, W, target[lo:hi], weights=(wts[0], wts[1], wts[2] / size), norm=norm)
    dP = torch.from_numpy(r["dP"].astype(np.float32))
    dW = torch.from_numpy(r["dW"].astype(np.float32))
    scalars = torch.tensor([r["loss"], r["cross_entropy"], r["kld_loss"], r["l1"], r["n_correct"],
                            norm[0], norm[1], 0.0], dtype=torch.float64)
    parallel.allreduce_head_grads(dP, dW, scalars)
    return dP.numpy(), dW.numpy(), scalars.numpy(), r["dz"], (lo, hi)


def test_data_parallel_step_equals_single_gpu_step():
    z, P, W, target, wts = _dp_case()
    full = orc.step_loss(z, P, W, target, weights=wts)
    out = spawn(_dp_worker)
    for dP, dW, scalars, dz, (lo, hi) in out:
        assert np.allclose(dP, full["dP"], rtol=1e-5, atol=1e-7)
        assert np.allclose(dW, full["dW"], rtol=1e-5, atol=1e-7)
        assert np.allclose(dz, full["dz"][lo:hi], rtol=1e-9, atol=1e-12)
        assert scalars[0] == pytest.approx(full["loss"], rel=1e-6)
        assert scalars[1] == pytest.approx(full["cross_entropy"], rel=1e-6)
        assert scalars[2] == pytest.approx(full["kld_loss"], rel=1e-6)
        assert scalars[3] == pytest.approx(full["l1"], rel=1e-6)            # replicated, not summed
        assert scalars[4] == full["n_correct"] and scalars[5] == full["n_patches"]
        assert scalars[6] == full["n_kld_terms"]


def test_image_ranges_are_contiguous_and_cover():
    for n, size in ((2975, 8), (5, 2), (3, 4), (0, 2)):
        spans = [parallel.image_range(n, r, size) for r in range(size)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


# ------------------------------------------------------------------ k nearest patches, images sharded over ranks
def _nearest_case():
    n, K, h, w, N, C, H, W, k = 7, 16, 6, 7, 12, 4, 48, 56, 3
    z, P, _ = synth.head_inputs(n, K, h, w, N, C, seed=31)
    labels = synth.block_labels(n, H, W, C, block=8, seed=32)
    labels[2, :, :30] = 0
    pc = (np.arange(N) // (N // C)).astype(np.int64)
    return z, P, labels, pc, (n, h, w, N, k)


def _image_minima_cpu(z, P, labels, h, w):
    """what adlm_push_scan_images returns for a void mask, from the oracle's distance maps"""
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    n, N = z.shape[0], P.shape[0]
    H, W = labels.shape[1:]
    rl, cl = nearest.pil_nearest_lut(H, h), nearest.pil_nearest_lut(W, w)
    packed = np.full((n, N), _lib.PUSH_NONE, dtype=np.int64)
    for i in range(n):
        d = orc.l2_distances(z[i:i + 1], P)[0].reshape(N, h * w)
        ok = (labels[i][rl][:, cl] != 0).reshape(-1)
        for j in range(N):
            if ok.any():
                dj = np.where(ok, d[j], np.inf)
                a = int(dj.argmin())
                packed[i, j] = (int(np.float32(dj[a]).view(np.uint32)) << 32) | a
    return packed


def _nearest_worker(rank, size):
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    z, P, labels, pc, (n, h, w, N, k) = _nearest_case()
    lo, hi = parallel.image_range(n, rank, size)
    packed = torch.from_numpy(_image_minima_cpu(z[lo:hi], P, labels[lo:hi], h, w))
    out = nearest.top_k_and_labels(packed, k, lambda i: labels[lo + i], pc, h, w, image_base=lo, n_total=n)
    return {key: v.copy() for key, v in out.items()}


def test_k_nearest_sharded_equals_single_process_and_oracle():
    from adlm_inherently_interpretable_segmentation_b200 import nearest
    z, P, labels, pc, (n, h, w, N, k) = _nearest_case()
    packed = torch.from_numpy(_image_minima_cpu(z, P, labels, h, w))
    single = nearest.top_k_and_labels(packed, k, lambda i: labels[i], pc, h, w)
    for out in spawn(_nearest_worker):                # identical on both ranks and equal to the unsharded result
        for key in single:
            assert np.array_equal(out[key], single[key]), key
    ol, od, oi, op = orc.find_k_nearest(z, P, labels, pc, k)
    assert np.array_equal(single["labels"], ol) and np.array_equal(single["images"], oi)
    assert np.array_equal(single["patch_i"] * w + single["patch_j"], op)
    assert np.allclose(single["distances"], od, rtol=1e-6)
