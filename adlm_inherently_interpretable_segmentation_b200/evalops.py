"""Evaluation epilogue of the reference on the GPU.

segmentation/eval_valid.py:171-199 upsamples the logits (B,C,h,w) and the prototype distances (B,N,h,w) to the
annotation size with F.interpolate(mode='bilinear', align_corners=False), takes argmax / argmin over the planes,
copies everything to the host and counts correct pixels, per-class intersections / unions and how often each
prototype is the nearest one of its predicted class.  upsample_argmax() does all of that in one kernel
(adlm_eval_upsample): the upsampled stacks (1.6 GB per Cityscapes image) are never materialised.
"""
from ctypes import c_int

import numpy as np
import torch

from . import _lib


def source_tables(in_size, out_size):
    """torch's area_pixel_compute_source_index for align_corners=False, in float32 like the CUDA kernel of
    F.interpolate: src = max(scale * (dst + 0.5) - 0.5, 0), scale = in / out.
    -> (index int32 [out], lambda float32 [out]); the second tap is min(index + 1, in - 1)."""
    scale = np.float32(in_size) / np.float32(out_size)
    dst = np.arange(out_size, dtype=np.float32)
    src = np.maximum(scale * (dst + np.float32(0.5)) - np.float32(0.5), np.float32(0.0)).astype(np.float32)
    idx = np.minimum(src.astype(np.int32), in_size - 1)
    lam = (src - idx.astype(np.float32)).astype(np.float32)
    return idx, lam


def tile_patch_extent(idx, in_size, tile):
    """largest number of source rows (columns) an output tile of `tile` rows (columns) touches"""
    out = len(idx)
    best = 1
    for a in range(0, out, tile):
        b = min(a + tile, out) - 1
        best = max(best, min(int(idx[b]) + 1, in_size - 1) - int(idx[a]) + 1)
    return best


_TABLES = {}


def _tables(h, w, H, W, device):
    key = (h, w, H, W, str(device))
    if key not in _TABLES:
        y0, ly = source_tables(h, H)
        x0, lx = source_tables(w, W)
        _TABLES[key] = (torch.from_numpy(y0).to(device), torch.from_numpy(ly).to(device),
                        torch.from_numpy(x0).to(device), torch.from_numpy(lx).to(device),
                        tile_patch_extent(y0, h, _lib.EVAL_TILE_Y), tile_patch_extent(x0, w, _lib.EVAL_TILE_X))
    return _TABLES[key]


def counter_views(counters, C, N):
    """split the flat counter vector of upsample_argmax into the reference's quantities"""
    c = counters
    return dict(correct_pixels=c[0], total_pixels=c[1], CLS_I=c[2:2 + C], CLS_U=c[2 + C:2 + 2 * C],
                proto_counts=c[2 + 2 * C:2 + 2 * C + N])


def upsample_argmax(logits, size, distances=None, labels=None, proto_class=None, counters=None,
                    want_pred=True, want_nearest=True):
    """logits (B,h,w,C) as returned by PPNet.forward; distances (B,N,h,w) or None; size = (H, W) of the annotation.
    -> dict(pred (B,H,W) uint8 = torch.argmax(F.interpolate(logits.permute(0,3,1,2), size, 'bilinear'), 1),
            nearest (B,H,W) int16 = torch.argmin(F.interpolate(distances, size, 'bilinear'), 1) | None,
            counters (2 + 2C + N,) int64 | None)
    labels (B,H,W) uint8 (0 = void, class + 1 otherwise) switches the accuracy / IoU counters on; proto_class (N,)
    int32 the prototype-usage counters; `counters` accumulates across calls when passed back in."""
    lib = _lib.load()
    if not logits.is_cuda or logits.dtype != torch.float32 or logits.dim() != 4:
        raise ValueError("logits must be a CUDA float32 tensor (B,h,w,C): the evaluation kernel has no CPU path")
    logits = logits.contiguous()
    B, h, w, C = logits.shape
    H, W = int(size[0]), int(size[1])
    dev = logits.device
    N = 0
    if distances is not None:
        if not distances.is_cuda or distances.dtype != torch.float32 or tuple(distances.shape[0:1] + distances.shape[2:]) != (B, h, w):
            raise ValueError("distances must be a CUDA float32 tensor (B,N,h,w)")
        distances = distances.contiguous()
        N = distances.shape[1]
    if labels is not None:
        if tuple(labels.shape) != (B, H, W):
            raise ValueError(f"labels must be (B,H,W) = {(B, H, W)}")
        labels = labels.to(device=dev, dtype=torch.uint8).contiguous()
    if proto_class is not None:
        proto_class = proto_class.to(device=dev, dtype=torch.int32).contiguous()
    y0, ly, x0, lx, pr, pcols = _tables(h, w, H, W, dev)
    want_counts = labels is not None or (proto_class is not None and distances is not None)
    with torch.cuda.device(dev):
        pred = torch.empty((B, H, W), device=dev, dtype=torch.uint8) if want_pred else None
        nearest = torch.empty((B, H, W), device=dev, dtype=torch.int16) if (want_nearest and distances is not None) else None
        if want_counts and counters is None:
            counters = torch.zeros(2 + 2 * C + N, device=dev, dtype=torch.int64)
        rc = lib.adlm_eval_upsample(_lib.ptr(logits), _lib.ptr(distances), B, h, w, C, c_int(max(N, 1)), H, W,
                                    _lib.ptr(y0), _lib.ptr(ly), _lib.ptr(x0), _lib.ptr(lx), pr, pcols,
                                    _lib.ptr(pred), _lib.ptr(nearest), _lib.ptr(labels), _lib.ptr(proto_class),
                                    _lib.ptr(counters if want_counts else None), _lib.stream_ptr())
    _lib.check(rc, "adlm_eval_upsample")
    return dict(pred=pred, nearest=nearest, counters=counters if want_counts else None)
