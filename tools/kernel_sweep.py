#!/usr/bin/env python
"""Per-kernel timing sweep over the batch size: separates fixed cost (launch, per-CTA setup) from per-tile cost.

    python tools/kernel_sweep.py [--precision tf32] [--batches 16,64,256] [--reps 20]

Prints one line per (batch, phase): CUDA-event milliseconds of head_forward / head_backward / fused_step and
the library's own per-launch brackets (adlm_timing_begin/end) for the tagged kernels.
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import _lib, head, synth  # noqa: E402


def ev_time(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--batches", default="16,64,256")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--once", action="store_true", help="one call of each phase at the first batch size (for ncu)")
    ap.add_argument("--protos", default="190,19", help="N,C: 190,19 = Cityscapes (cfg2), 30,3 = mds_new pancreas (cfg4)")
    ap.add_argument("--infer", action="store_true", help="inference only: logits + per-pixel argmax (cfg4)")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    K, h, w = 64, 65, 65
    N, C = [int(x) for x in a.protos.split(",")]
    _, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
    P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
    pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
    for B in [int(x) for x in a.batches.split(",")]:
        g = torch.Generator(device=dev).manual_seed(1)
        z = torch.sigmoid(torch.randn(B, K, h, w, device=dev, generator=g))
        tgt = torch.from_numpy(synth.block_labels(B, h, w, C, block=8, seed=3).astype(np.int32)).to(dev)
        dl = torch.randn(B, h * w, C, device=dev, generator=g) / (B * h * w)
        M = B * h * w
        fwd = lambda: head.head_forward(z, P, W, precision=a.precision, want_dist=False, want_argmax=True)
        bwd = lambda: head.head_backward(z, P, W, dlogits=dl, precision=a.precision)
        step = lambda: head.fused_step(z, P, W, tgt, pc, precision=a.precision)
        if a.once:
            fwd(); bwd(); step()
            torch.cuda.synchronize()
            return
        fb, bb = 4 * K + 4 * C + (4 if a.infer else 0), 8 * K + 4 * C   # algorithmic bytes / pixel (+4: argmax)
        phases = (("fwd", fwd, fb * M),) if a.infer else (("fwd", fwd, fb * M), ("bwd", bwd, bb * M), ("step", step, (fb + bb) * M))
        for name, fn, nbytes in phases:
            ms = ev_time(fn, a.reps)
            print(f"B={B:4d} M={M:8d} {name:5s} {ms * 1e3:9.1f} us  {M / ms / 1e3:8.1f} Mpix/s  {nbytes / ms / 1e6:7.1f} GB/s(alg)",
                  flush=True)
        if a.infer:
            continue
        for tag, nm in ((_lib.K_STEP_PREPARE, "k_prepare"), (_lib.K_HEAD_FWD, "k_fwd"), (_lib.K_KLD, "k_kld"),
                        (_lib.K_HEAD_BWD, "k_bwd")):
            with _lib.kernel_timing(tag, 64) as t:
                for _ in range(5):
                    step()
                torch.cuda.synchronize()
            print(f"B={B:4d} inside step {nm}: {np.mean(t.ms) * 1e3:8.1f} us", flush=True)


if __name__ == "__main__":
    main()
