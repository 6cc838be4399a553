#!/usr/bin/env python
"""Add-on Sigmoid fused into the head vs stock torch sigmoid around the head step.

    python tools/sigmoid_fusion_bench.py [--batches 16,256]
unfused: z = torch.sigmoid(pre); step(z); dpre = dz * z * (1 - z)   (what autograd does around the head)
fused:   step(pre, pre_sigmoid=True)                                (adlm_step_fwd_bwd_pre)
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import head, synth  # noqa: E402


def ev_time(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="16,256")
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    K, h, w, N, C = 64, 65, 65, 190, 19
    _, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
    P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
    pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
    for B in [int(x) for x in a.batches.split(",")]:
        pre = torch.randn(B, K, h, w, device=dev) * 1.5
        tgt = torch.from_numpy(synth.block_labels(B, h, w, C, block=8, seed=3).astype(np.int32)).to(dev)

        def unfused():
            z = torch.sigmoid(pre)
            r = head.fused_step(z, P, W, tgt, pc, precision="tf32")
            return r.dz * z * (1 - z)

        def fused():
            return head.fused_step(pre, P, W, tgt, pc, precision="tf32", pre_sigmoid=True).dz

        def plain():
            return head.fused_step(pre, P, W, tgt, pc, precision="tf32").dz

        tu, tf, tp = ev_time(unfused, a.reps), ev_time(fused, a.reps), ev_time(plain, a.reps)
        print(f"B={B:4d}: torch.sigmoid + step + sigmoid backward {tu:8.1f} us | fused {tf:8.1f} us "
              f"({tu / tf:.2f}x) | step alone on given z {tp:8.1f} us", flush=True)


if __name__ == "__main__":
    main()
