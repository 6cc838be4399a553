#!/usr/bin/env python
"""Fused evaluation epilogue vs the reference's torch calls on the GPU (eval_valid.py:165-199, one image at a time).

    python tools/eval_bench.py [--images 4]
"""
import argparse
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import evalops  # noqa: E402


def ev(fn, reps):
    for _ in range(2):
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
    ap.add_argument("--images", type=int, default=4)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    B, h, w, C, N, H, W = a.images, 129, 257, 19, 190, 1024, 2048
    logits = torch.randn(B, h, w, C, device=dev)
    dist = torch.rand(B, N, h, w, device=dev) * 10
    ann = torch.randint(0, C + 1, (B, H, W), device=dev, dtype=torch.uint8)
    pc = (torch.arange(N, device=dev) // (N // C)).to(torch.int32)

    def fused():
        return evalops.upsample_argmax(logits, (H, W), distances=dist, labels=ann, proto_class=pc)

    def torch_gpu():                      # the reference's calls, kept on the device (no .cpu() copies)
        out = []
        lp = logits.permute(0, 3, 1, 2)
        for i in range(B):
            lg = F.interpolate(lp[i:i + 1], size=(H, W), mode="bilinear", align_corners=False)[0]
            ds = F.interpolate(dist[i:i + 1], size=(H, W), mode="bilinear", align_corners=False)[0]
            out.append((torch.argmax(lg, 0), torch.argmin(ds, 0)))
        return out

    def torch_ref():                      # ... and with the host copies the reference makes (eval_valid.py:174-176)
        lp = logits.permute(0, 3, 1, 2)
        for i in range(B):
            lg = F.interpolate(lp[i:i + 1], size=(H, W), mode="bilinear", align_corners=False)[0]
            ds = F.interpolate(dist[i:i + 1], size=(H, W), mode="bilinear", align_corners=False)[0]
            torch.argmin(ds, 0).cpu(); ds.cpu(); torch.argmax(lg, 0).cpu()

    t_f = ev(fused, a.reps)
    t_g = ev(torch_gpu, a.reps)
    t_r = ev(torch_ref, max(1, a.reps // 5))
    out_bytes = B * H * W * 3
    print(f"{B} images {h}x{w} -> {H}x{W}, C={C}, N={N}", flush=True)
    print(f"fused kernel                         {t_f / B * 1e3:9.1f} us/image  ({out_bytes / t_f / 1e6:7.1f} GB/s of output planes)")
    print(f"torch interpolate+argmax on device   {t_g / B * 1e3:9.1f} us/image  ({t_g / t_f:6.1f}x)")
    print(f"reference sequence incl. host copies {t_r / B * 1e3:9.1f} us/image  ({t_r / t_f:6.1f}x)")


if __name__ == "__main__":
    main()
