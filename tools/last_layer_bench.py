#!/usr/bin/env python
"""Post-push last-layer phase (train_and_test.last_only): step time with all gradients vs need_grads="last_layer",
cfg2 head shape, per precision.   python tools/last_layer_bench.py"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import head, synth  # noqa: E402
dev = torch.device("cuda", 0)
B, K, h, w, N, C = 16, 64, 65, 65, 190, 19
zs = [torch.sigmoid(torch.randn(B, K, h, w, device=dev)) for _ in range(8)]
_, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=5)
P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
tg = torch.from_numpy(synth.block_labels(B, h, w, C, block=8, seed=6).astype(np.int32)).to(dev)
pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
for prec in ("tf32", "tf32x3", "fp32"):
    for mode in (True, "last_layer"):
        ws = None
        for i in range(3):
            r = head.fused_step(zs[i % 8], P, W, tg, pc, precision=prec, need_grads=mode, workspace=ws)
            ws = r.workspace
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(50):
            head.fused_step(zs[i % 8], P, W, tg, pc, precision=prec, need_grads=mode, workspace=ws)
        e1.record()
        torch.cuda.synchronize()
        print(f"{prec:7s} need_grads={mode!s:10s} {e0.elapsed_time(e1) / 50 * 1e3:8.1f} us per step")
