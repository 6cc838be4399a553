#!/usr/bin/env python
"""labels -> patch class mask kernel alone (adlm_patch_class_mask): library-bracketed kernel time and GB/s for
n images of 1024x2048 uint8 labels -> 129x257 patches.   python tools/mask_bench.py [--n 512] [--block 32]
ADLM_MASK_CTA_KERNEL=1 selects the one-CTA-per-patch-row kernel instead of the thread-per-patch one."""
import argparse, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import _lib, push, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=512)
ap.add_argument("--block", type=int, default=32)
a = ap.parse_args()
dev = torch.device("cuda", 0)
H, W, h, w, N, K, C = 1024, 2048, 129, 257, 190, 64, 19
lab = torch.from_numpy(synth.block_labels(12, H, W, C, block=a.block, seed=3)).to(dev)
labels = lab.repeat((a.n + 11) // 12, 1, 1)[:a.n].contiguous()
s = push.PushSearch(torch.rand(N, K, device=dev), torch.arange(N, dtype=torch.int32, device=dev) // 10, precision="fp32")
m = s.class_mask(labels, h, w)
torch.cuda.synchronize()
with _lib.kernel_timing(_lib.K_PATCH_MASK, 16) as t:
    for _ in range(5):
        s.class_mask(labels, h, w)
    torch.cuda.synchronize()
ms = float(np.mean(t.ms))
print(f"n={a.n} block={a.block}: kernel {ms*1e3:.1f} us, {labels.numel() / (ms * 1e-3) / 1e9:.0f} GB/s (labels read), "
      f"checksum {int(m.long().sum())}")
