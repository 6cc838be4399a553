#!/usr/bin/env python
"""One push scan over a resident synthetic shard (for ncu): python tools/push_once.py [n_img] [precision]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import push, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
prec = sys.argv[2] if len(sys.argv) > 2 else "tf32x3"
dev = torch.device("cuda", 0)
K, h, w, H, W, N, C = 64, 129, 257, 1024, 2048, 190, 19
z = torch.sigmoid(torch.randn(n, K, h, w, device=dev))
labels = torch.from_numpy(synth.block_labels(8, H, W, C, block=32, seed=1)).to(dev).repeat(n // 8, 1, 1).contiguous()
P = torch.rand(N, K, device=dev)
pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
for _ in range(2):
    s = push.PushSearch(P, pc, precision=prec)
    s.scan(z, labels, 0)
torch.cuda.synchronize()
print("winners", int((s.packed != 0x7fffffffffffffff).sum()))
