#!/usr/bin/env python
"""k nearest patches (find_nearest.py) over a resident synthetic shard: python tools/nearest_bench.py [n_img]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import nearest, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
K, h, w, H, W, N, C, k = 64, 129, 257, 1024, 2048, 190, 19, 5
z = torch.sigmoid(torch.randn(n, K, h, w, device=dev))
labels = torch.from_numpy(synth.block_labels(8, H, W, C, block=32, seed=1)).to(dev).repeat(n // 8, 1, 1).contiguous()
labels_host = labels.cpu().numpy()
P = torch.rand(N, K, device=dev)
pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = nearest.find_k_nearest_tensors(z, labels, P, pc, k=k, labels_host=labels_host)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
print(f"{n} images {h}x{w}, N={N}, k={k}, 3xTF32: {dt * 1e3:.1f} ms total = {dt / n * 1e6:.0f} us per image "
      f"(scan launches + void masks + top-k + {N * k} label regions on the host)")
