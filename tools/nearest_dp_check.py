#!/usr/bin/env python
"""k nearest patches with the images sharded over GPUs equals the single-GPU result:
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/nearest_dp_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import nearest, parallel, synth  # noqa: E402

rank, local, size = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n, K, h, w, H, W, N, C, k = 24, 64, 65, 129, 512, 1024, 190, 19, 5
g = torch.Generator(device=dev).manual_seed(3)                      # same data on every rank
z = torch.sigmoid(torch.randn(n, K, h, w, device=dev, generator=g))
P = torch.rand(N, K, device=dev, generator=g)
labels_np = synth.block_labels(n, H, W, C, block=32, seed=4)
labels = torch.from_numpy(labels_np).to(dev)
pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
whole = nearest.find_k_nearest_tensors(z, labels, P, pc, k=k, labels_host=labels_np)
lo, hi = parallel.image_range(n, rank, size)
part = nearest.find_k_nearest_tensors(z[lo:hi], labels[lo:hi], P, pc, k=k, labels_host=labels_np[lo:hi],
                                      image_base=lo, n_total=n)
ok = all(np.array_equal(whole[key], part[key]) for key in whole)
t = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"k-nearest over {size} GPUs ({n} images, shards of {hi - lo}): "
          f"{'identical to the single-GPU result on every rank' if int(t) else 'MISMATCH'}")
    print("NEAREST_DP_CHECK", "PASS" if int(t) else "FAIL")
dist.destroy_process_group()
