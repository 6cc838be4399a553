#!/usr/bin/env python
"""Legacy min-pool classification branch (model.py:284-299): the reference's op sequence with stock torch on the same
GPU (two 1x1 convs, relu, max_pool2d, then the (B,N) tail; forward + backward) against adlm_min_pool_fwd/bwd.
python tools/minpool_bench.py [B]"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import head  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda", 0)
K, h, w, N, C = 64, 65, 65, 190, 19
z = torch.sigmoid(torch.randn(B, K, h, w, device=dev)).requires_grad_(True)
P = torch.rand(N, K, 1, 1, device=dev, requires_grad=True)
W = torch.randn(C, N, device=dev, requires_grad=True)
ones = torch.ones(N, K, 1, 1, device=dev)
label = torch.randint(0, C, (B,), device=dev)


gmin = torch.randn(B, N, device=dev) / (B * N)
TAIL = True


def tail(mind):
    """the (B,N)-sized rest of the step (similarity, last layer, cross entropy, a cluster cost): stock torch in both"""
    if not TAIL:
        return mind.backward(gmin)
    act = torch.log((mind + 1) / (mind + 1e-4))
    (F.cross_entropy(act @ W.t(), label) + 0.8 * mind.min(dim=1).values.mean()).backward()


def stock():
    x2 = F.conv2d(z ** 2, ones)
    xp = F.conv2d(z, P)
    d = F.relu(x2 - 2 * xp + (P ** 2).sum(dim=(1, 2, 3)).view(-1, 1, 1))
    mind = -F.max_pool2d(-d, kernel_size=(h, w)).view(-1, N)
    tail(mind)


def fused():
    mind, _ = head.MinPoolFunction.apply(z, P, None)
    tail(mind)


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        z.grad = None; P.grad = None; W.grad = None
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for TAIL in (False, True):
    what = "min-pool + loss tail" if TAIL else "min-pool only (given d loss / d min_distances)"
    for tf32 in (True, False):
        torch.backends.cudnn.allow_tf32 = tf32
        print(f"{what}: stock torch (cudnn allow_tf32={tf32}): {timeit(stock):.0f} us per fwd+bwd, "
              f"B={B} {K}x{h}x{w}, N={N}")
    print(f"{what}: adlm_min_pool_fwd/bwd (3xTF32 scan): {timeit(fused):.0f} us per fwd+bwd")
z.grad = None; P.grad = None; W.grad = None
stock(); gz, gP = z.grad.clone(), P.grad.clone()
z.grad = None; P.grad = None; W.grad = None
fused()
print("max |dz - stock| / max|dz| =", ((z.grad - gz).abs().max() / gz.abs().max()).item(),
      " dP:", ((P.grad - gP).abs().max() / gP.abs().max()).item())
