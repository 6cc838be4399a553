#!/usr/bin/env python
"""Determinism and agreement of the fused backward with the two-kernel backward (debugging aid)."""
import os, sys, subprocess
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import head
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
K, h, w, N, C = 64, 65, 65, 190, 19
g = torch.Generator(device="cuda").manual_seed(5)
z = torch.sigmoid(torch.randn(B, K, h, w, device="cuda", generator=g))
P = torch.rand(N, K, device="cuda", generator=g)
W = torch.randn(C, N, device="cuda", generator=g) * 0.3
dl = torch.randn(B, h * w, C, device="cuda", generator=g) / (B * h * w)
outs = [head.head_backward(z, P, W, dlogits=dl, precision="tf32") for _ in range(4)]
torch.cuda.synchronize()
for i in range(1, 4):
    for name, a, b in zip(("dz", "dP", "dW"), outs[0], outs[i]):
        d = (a - b).abs()
        print(f"run {i} vs 0: {name} equal={torch.equal(a, b)} max diff {float(d.max()):.3e} (max |x| {float(a.abs().max()):.3e}) n_diff {int((d > 0).sum())}")
if os.environ.get("ADLM_BWD_TWO_KERNELS") != "1":
    torch.save([o.cpu() for o in outs[0]], "/tmp/fused.pt")
    env = dict(os.environ, ADLM_BWD_TWO_KERNELS="1")
    subprocess.run([sys.executable, __file__, str(B)], env=env)
else:
    ref = torch.load("/tmp/fused.pt")
    for name, a, b in zip(("dz", "dP", "dW"), ref, outs[0]):
        d = (a.cuda() - b).abs()
        print(f"fused vs two-kernel: {name} max diff {float(d.max()):.3e} rel {float(d.max() / b.abs().max()):.3e}")
    sub = head.head_backward(z[32:64], P, W, dlogits=dl[32:64], precision="tf32")
    d = (sub[0] - outs[0][0][32:64]).abs()
    print("two-kernel sub-batch dz equal:", torch.equal(sub[0], outs[0][0][32:64]))
