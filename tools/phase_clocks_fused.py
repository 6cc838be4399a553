#!/usr/bin/env python
"""Phase timeline of CTA 0 of k_tc_bwd_fused (library built with ADLM_EXTRA_NVCC_FLAGS=-DADLM_PHASE_CLOCKS).
    python tools/phase_clocks_fused.py [B]"""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import _lib, head, synth  # noqa: E402
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda", 0)
K, h, w, N, C = 64, 65, 65, 190, 19
_, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
z = torch.sigmoid(torch.randn(B, K, h, w, device=dev))
dl = torch.randn(B, h * w, C, device=dev) / (B * h * w)
lib = _lib.load()
for rep in range(3):
    head.head_backward(z, P, W, dlogits=dl, precision="tf32")
    torch.cuda.synchronize()
buf = (ctypes.c_longlong * 192)()
assert lib.adlm_debug_dp_phase_clocks(buf) == 0
a = np.array(buf[:]).reshape(3, 64)
t0 = a[a > 0].min()
names = {0: ["start", "kfree+seen", "K done", "free", "T done", "prefetched", "writer: free seen"],
         1: ["gs b0", "A0 issued", "D next", "-", "gs b1", "A1 issued", "D next"],
         2: ["wait acc0", "acc0", "blk0 done", "-", "wait acc1", "acc1", "blk1 done"]}
for role, rn in enumerate(["builder", "mma", "epilogue w0"]):
    print(rn)
    for it in range(7):
        row = a[role, 1 + it * 8: 8 + it * 8] - t0
        print(f"   tile {it}: " + "  ".join(f"{n}={v}" for n, v in zip(names[role], row) if n != "-" and v > -t0))

