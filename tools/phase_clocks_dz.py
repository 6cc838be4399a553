#!/usr/bin/env python
"""Phase timeline of CTA 0 of k_tc_bwd_dz, both warpgroups (library built with ADLM_EXTRA_NVCC_FLAGS=-DADLM_PHASE_CLOCKS):
    python tools/phase_clocks_dz.py [B]"""
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
buf = (ctypes.c_longlong * 128)()
assert lib.adlm_debug_dz_phase_clocks(buf) == 0
a = np.array(buf[:]).reshape(2, 64)
t0 = a[:, 0].min()
names = ["start", "operands", "D/Gl h0", "epi h0", "D/Gl h1", "epi h1", "dZ done", "stored"]
for wg in range(2):
    print(f"warpgroup {wg}: loop entered at {a[wg, 0] - t0}, left at {a[wg, 63] - t0}")
    for it in range(4):
        row = a[wg, 1 + it * 8: 9 + it * 8] - t0
        if row[0] <= 0:
            break
        print(f"   tile {it}: " + "  ".join(f"{n}={v}" for n, v in zip(names, row)))
