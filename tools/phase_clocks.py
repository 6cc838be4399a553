#!/usr/bin/env python
"""Phase timeline of CTA 0 of k_tc_fwd (needs a library built with ADLM_EXTRA_NVCC_FLAGS=-DADLM_PHASE_CLOCKS):

    ADLM_EXTRA_NVCC_FLAGS=-DADLM_PHASE_CLOCKS python -m adlm_inherently_interpretable_segmentation_b200.build --force
    python tools/phase_clocks.py [B]
"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import _lib, head, synth  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
prec = sys.argv[2] if len(sys.argv) > 2 else "tf32"
dev = torch.device("cuda", 0)
K, h, w, N, C = 64, 65, 65, 190, 19
_, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
z = torch.sigmoid(torch.randn(B, K, h, w, device=dev))
lib = _lib.load()
names = ["tile start", "converted+bar", "MMA1 issued", "prefetch issued", "MMA1 done", "epilogue1 done", "MMA2 done", "stored"]
for rep in range(3):
    head.head_forward(z, P, W, precision=prec, want_dist=False, want_argmax=True)
    torch.cuda.synchronize()
buf = (ctypes.c_longlong * 128)()
assert lib.adlm_debug_phase_clocks(buf) == 0
a = np.array(buf[:]).reshape(2, 64)
for wg in range(2):
    t0 = a[wg, 0]
    print(f"WG{wg}: setup {a[wg, 1] - t0} cycles; kernel end at {a[wg, 63] - t0}")
    for t in range(4):
        base = 2 + 8 * t
        if base + 7 >= 63 or a[wg, base] < t0:
            break
        row = a[wg, base:base + 8] - t0
        if row[0] <= 0 or (t > 0 and row[0] < a[wg, base - 1] - t0):
            break
        print(f"  tile {t}: " + "  ".join(f"{n}={v}" for n, v in zip(names, row)))
