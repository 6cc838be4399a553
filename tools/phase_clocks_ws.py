#!/usr/bin/env python
"""Phase timeline of CTA 0 of k_tc_fwd_ws (library built with ADLM_EXTRA_NVCC_FLAGS=-DADLM_PHASE_CLOCKS):
    python tools/phase_clocks_ws.py [B [N C]]        # N C = 30 3: the cfg4 head"""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import _lib, head, synth  # noqa: E402
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda", 0)
K, h, w, N, C = 64, 65, 65, 190, 19
if len(sys.argv) > 3:
    N, C = int(sys.argv[2]), int(sys.argv[3])
_, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
z = torch.sigmoid(torch.randn(B, K, h, w, device=dev))
lib = _lib.load()
for rep in range(3):
    head.head_forward(z, P, W, precision="tf32", want_dist=False, want_argmax=True)
    torch.cuda.synchronize()
buf = (ctypes.c_longlong * 256)()
assert lib.adlm_debug_ws_phase_clocks(buf) == 0
a = np.array(buf[:]).reshape(4, 64)
t0 = a[:, 0].min()
names = {0: ["wait", "D1 ready", "E1 done", "E2 done"], 1: ["wait", "D1 ready", "E1 done", "E2 done"],
         2: ["start", "stage free", "written", "prefetched"], 3: ["start", "A full", "MMA1 issued", "MMA2 issued"]}
for role, rn in enumerate(["epilogue half0", "epilogue half1", "producer", "mma"]):
    print(f"{rn}: setup done at {a[role, 0] - t0}")
    for it in range(12):
        row = a[role, 1 + it * 4: 5 + it * 4] - t0
        if row[0] <= 0: break
        print(f"   tile {it}: " + "  ".join(f"{n}={v}" for n, v in zip(names[role], row)))
