#!/usr/bin/env python
"""Golden vectors for the evaluation epilogue.

TEST INFRASTRUCTURE ONLY.  The reference's evaluation loop (segmentation/eval_valid.py:165-199) is inline code
inside run_evaluation (it needs the datasets), so this script runs the same torch calls the reference makes —
F.interpolate(..., mode='bilinear', align_corners=False), torch.argmax, torch.argmin — on seeded inputs with the
torch build in this container and stores inputs + outputs in tests/golden/eval_small.npz.  The arithmetic lives in
torch (third-party, pinned by the reference at 1.10/1.11; here torch 2.11 CPU); oracle/head_oracle.py restates it
and tests/test_oracle_golden.py checks the restatement against this fixture.

Usage: python oracle/make_golden_eval.py
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from adlm_inherently_interpretable_segmentation_b200 import synth  # noqa: E402


def main():
    rng = np.random.default_rng(synth.BASE_SEED + 900)
    B, h, w, C, N, H, W = 2, 9, 13, 4, 20, 70, 100           # odd map, non-integer scale (like 129x257 -> 1024x2048)
    logits = rng.normal(size=(B, h, w, C)).astype(np.float32)
    dist = rng.uniform(0.0, 8.0, size=(B, N, h, w)).astype(np.float32)
    ann = synth.block_labels(B, H, W, C, block=9, seed=synth.BASE_SEED + 901).astype(np.uint8)   # 0 = void
    proto_class = (np.arange(N) // (N // C)).astype(np.int32)
    tl = torch.from_numpy(logits).permute(0, 3, 1, 2)
    td = torch.from_numpy(dist)
    out = {}
    ups_l, ups_d, preds, nears = [], [], [], []
    for i in range(B):                                        # eval_valid.py:165-176, one sample at a time
        lg = F.interpolate(tl[i:i + 1], size=(H, W), mode='bilinear', align_corners=False)[0]
        ds = F.interpolate(td[i:i + 1], size=(H, W), mode='bilinear', align_corners=False)[0]
        ups_l.append(lg.numpy()); ups_d.append(ds.numpy())
        preds.append(torch.argmax(lg, dim=0).numpy()); nears.append(torch.argmin(ds, dim=0).numpy())
    pred, near = np.stack(preds), np.stack(nears)
    correct = int(np.sum(((pred + 1) == ann) & (ann != 0)))  # :178-180
    total = int(np.sum(ann != 0))
    I, U = np.zeros(C, np.int64), np.zeros(C, np.int64)
    for c in range(C):                                        # :182-189
        pr, gt = pred == c, ann == c + 1
        I[c] = np.sum(pr & gt)
        U[c] = np.sum((pr | gt) & (ann != 0))
    counts = np.zeros(N, np.int64)                            # :192-197
    for c in range(C):
        is_class_proto = (pred == c) & (proto_class[near] == c)
        for n in np.nonzero(proto_class == c)[0]:
            counts[n] += np.sum(is_class_proto & (near == n))
    path = os.path.join(ROOT, "tests", "golden", "eval_small.npz")
    np.savez_compressed(path, logits=logits, dist=dist, ann=ann, proto_class=proto_class,
                        up_logits=np.stack(ups_l), up_dist=np.stack(ups_d), pred=pred.astype(np.int16),
                        nearest=near.astype(np.int16), correct_pixels=correct, total_pixels=total, CLS_I=I, CLS_U=U,
                        proto_counts=counts, torch_version=torch.__version__)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
