#!/usr/bin/env python
"""BASELINE configs[4]: 1000 prototypes x 256-d, 100 classes on a full 1024 x 2048 feature map (one map per GPU per
iteration), forward + per-pixel argmax through adlm_head_fwd_big; reports algorithmic TFLOP/s (2NK + 2NC per pixel,
SURVEY 8(d)) against the TF32 matmul peak measured in the same process.
    python tools/cfg5_bench.py [--hw 1024x2048] [--reps 5]"""
import argparse, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import _lib, head  # noqa: E402


def tf32_peak(dev):
    n = 8192
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    a, b = torch.randn(n, n, device=dev), torch.randn(n, n, device=dev)
    for _ in range(3):
        a @ b
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    torch.backends.cuda.matmul.allow_tf32 = old
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hw", default="1024x2048")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--protos", default="1000,256,100")
    a = ap.parse_args()
    h, w = [int(x) for x in a.hw.split("x")]
    N, K, C = [int(x) for x in a.protos.split(",")]
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(5)
    z = torch.empty(1, K, h, w, device=dev)
    for k0 in range(0, K, 32):
        z[:, k0:k0 + 32].normal_(generator=g).sigmoid_()
    P = torch.rand(N, K, device=dev, generator=g)
    W = torch.randn(C, N, device=dev, generator=g) * 0.1
    fn = lambda: head.head_forward(z, P, W, precision="tf32", want_dist=False, want_argmax=True)
    out = fn()
    torch.cuda.synchronize()
    with _lib.kernel_timing(_lib.K_HEAD_FWD, 64) as t:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
    ms_call = e0.elapsed_time(e1) / a.reps
    ms_k = float(np.mean(t.ms))
    M = h * w
    flops = (2.0 * N * K + 2.0 * N * C) * M
    peak = tf32_peak(dev)
    # spot check against torch on a strip of pixels
    sl = slice(0, 4096)
    zf = z[0].reshape(K, -1)[:, sl].t().double()
    d = ((zf[:, None, :] - P.double()[None]) ** 2).sum(-1)
    ref = (torch.log((d + 1) / (d + 1e-4)) @ W.double().t()).float()
    got = out["logits"].reshape(-1, C)[sl]
    err = float((got - ref).abs().max() / ref.abs().max())
    agree = float((got.argmax(1) == ref.argmax(1)).float().mean())
    print(json.dumps({"workload": f"configs[4]: {N} prototypes x {K}-d, {C} classes, one {h}x{w} map, forward + argmax",
                      "ms_per_call": ms_call, "kernel_ms": ms_k, "gpixels_per_s": M / (ms_k * 1e-3) / 1e9,
                      "algorithmic_tflops": flops / (ms_k * 1e-3) / 1e12, "tf32_peak_tflops": peak,
                      "frac_of_tf32_peak": flops / (ms_k * 1e-3) / 1e12 / peak,
                      "logits_rel_err_vs_fp64_strip": err, "argmax_agreement_strip": agree}))


if __name__ == "__main__":
    main()
