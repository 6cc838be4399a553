#!/usr/bin/env python
"""Multi-GPU check of the peer-memory data-parallel step (adlm_step_fwd_bwd_dp) against the whole-batch
single-GPU step and against the NCCL transport; then times both transports.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/dp_check.py [--steps 200]
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from adlm_inherently_interpretable_segmentation_b200 import head, parallel, synth  # noqa: E402


def rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--per-rank", type=int, default=16)
    a = ap.parse_args()
    rank, local, size = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    K, h, w, N, C = 64, 65, 65, 190, 19
    Bl = a.per_rank
    B = Bl * size
    g = torch.Generator(device=dev).manual_seed(7)                 # same whole batch on every rank
    z = torch.sigmoid(torch.randn(B, K, h, w, device=dev, generator=g))
    tgt = torch.from_numpy(synth.block_labels(B, h, w, C, block=8, seed=3).astype(np.int32)).to(dev)
    _, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
    P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
    pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
    wts = (1.0, 0.25, 1e-4)
    full = head.fused_step(z, P, W, tgt, pc, weights=wts, precision="tf32")
    lo, hi = rank * Bl, (rank + 1) * Bl
    zs, ts = z[lo:hi].contiguous(), tgt[lo:hi].contiguous()
    ok = True
    for name, exch in (("peer-memory", "auto"), ("nccl", None)):
        res = None
        for _ in range(5):                                         # several steps: both slot parities
            res = parallel.dp_fused_step((P, W), zs, ts, pc, weights=wts, precision="tf32", exchange=exch)
        torch.cuda.synchronize()
        e = {"dP": rel(res.dP.reshape(N, K), full.dP.reshape(N, K)), "dW": rel(res.dW, full.dW),
             "dz": rel(res.dz, full.dz[lo:hi]), "scalars": rel(res.scalars[:7], full.scalars[:7]),
             "timeouts": float(res.scalars[7])}
        flat = torch.cat([res.dP.reshape(-1), res.dW.reshape(-1)])
        gathered = [torch.empty_like(flat) for _ in range(size)]
        dist.all_gather(gathered, flat)
        e["bit_identical_across_ranks"] = all(torch.equal(gathered[0], x) for x in gathered)
        good = max(e["dP"], e["dW"], e["dz"]) <= 1e-3 and e["scalars"] <= 1e-5 and e["timeouts"] == 0
        ok = ok and good
        if rank == 0:
            print(f"[{name}] world={size} vs whole-batch single-GPU step: {e} -> {'OK' if good else 'MISMATCH'}", flush=True)
        # timing
        for _ in range(10):
            parallel.dp_fused_step((P, W), zs, ts, pc, weights=wts, precision="tf32", exchange=exch)
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            parallel.dp_fused_step((P, W), zs, ts, pc, weights=wts, precision="tf32", exchange=exch)
        e1.record()
        dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"[{name}] {float(t) * 1e3:.1f} us per data-parallel step (max over ranks, {a.steps} steps)", flush=True)
    # single-GPU step on one shard for reference
    for _ in range(10):
        head.fused_step(zs, P, W, ts, pc, weights=wts, precision="tf32")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        head.fused_step(zs, P, W, ts, pc, weights=wts, precision="tf32")
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"[single] {e0.elapsed_time(e1) / a.steps * 1e3:.1f} us per step of one shard without any exchange", flush=True)
        print("DP_CHECK", "PASS" if ok else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
