#!/usr/bin/env python
"""Benchmark of the ProtoSeg prototype-head hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp32|tf32|tf32x3]
                    [--workload cfg2|cfg5] [--no-push] [--no-cpu] [--no-extras]

One "step" = one fused training step of the head (forward + CE + KLD + masked L1 + backward,
segmentation/module.py:137-222,236) over one synthetic batch of BASELINE configs[1]'s head shape:
16 images of 64 x 65 x 65 DeepLab features (513x513 crops), 190 prototypes x 64-d, 19 classes.  The
DeepLabv2/ResNet-101 backbone is out of scope (BASELINE north_star) and is not timed.

Prints ONE JSON line (rank 0):
  value          Mpixels/s over all ranks, inputs resident in HBM, CUDA-event timed, max over ranks
  e2e            same metric through the public C-ABI-backed API with HOST buffers: pinned H2D of the step's
                 features + labels and D2H of its results (loss scalars, dz, dP, dW) inside the timed region,
                 triple-buffered over three streams so both PCIe directions overlap the kernels
  roofline       dominant kernel group's achieved HBM GB/s (algorithmic bytes / CUDA-event launch time) vs
                 MEASURED_PEAKS.json, plus `tensor`: the step's algorithmic TFLOP/s against the TF32 matmul peak measured
                 in this run (torch.matmul fp32, allow_tf32, 8192^3 — SURVEY 8(d))
  cpu_baseline   the reference's own PatchClassificationModule._step (oracle/_ref, kind "reference") on this box's host
                 cores; the operator-sequence port (oracle/torch_port.py, kind "port") only if oracle/_ref is absent
  stock_torch_gpu  the reference FORMULATION (model.py:203-221,231-237,161-162: two 1x1 convolutions, relu, log, linear,
                 cross entropy, autograd; with and without the module.py:167-208 KLD loops) on the same B200 with stock
                 torch / cuDNN / cuBLAS, allow_tf32 on and off — the "beat this" number, there being no native kernel
  cfg4           BASELINE configs[3] (mds_new: 30 prototypes, 3 classes, batch 64 of 65x65, inference + per-pixel argmax)
  cfg5           BASELINE configs[4] (1000 prototypes x 256-d, 100 classes, one 1024x2048 map): forward + argmax on the
                 CTA-pair tensor-core kernel, algorithmic TFLOP/s against the TF32 matmul peak measured in this run;
                 `--workload cfg5` makes it the line itself (batch-sharded over the ranks: one map per GPU, no collective)
  push           secondary metric of BASELINE.json: push imgs/s over the 2975-image Cityscapes-train-sized synthetic set,
                 sharded over the ranks (strong scaling: N=1 scans all 2975 images, N=8 372 each), with the pass breakdown

--impl reference times the reference's own _step (oracle/_ref) on the same config on host cores.
Under torchrun (N > 1): batch-sharded weak scaling; the loss normalisers and [dP | dW | loss sums] are exchanged
inside the step's own kernels over peer-mapped memory (adlm_step_fwd_bwd_dp via parallel.dp_fused_step; NCCL
all-reduces only if symmetric memory cannot be set up, or with ADLM_DP_TRANSPORT=nccl) — `config.dp_transport` says which.
"""
import argparse
import json
import os
import sys
import threading
import time

# The contract is ONE JSON line on stdout.  NCCL prints its banner ("NCCL version ...", when the box exports
# NCCL_DEBUG) with its own printf on file descriptor 1, so for the whole run descriptor 1 is pointed at stderr and the
# result line is written to a saved copy of the real stdout.
sys.stdout.flush()
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(B=16, K=64, h=65, w=65, N=190, C=19)          # BASELINE configs[1], head only
CFG4 = dict(B=64, K=64, h=65, w=65, N=30, C=3)           # BASELINE configs[3] (mds_new.gin:6-7)
CFG5 = dict(B=1, K=256, h=1024, w=2048, N=1000, C=100)   # BASELINE configs[4]: large head, one full-resolution map per GPU
WEIGHTS = (1.0, 0.25, 1e-4)                              # cityscapes_kld_imnet.gin:35,38-39
BYTES_FWD = lambda K, C: 4 * K + 4 * C                   # z in, logits out
BYTES_BWD = lambda K, C: 4 * K + 4 * C + 4 * K           # z + dlogits in, dz out
FLOPS_FWD = lambda N, K, C: 2 * N * K + 2 * N * C
FLOPS_BWD = lambda N, K, C: 4 * N * K + 4 * N * C
PUSH_CFG = dict(n_total=2975, K=64, h=129, w=257, H=1024, W=2048, N=190, C=19)  # configs[2]: Cityscapes train size


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=None)
    ap.add_argument("--no-push", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the stock-torch, cfg4, cfg5 and TF32-peak legs")
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg5"],
                    help="cfg2 (default): the fused training step, BASELINE's metric.  cfg5: the large-head forward "
                         "(1000 prototypes x 256-d, 100 classes, one 1024x2048 map per GPU) as the line's own metric")
    ap.add_argument("--push-images", type=int, default=None, help="size of the push set (default 2975)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=float(p["hbm_gbs"]), bf16=float(p.get("bf16_tflops", 0.0)),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, source="fallback (B200_PROFILING.md)")


def measure_tf32_peak(torch, dev):
    """TF32 dense matmul peak of this GPU, the way MEASURED_PEAKS.json measures bf16: torch.matmul
    fp32 with allow_tf32 at 8192^3, best of 10 (burst) and the mean of a 1.5 s back-to-back loop (sustained)"""
    n = 8192
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        for _ in range(3):
            torch.matmul(a, b)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(10, int(1500.0 / best))
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        sus = e0.elapsed_time(e1) / reps
        fl = 2.0 * n ** 3
        return {"tf32_tflops": fl / (best * 1e-3) / 1e12, "tf32_tflops_sustained": fl / (sus * 1e-3) / 1e12,
                "how": "torch.matmul fp32, allow_tf32=True, 8192^3: best of 10 (burst) / back to back for 1.5 s (sustained)"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


# ------------------------------------------------------------------------------------------------ host placement
def bind_near_gpu(index):
    """Pin this process (and, by first touch, its pinned staging buffers) to the CPU cores of the NUMA node the
    GPU hangs off: with one process per GPU, eight ranks staging through one node's memory controller is what
    limits the end-to-end leg.  Returns a short description for the JSON line; a box that hides the topology
    (numa_node = -1, no sysfs) is left alone."""
    if os.environ.get("ADLM_BENCH_BIND", "1") == "0":
        return "off"
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        dev = f"{int(dom, 16):04x}:{rest.lower()}"
        node = int(open(f"/sys/bus/pci/devices/{dev}/numa_node").read())
        if node < 0:
            return f"gpu {dev}: numa_node -1 (not bound)"
        cpulist = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"gpu {dev}: node {node} has no allowed cpus (not bound)"
        os.sched_setaffinity(0, cpus)
        return f"gpu {dev}: node {node}, {len(cpus)} cpus"
    except Exception as e:  # noqa: BLE001 - placement is best effort
        return f"not bound ({type(e).__name__}: {e})"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """samples SM clock + throttle reasons with NVML while the timed region runs"""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80, "sync_boost": 0x10,
                 "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        return False

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------ CPU legs
def host_threads():
    """all host threads this process may use — torchrun exports OMP_NUM_THREADS=1 to every rank, which would otherwise
    throttle the CPU arm (rank 0 is the only rank that runs it)"""
    import torch
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def ref_inputs(n_images):
    """the benchmark batch (or its first n_images) as numpy: features, prototypes, last layer, 513x513 labels"""
    from adlm_inherently_interpretable_segmentation_b200 import synth
    c = CFG
    z, P, W = synth.head_inputs(c["B"], c["K"], c["h"], c["w"], c["N"], c["C"], w_noise=0.1, seed=synth.BASE_SEED)
    labels = synth.block_labels(c["B"], 513, 513, c["C"], block=32, seed=synth.BASE_SEED)
    return z[:n_images], P, W, labels[:n_images]


def reference_step_fn(n_images):
    """(step, pixels, kind, description): the reference's own _step from oracle/_ref when it travelled to this box,
    else the operator-sequence port"""
    c = CFG
    from oracle import ref_runner
    z, P, W, labels = ref_inputs(n_images)
    if ref_runner.available():
        step, info = ref_runner.make_step(z, P, W, labels, weights=WEIGHTS, ignore_void=True)
        return step, info["pixels"], "reference", ("segmentation.module.PatchClassificationModule._step of the "
                                                    "unmodified reference (oracle/_ref) incl. resize_label, KLD loops, "
                                                    "autograd")
    import torch
    from oracle import torch_port as tp
    from adlm_inherently_interpretable_segmentation_b200 import synth
    target = torch.from_numpy(synth.block_labels(n_images, c["h"], c["w"], c["C"], block=8,
                                                 seed=synth.BASE_SEED).astype(np.int64))
    ident = tp.class_identity(c["N"], c["C"])
    zt = torch.from_numpy(z).requires_grad_(True)
    P4 = torch.from_numpy(P).reshape(c["N"], c["K"], 1, 1).requires_grad_(True)
    Wt = torch.from_numpy(W).requires_grad_(True)

    def step():
        for t in (zt, P4, Wt):
            t.grad = None
        return tp.step(zt, P4, Wt, target, ident, ignore_void=True, weights=WEIGHTS)
    return step, n_images * c["h"] * c["w"], "port", "oracle/torch_port.step (reference op sequence; oracle/_ref absent)"


def _pick_reference_sample(seconds_per_step):
    """largest number of the batch's 16 images whose reference _step fits `seconds_per_step` on this host.  The
    reference's cost is QUADRATIC in the batch: autograd of every `dist[b, p]` slice taken in the KLD loops
    (module.py:176-183) zero-fills a gradient the size of the whole (B,N,h,w) distance map (51 MB at batch 16, 3040 slices
    per step), so it is probed at 4 images and extrapolated with (n/4)^2."""
    step, pixels, kind, what = reference_step_fn(4)
    step()
    t0 = time.perf_counter()
    step()
    t4 = time.perf_counter() - t0
    for n in (16, 8):
        if t4 * (n / 4.0) ** 2 <= seconds_per_step:
            return (n,) + reference_step_fn(n)
    if t4 <= seconds_per_step:
        return 4, step, pixels, kind, what
    for n in (2, 1):
        if t4 * (n / 4.0) ** 2 <= seconds_per_step or n == 1:
            return (n,) + reference_step_fn(n)


def _sample_text(n_img, what):
    if n_img == CFG["B"]:
        return f"each step = the full batch of 16 images: {what}"
    return (f"each step = the first {n_img} of the batch's 16 images (the reference step costs ~ batch^2 on the host, see "
            f"bench.py _pick_reference_sample; a smaller batch FAVOURS the reference's Mpixels/s): {what}")


def cpu_baseline(budget_s=20.0):
    import torch
    cores = host_threads()
    n_img, step, pixels, kind, what = _pick_reference_sample(budget_s / 3.0)
    step()
    t0 = time.perf_counter()
    n = 0
    while n < 2 or time.perf_counter() - t0 < budget_s * 0.6:
        step()
        n += 1
    dt = time.perf_counter() - t0
    return {"value": pixels * n / dt / 1e6, "unit": "Mpixels/s", "cores": cores, "kind": kind,
            "sample": f"{n} steps, {_sample_text(n_img, what)}; {dt:.1f} s, torch {torch.__version__}",
            "ms_per_step": dt / n * 1e3, "images_per_step": n_img}


def run_reference(args):
    """--impl reference: the reference's own _step on the same config on host cores; rank 0 only"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_threads()
    total = args.steps + args.warmup
    n_img, step, pixels, kind, what = _pick_reference_sample(280.0 / max(total, 1))   # the whole run ends within minutes
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = pixels * args.steps / dt / 1e6
    line = {
        "impl": "reference", "metric": "proto_head_fwd_bwd_mpixels_per_s", "value": val, "unit": "Mpixels/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict("fp32", None, None),
        "cpu_baseline": {"value": val, "unit": "Mpixels/s", "cores": cores, "kind": kind,
                         "sample": _sample_text(n_img, what), "images_per_step": n_img},
        "e2e": {"value": val, "unit": "Mpixels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def config_dict(prec, ring, transport):
    c = CFG
    return {"workload": "configs[1] head: fused prototype-head training step (fwd + CE + KLD + L1 + bwd), "
                        "batch 16 of 64x65x65 features (513x513 crops), 190 prototypes x 64-d, 19 classes; "
                        "backbone out of scope",
            "global_batch_per_gpu": c["B"], "pixels_per_step_per_gpu": c["B"] * c["h"] * c["w"],
            "precision": prec,
            "parallelism": "dp (batch-sharded; normalisers + prototype/last-layer gradients exchanged once per step)",
            "dp_transport": transport,
            "l2_policy": (f"ring of {ring} distinct resident batches ({ring * c['B'] * c['K'] * c['h'] * c['w'] * 4 / 2**20:.0f} MiB "
                          "> 2x L2) so every step reads inputs evicted from L2") if ring else "n/a"}


# ------------------------------------------------------------------------------------------------ GPU legs
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from adlm_inherently_interpretable_segmentation_b200 import _lib, head, parallel, push, synth

    if args.workload == "cfg5":
        return run_cfg5(args, torch, dist, head, synth)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    binding = bind_near_gpu(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    prec = args.precision or "tf32"
    c = CFG
    B, K, h, w, N, C = c["B"], c["K"], c["h"], c["w"], c["N"], c["C"]
    M = B * h * w
    K_steps, W_steps = args.steps, max(args.warmup, 3)
    sm_hz = torch.cuda.get_device_properties(dev).clock_rate * 1e3 if hasattr(torch.cuda.get_device_properties(dev), "clock_rate") else 1.9e9

    # ---- synthetic, seeded inputs: a ring of distinct batches larger than L2
    ring = 16
    g = torch.Generator(device=dev).manual_seed(synth.BASE_SEED + rank)
    zs = [torch.sigmoid(torch.randn(B, K, h, w, device=dev, generator=g)) for _ in range(ring)]
    labs = synth.block_labels(ring * B, h, w, C, block=8, seed=synth.BASE_SEED + rank).astype(np.int32)
    tgts = [torch.from_numpy(labs[i * B:(i + 1) * B]).to(dev) for i in range(ring)]
    _, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED)
    P = torch.from_numpy(Pn).to(dev)
    Wt = torch.from_numpy(Wn).to(dev)
    pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
    ws = torch.empty(int(_lib.load().adlm_step_workspace_bytes(B, K, h * w, N, C)), device=dev, dtype=torch.uint8)

    transport = None
    exchange = None
    if world > 1:
        if parallel.dp_supported_fused(K, N, C, prec):
            exchange = parallel.default_exchange(K, N, C, dev)
        transport = "peer" if exchange is not None else "nccl"

    def step(i):
        j = i % ring
        if world > 1:
            return parallel.dp_fused_step((P, Wt), zs[j], tgts[j], pc, weights=WEIGHTS, precision=prec, workspace=ws,
                                          exchange=exchange)
        return head.fused_step(zs[j], P, Wt, tgts[j], pc, weights=WEIGHTS, precision=prec, workspace=ws)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n, sampler=None, drain=None, gate_us=0.0):
        """n calls of fn between two events on the launching stream, bracketed by barrier + synchronize.  gate_us > 0
        puts a spin kernel of that length in front of the first event, so that the host has the first steps enqueued
        when the timed region starts on the device (every rank starts with a full queue; without it the first launches
        of a short window run one by one behind the host, and at N > 1 the ranks start skewed by their host jitter).
        At N > 1 one extra untimed step follows the gate (reported in `warmup`)."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ctx = sampler if sampler is not None else _Null()
        with ctx:
            if gate_us > 0:
                torch.cuda._sleep(int(gate_us * 1e-6 * sm_hz))
                if world > 1:
                    # one more untimed warm-up step behind the gate: its in-kernel exchange is a device-side barrier, so
                    # the ranks enter the timed region within microseconds of each other instead of a host barrier's
                    # ~0.5 ms (which a 20-step window would otherwise charge to its first step)
                    fn(n)
            timed.launches_before = _lib.launch_count()
            e0.record()
            for i in range(n):
                fn(i)
            if drain is not None:
                drain()
            e1.record()
            barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms

    for i in range(W_steps):
        step(i)
    # host cost of issuing one step (python + ctypes + six launches), measured with an empty queue behind a long gate
    torch.cuda.synchronize()
    torch.cuda._sleep(int(0.02 * sm_hz))
    t0 = time.perf_counter()
    for i in range(20):
        step(i)
    host_us = (time.perf_counter() - t0) / 20 * 1e6
    torch.cuda.synchronize()
    gate_us = min(20000.0, max(200.0, 0.5 * host_us * K_steps))
    sampler = ClockSampler(local)
    ms = timed(step, K_steps, sampler, gate_us=gate_us)
    launches = _lib.launch_count() - timed.launches_before
    value = world * M * K_steps / (ms * 1e-3) / 1e6

    # ---- roofline leg: per-launch CUDA events around the kernel groups, same loop
    pk = peaks()
    roof = None
    kern = {}
    for tag, name, nbytes, nflop in ((_lib.K_HEAD_BWD, "head_bwd", BYTES_BWD(K, C) * M, FLOPS_BWD(N, K, C) * M),
                                     (_lib.K_HEAD_FWD, "head_fwd", BYTES_FWD(K, C) * M, FLOPS_FWD(N, K, C) * M),
                                     (_lib.K_KLD, "kld", None, None), (_lib.K_STEP_PREPARE, "prepare", None, None)):
        with _lib.kernel_timing(tag, K_steps) as t:
            timed(step, K_steps, gate_us=gate_us)
        if t.ms:
            avg = float(np.mean(t.ms))
            kern[name] = {"avg_ms": avg, "launches": len(t.ms)}
            if nbytes:
                kern[name].update(algorithmic_bytes=nbytes, achieved_gbs=nbytes / (avg * 1e-3) / 1e9,
                                  algorithmic_tflops=nflop / (avg * 1e-3) / 1e12)
    tf32 = None
    if rank == 0 and not args.no_extras:
        tf32 = measure_tf32_peak(torch, dev)
    if kern:
        dom = max((k for k in kern if "achieved_gbs" in kern[k]), key=lambda k: kern[k]["avg_ms"])
        traffic, traffic_src = ncu_traffic(dom, prec)
        roof = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["achieved_gbs"], "peak": pk["hbm"],
                "unit": "GB/s", "frac": kern[dom]["achieved_gbs"] / pk["hbm"],
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": pk["source"], "kernels": kern,
                "share_of_step": kern[dom]["avg_ms"] / (ms / K_steps)}
        if tf32 is not None:
            step_tflops = (FLOPS_FWD(N, K, C) + FLOPS_BWD(N, K, C)) * M / (ms / K_steps * 1e-3) / 1e12   # per GPU
            roof["tensor"] = {"bound": "tensor", "unit": "TFLOP/s", "peak": tf32["tf32_tflops"],
                              "peak_sustained": tf32["tf32_tflops_sustained"], "peak_source": tf32["how"],
                              "achieved_step": step_tflops, "frac_step": step_tflops / tf32["tf32_tflops"],
                              "achieved_dominant": kern[dom]["algorithmic_tflops"],
                              "frac_dominant": kern[dom]["algorithmic_tflops"] / tf32["tf32_tflops"],
                              "flops_per_pixel": FLOPS_FWD(N, K, C) + FLOPS_BWD(N, K, C),
                              "note": "algorithmic flops (6NK + 6NC per pixel, SURVEY 8(d)); recomputation in the "
                                      "backward and lane padding are not counted"}

    # ---- e2e leg: host buffers in, host buffers out, through the public API.  Triple-buffered over three streams
    # (H2D | kernels | D2H) so the two PCIe directions and the kernels overlap; the caller receives step i-2's
    # results on the host while step i is issued.  Every step's copies are inside the timed region.
    n_host, NB = 4, 3
    hz = [zs[i].cpu().pin_memory() for i in range(n_host)]
    ht = [tgts[i].cpu().pin_memory() for i in range(n_host)]
    n_grad = N * K + C * N
    dz_h = [torch.empty((B, K, h, w), dtype=torch.float32).pin_memory() for _ in range(NB)]
    g_h = [torch.empty(n_grad, dtype=torch.float32).pin_memory() for _ in range(NB)]
    sc_h = [torch.empty(_lib.STEP_NSCALARS, dtype=torch.float64).pin_memory() for _ in range(NB)]
    z_d = [torch.empty((B, K, h, w), device=dev, dtype=torch.float32) for _ in range(NB)]
    t_d = [torch.empty((B, h, w), device=dev, dtype=torch.int32) for _ in range(NB)]
    dz_d = [torch.empty((B, K, h, w), device=dev, dtype=torch.float32) for _ in range(NB)]
    g_d = [torch.empty(n_grad, device=dev, dtype=torch.float32) for _ in range(NB)]
    sc_d = [torch.empty(_lib.STEP_NSCALARS, device=dev, dtype=torch.float64) for _ in range(NB)]
    h2d = hz[0].numel() * 4 + ht[0].numel() * 4
    d2h = dz_h[0].numel() * 4 + n_grad * 4 + _lib.STEP_NSCALARS * 8
    cur = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    ev_in = [torch.cuda.Event() for _ in range(NB)]
    ev_cmp = [torch.cuda.Event() for _ in range(NB)]
    ev_out = [torch.cuda.Event() for _ in range(NB)]
    for b in range(NB):
        ev_cmp[b].record(cur)
        ev_out[b].record(cur)

    def e2e_step(i):
        j, b = i % n_host, i % NB
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_cmp[b])                 # the kernels of step i-NB have consumed z_d[b]
            z_d[b].copy_(hz[j], non_blocking=True)
            t_d[b].copy_(ht[j], non_blocking=True)
            ev_in[b].record(s_in)
        cur.wait_event(ev_in[b])
        cur.wait_event(ev_out[b])                      # the D2H of step i-NB has drained dz_d[b] / g_d[b]
        if world > 1:
            r = parallel.dp_fused_step((P, Wt), z_d[b], t_d[b], pc, weights=WEIGHTS, precision=prec, workspace=ws,
                                       dz_out=dz_d[b], exchange=exchange)
            g_d[b][:N * K].copy_(r.dP.reshape(-1))
            g_d[b][N * K:].copy_(r.dW.reshape(-1))
        else:
            r = head.fused_step(z_d[b], P, Wt, t_d[b], pc, weights=WEIGHTS, precision=prec, workspace=ws,
                                dz_out=dz_d[b], grads_out=g_d[b])
        sc_d[b].copy_(r.scalars)
        ev_cmp[b].record(cur)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_cmp[b])
            dz_h[b].copy_(dz_d[b], non_blocking=True)
            g_h[b].copy_(g_d[b], non_blocking=True)
            sc_h[b].copy_(sc_d[b], non_blocking=True)
            ev_out[b].record(s_out)
        ev_out[(i - 2) % NB].synchronize()             # the caller now holds step i-2's results on the host

    def e2e_drain():
        for b in range(NB):
            cur.wait_event(ev_out[b])                  # the closing event sees the last D2H

    for i in range(6):
        e2e_step(i)
    ms_e2e = timed(e2e_step, K_steps, drain=e2e_drain)
    if not all(bool(torch.isfinite(x).all()) for x in sc_h + dz_h):
        raise RuntimeError("e2e leg: non-finite results on the host")
    e2e = {"value": world * M * K_steps / (ms_e2e * 1e-3) / 1e6, "unit": "Mpixels/s",
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / K_steps,
           "host_binding": binding,
           "pipeline": "3 buffers, 3 streams (H2D | kernels | D2H); results of step i-2 are on the host when step i has been issued"}
    del hz, ht, dz_h, g_h, sc_h, z_d, t_d, dz_d, g_d, sc_d

    # ---- what the same GPU does with the reference formulation in stock torch, and configs[3]
    stock = cfg4 = cfg5 = None
    if rank == 0 and not args.no_extras:
        stock = bench_stock_torch(torch, zs[0], P, Wt, tgts[0], pc)
        cfg4 = bench_cfg4(torch, head, synth, dev, pk, prec)
    del zs
    torch.cuda.empty_cache()
    if rank == 0 and not args.no_extras and prec == "tf32":
        cfg5 = bench_cfg5(torch, head, synth, dev, tf32)
        torch.cuda.empty_cache()

    # ---- secondary metric: push imgs/s over the Cityscapes-train-sized synthetic set (configs[2])
    push_res = None
    if not args.no_push:
        push_res = bench_push(torch, dist, push, synth, dev, rank, world, "tf32x3", pk,
                              args.push_images or PUSH_CFG["n_total"])

    cpu = None
    if rank == 0 and not args.no_cpu:
        cpu = cpu_baseline()

    if rank == 0:
        line = {
            "metric": "proto_head_fwd_bwd_mpixels_per_s", "value": value, "unit": "Mpixels/s", "n_gpus": world,
            "steps": K_steps, "warmup": W_steps + (1 if world > 1 else 0), "ms_per_step": ms / K_steps,
            "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"fp32": "f32", "tf32": "tf32", "tf32x3": "tf32x3"}[prec],
            "data": "synthetic", "config": config_dict(prec, ring, transport), "clocks": sampler.summary(), "e2e": e2e,
            "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "push": push_res,
            "stock_torch_gpu": stock, "cfg4": cfg4, "cfg5": cfg5, "tf32_peak": tf32,
            "host_issue_us_per_step": host_us, "start_gate_us": gate_us,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def ncu_traffic(kernel_group, prec):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel group, from the committed summary of the
    round's `ncu --set full` capture (profiles/r2_dram_traffic.json, written by tools/ncu_traffic.py from the capture
    itself); null when there is no capture for this build / precision"""
    path = os.path.join(ROOT, "profiles", "r2_dram_traffic.json")
    try:
        rec = json.load(open(path))
        if rec.get("precision") == prec and kernel_group in rec.get("groups", {}):
            return float(rec["groups"][kernel_group]), f"profiles/r2_dram_traffic.json ({rec.get('capture')})"
    except Exception:  # noqa: BLE001
        pass
    return None, None


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def _cuda_time(torch, fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def bench_stock_torch(torch, z0, P, W, target, pc):
    """The reference formulation of the head step on this GPU with stock torch (cuDNN / cuBLAS): the op sequence of
    model.py:203-221 (two 1x1 convs, one against `ones`), :231-237, :161-162, cross entropy on the non-void patches
    (module.py:156-165), masked L1 (:213-218) and autograd — first without the KLD term, then with the Python loops of
    module.py:167-208 as the reference runs them (masked_select / log_softmax / kl_div per (image, label, pair))."""
    import torch.nn.functional as F
    B, K, h, w = z0.shape
    N, C = P.shape[0], W.shape[0]
    M = B * h * w
    ones = torch.ones(N, K, 1, 1, device=z0.device)
    ident = torch.zeros(N, C, device=z0.device)
    ident[torch.arange(N, device=z0.device), pc.long()] = 1
    tgt = target.long()

    def run(with_kld):
        z = z0.detach().clone().requires_grad_(True)
        P4 = P.detach().reshape(N, K, 1, 1).clone().requires_grad_(True)
        Wl = W.detach().clone().requires_grad_(True)
        x2 = F.conv2d(z ** 2, ones)
        p2 = torch.sum(P4 ** 2, dim=(1, 2, 3)).view(-1, 1, 1)
        xp = F.conv2d(z, P4)
        dist = F.relu(x2 + (-2 * xp + p2))
        flat = dist.permute(0, 2, 3, 1).contiguous().reshape(-1, N)
        act = torch.log((flat + 1) / (flat + 1e-4))
        logits = F.linear(act, Wl)
        t = tgt.reshape(-1)
        keep = (t != 0).nonzero().squeeze()
        ce = F.cross_entropy(logits[keep], t[keep] - 1)
        l1 = (Wl * (1 - ident.t())).norm(p=1)
        loss = WEIGHTS[0] * ce + WEIGHTS[2] * l1
        if with_kld:
            kld = []
            dv = dist.reshape(B, N, h * w)
            for b in range(B):
                tb = tgt[b].reshape(-1)
                for cls in torch.unique(tb).cpu().numpy():
                    if cls < 0 or cls >= C:
                        continue
                    protos = torch.nonzero(ident[:, cls]).flatten().cpu().numpy()
                    if len(protos) < 2:
                        continue
                    mask = tb == int(cls)
                    if int(mask.sum()) < 2:
                        continue
                    logp = [F.log_softmax(torch.masked_select(dv[b, int(p)], mask), dim=0) for p in protos]
                    for i in range(len(protos)):
                        for j in range(i + 1, len(protos)):
                            k1 = F.kl_div(logp[i], logp[j], log_target=True, reduction='sum')
                            k2 = F.kl_div(logp[j], logp[i], log_target=True, reduction='sum')
                            kld.append(torch.exp(-(k1 + k2) / 2.0))
            if kld:
                loss = loss + WEIGHTS[1] * torch.stack(kld).mean()
        loss.backward()
        return loss

    out = {"what": "reference formulation in stock torch on this GPU (cuDNN 1x1 convs, elementwise, nn.Linear, "
                   "cross_entropy, autograd), batch 16 of 64x65x65, 190x64, C=19", "unit": "Mpixels/s"}
    old_c, old_m = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    try:
        for tf in (True, False):
            torch.backends.cudnn.allow_tf32 = tf
            torch.backends.cuda.matmul.allow_tf32 = tf
            ms = _cuda_time(torch, lambda: run(False), 10)
            out[f"no_kld_allow_tf32_{str(tf).lower()}"] = {"ms_per_step": ms, "value": M / (ms * 1e-3) / 1e6}
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cuda.matmul.allow_tf32 = False          # torch's defaults
        t0 = time.perf_counter()
        run(True)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        run(True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t1
        out["with_kld_loops_torch_defaults"] = {"ms_per_step": dt * 1e3, "value": M / dt / 1e6, "first_call_ms": (t1 - t0) * 1e3,
                                                "note": "wall clock of one step (the loops synchronise with the host)"}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old_c, old_m
    return out


def bench_cfg4(torch, head, synth, dev, pk, prec):
    """BASELINE configs[3]: mds_new pancreas head (30 prototypes = 3 classes x 10, mds_new.gin:6-7), batch 64 of 64x65x65
    slices, inference: logits + per-pixel argmax (eval_valid.py:159-176), no distance map.  HBM-bound (AI ~ 15)."""
    from adlm_inherently_interpretable_segmentation_b200 import _lib
    c = CFG4
    B, K, h, w, N, C = c["B"], c["K"], c["h"], c["w"], c["N"], c["C"]
    M = B * h * w
    g = torch.Generator(device=dev).manual_seed(synth.BASE_SEED + 4)
    ring = 5                                        # 5 x 69 MB of features > 2 x L2
    zs = [torch.sigmoid(torch.randn(B, K, h, w, device=dev, generator=g)) for _ in range(ring)]
    _, Pn, Wn = synth.head_inputs(1, K, 2, 2, N, C, w_noise=0.1, seed=synth.BASE_SEED + 4)
    P, W = torch.from_numpy(Pn).to(dev), torch.from_numpy(Wn).to(dev)
    i = [0]

    def fn():
        head.head_forward(zs[i[0] % ring], P, W, precision=prec, want_dist=False, want_argmax=True)
        i[0] += 1
    _cuda_time(torch, fn, 10)
    with _lib.kernel_timing(_lib.K_HEAD_FWD, 64) as t:
        ms_call = _cuda_time(torch, fn, 50, warm=0)
    ms_k = float(np.mean(t.ms)) if t.ms else ms_call
    nbytes = (4 * K + 4 * C + 4) * M
    return {"workload": "configs[3]: mds_new head, 30 prototypes x 64-d, 3 classes, batch 64 of 64x65x65 slices, "
                        "inference (logits + per-pixel argmax, no distance map)", "precision": prec,
            "value": M / (ms_call * 1e-3) / 1e6, "unit": "Mpixels/s", "ms_per_call": ms_call, "kernel_ms": ms_k,
            "roofline": {"bound": "hbm", "achieved": nbytes / (ms_k * 1e-3) / 1e9, "peak": pk["hbm"], "unit": "GB/s",
                         "frac": nbytes / (ms_k * 1e-3) / 1e9 / pk["hbm"], "algorithmic_bytes": nbytes,
                         "bytes_per_pixel": 4 * K + 4 * C + 4},
            "l2_policy": f"ring of {ring} batches ({ring * M * K * 4 / 2**20:.0f} MiB)"}


def _cfg5_inputs(torch, synth, dev, seed):
    c = CFG5
    B, K, h, w, N, C = c["B"], c["K"], c["h"], c["w"], c["N"], c["C"]
    g = torch.Generator(device=dev).manual_seed(seed)
    z = torch.empty(B, K, h, w, device=dev)
    for k0 in range(0, K, 32):                     # 2.1 GB of features, generated in slabs
        z[:, k0:k0 + 32].normal_(generator=g).sigmoid_()
    P = torch.rand(N, K, device=dev, generator=g)
    W = torch.randn(C, N, device=dev, generator=g) * 0.1
    return z, P, W


def _cfg5_check(torch, out, z, P, W, h, w):
    """spot check of a strip of pixels against float64 torch (model.py:203-283); raises when the kernel is off"""
    K, N, C = z.shape[1], P.shape[0], W.shape[0]
    zs = z[0].reshape(K, h * w)[:, :4096].t().double()
    d = ((zs[:, None, :] - P.double()[None]) ** 2).sum(-1)
    ref = torch.log((d + 1) / (d + 1e-4)) @ W.double().t()
    got = out["logits"].reshape(h * w, C)[:4096].double()
    rel = float((got - ref).abs().max() / ref.abs().max())
    agree = float((got.argmax(1) == ref.argmax(1)).float().mean())
    if not (rel < 2e-3 and agree > 0.99):
        raise RuntimeError(f"cfg5 leg: logits off by {rel:.2e} (relative), argmax agreement {agree:.4f}")
    return rel, agree


def bench_cfg5(torch, head, synth, dev, tf32, reps=5):
    """BASELINE configs[4]: 1000 prototypes x 256-d, 100 classes on a 1024x2048 feature map — forward + per-pixel argmax
    through adlm_head_fwd_big (CTA pairs, tcgen05 cta_group::2).  Tensor-bound (AI ~ 680 flop/B): the roofline is the
    TF32 matmul peak measured in this run; algorithmic flop = (2NK + 2NC) per pixel."""
    from adlm_inherently_interpretable_segmentation_b200 import _lib
    c = CFG5
    B, K, h, w, N, C = c["B"], c["K"], c["h"], c["w"], c["N"], c["C"]
    M = B * h * w
    z, P, W = _cfg5_inputs(torch, synth, dev, synth.BASE_SEED + 5)
    fn = lambda: head.head_forward(z, P, W, precision="tf32", want_dist=False, want_argmax=True)
    out = fn()
    rel, agree = _cfg5_check(torch, out, z, P, W, h, w)
    del out
    _cuda_time(torch, fn, 2, warm=1)
    with _lib.kernel_timing(_lib.K_HEAD_FWD, 64) as t:
        ms_call = _cuda_time(torch, fn, reps, warm=0)
    ms_k = float(np.mean(t.ms)) if t.ms else ms_call
    tfl = FLOPS_FWD(N, K, C) * M / (ms_k * 1e-3) / 1e12
    peak = tf32["tf32_tflops"] if tf32 else None
    return {"workload": "configs[4]: 1000 prototypes x 256-d, 100 classes, one 1024x2048 feature map, forward + per-pixel "
                        "argmax (logits written, no distance map)", "precision": "tf32",
            "value": M / (ms_call * 1e-3) / 1e6, "unit": "Mpixels/s", "ms_per_call": ms_call, "kernel_ms": ms_k,
            "roofline": {"bound": "tensor", "achieved": tfl, "peak": peak, "unit": "TFLOP/s",
                         "frac": tfl / peak if peak else None, "flop_per_pixel": FLOPS_FWD(N, K, C),
                         "peak_source": "TF32 matmul measured in this run (tf32_peak)"},
            "hbm_gbs_algorithmic": (4 * K + 4 * C + 4) * M / (ms_k * 1e-3) / 1e9,
            "check": {"logits_rel_err_vs_fp64_strip": rel, "argmax_agreement_strip": agree},
            "l2_policy": "one 2.1 GB map: every call streams its inputs from HBM"}


def run_cfg5(args, torch, dist, head, synth):
    """`--workload cfg5`: the large-head forward as the line's own metric.  One map per GPU per step, no collective
    (batch-sharded inference); barrier + synchronize around the timed steps, max over ranks."""
    from adlm_inherently_interpretable_segmentation_b200 import _lib
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    c = CFG5
    B, K, h, w, N, C = c["B"], c["K"], c["h"], c["w"], c["N"], c["C"]
    M = B * h * w
    K_steps, W_steps = min(args.steps, 20), max(min(args.warmup, 5), 3)
    z, P, W = _cfg5_inputs(torch, synth, dev, synth.BASE_SEED + 5 + rank)
    fn = lambda: head.head_forward(z, P, W, precision="tf32", want_dist=False, want_argmax=True)
    rel, agree = _cfg5_check(torch, fn(), z, P, W, h, w)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(W_steps):
        fn()
    sampler = ClockSampler(local)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with sampler, _lib.kernel_timing(_lib.K_HEAD_FWD, K_steps) as t:
        e0.record()
        for _ in range(K_steps):
            fn()
        e1.record()
        barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    ms_k = float(np.mean(t.ms)) if t.ms else ms / K_steps
    # e2e: argmax map back on the host every step; features are produced by the backbone on the device (out of scope),
    # so the H2D leg carries nothing here
    amax_h = torch.empty(B, h, w, dtype=torch.int32).pin_memory()
    barrier()
    e0.record()
    for _ in range(K_steps):
        amax_h.copy_(fn()["argmax"], non_blocking=True)
    e1.record()
    barrier()
    ms_e = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_e, op=dist.ReduceOp.MAX)
    ms_e = float(ms_e)
    tf32 = measure_tf32_peak(torch, dev) if rank == 0 else None
    if rank == 0:
        tfl = FLOPS_FWD(N, K, C) * M / (ms_k * 1e-3) / 1e12
        emit({"metric": "proto_head_fwd_mpixels_per_s", "value": world * M * K_steps / (ms * 1e-3) / 1e6,
              "unit": "Mpixels/s", "n_gpus": world, "steps": K_steps, "warmup": W_steps, "ms_per_step": ms / K_steps,
              "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "tf32", "data": "synthetic",
              "config": {"workload": "configs[4] head: 1000 prototypes x 256-d, 100 classes, one 1024x2048 feature map per GPU "
                                     "per step, forward + per-pixel argmax; backbone out of scope",
                         "parallelism": "batch-sharded inference, no collective",
                         "l2_policy": "one 2.1 GB map per step: inputs stream from HBM"},
              "clocks": sampler.summary(),
              "e2e": {"value": world * M * K_steps / (ms_e * 1e-3) / 1e6, "unit": "Mpixels/s", "h2d_bytes_per_step": 0,
                      "d2h_bytes_per_step": int(amax_h.numel() * 4), "ms_per_step": ms_e / K_steps,
                      "note": "argmax map to pinned host memory every step; the features come from the backbone on the device"},
              "gpu_launches": 2 * K_steps,
              "roofline": {"bound": "tensor", "kernel": "head_fwd (k_tc_fwd_big)", "achieved": tfl,
                           "peak": tf32["tf32_tflops"], "unit": "TFLOP/s", "frac": tfl / tf32["tf32_tflops"],
                           "kernel_ms": ms_k, "flop_per_pixel": FLOPS_FWD(N, K, C), "traffic": None,
                           "peak_source": "TF32 matmul measured in this run"},
              "cpu_baseline": None, "tf32_peak": tf32,
              "check": {"logits_rel_err_vs_fp64_strip": rel, "argmax_agreement_strip": agree}})
    if world > 1:
        dist.destroy_process_group()


def bench_push(torch, dist, push, synth, dev, rank, world, prec, pk, n_total):
    """push search over the whole synthetic set, sharded like push.push_prototypes shards it (contiguous image ranges,
    earliest image wins): labels -> class mask -> scan -> packed-min all-reduce -> gather (+ sum all-reduce)"""
    from adlm_inherently_interpretable_segmentation_b200 import _lib
    c = PUSH_CFG
    K, h, w, H, W, N, C = c["K"], c["h"], c["w"], c["H"], c["W"], c["N"], c["C"]
    per = (n_total + world - 1) // world
    lo, hi = min(n_total, rank * per), min(n_total, (rank + 1) * per)
    n = hi - lo
    g = torch.Generator(device=dev).manual_seed(synth.BASE_SEED + 100 + rank)
    z = torch.empty(n, K, h, w, device=dev)
    for i in range(0, n, 64):                       # in place, chunked: 25.3 GB at N = 1
        blk = z[i:i + 64]
        blk.normal_(generator=g)
        blk.sigmoid_()
    lab1 = synth.block_labels(12, H, W, C, block=32, seed=synth.BASE_SEED + rank)
    base = torch.from_numpy(lab1).to(dev)
    labels = base.repeat((n + 11) // 12, 1, 1)[:n].contiguous()      # 6.2 GB of uint8 at N = 1
    P = torch.rand(N, K, device=dev, generator=g)
    pc = torch.arange(N, dtype=torch.int32, device=dev) // (N // C)
    chunk = 512                                     # images per labels->mask / scan / gather call
    out = torch.zeros((N, K), device=dev)

    def one_pass(timers=None):
        def mark(k):
            if timers is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                timers.append((k, e))
        mark("start")
        s = push.PushSearch(P, pc, precision=prec)
        masks = [s.class_mask(labels[i:i + chunk], h, w) for i in range(0, n, chunk)]
        mark("mask")
        for ci, i in enumerate(range(0, n, chunk)):
            s.scan(z[i:i + chunk], None, lo + i, mask=masks[ci])
        mark("scan")
        s.reduce()
        mark("reduce")
        out.zero_()
        for i in range(0, n, chunk):
            s.gather(z[i:i + chunk], lo + i, out)
        if world > 1:
            dist.all_reduce(out)
        mark("gather")
        return out

    one_pass()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    reps = 3
    e0.record()
    for _ in range(reps):
        one_pass()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    timers = []
    one_pass(timers)
    torch.cuda.synchronize()
    breakdown = {timers[i][0] + "_ms": timers[i - 1][1].elapsed_time(timers[i][1]) for i in range(1, len(timers))}
    with _lib.kernel_timing(_lib.K_PUSH_SCAN, 64) as t:
        one_pass()
        torch.cuda.synchronize()
    scan_ms = float(np.sum(t.ms)) if t.ms else None
    bytes_scan = n * h * w * (4 * K + 4)
    flops_scan = 2.0 * N * K * n * h * w          # algorithmic: one pass of the z.p contraction (3xTF32 issues 3x)
    return {"metric": "push_imgs_per_s", "value": n_total / (ms * 1e-3), "unit": "imgs/s", "scaling": "strong",
            "images_total": n_total, "images_this_gpu": n, "ms_per_pass": ms, "precision": prec,
            "breakdown_rank0": breakdown,
            "workload": f"configs[2]: {n_total}-image Cityscapes-train-sized synthetic set sharded over {world} GPU(s) "
                        "(contiguous ranges), 1024x2048 uint8 labels, 64x129x257 features, 190 prototypes; labels->mask, "
                        "scan, packed-min reduce, gather (features resident; backbone out of scope)",
            "scan_kernel": None if scan_ms is None else {
                "ms_per_pass": scan_ms, "achieved_gbs": bytes_scan / (scan_ms * 1e-3) / 1e9,
                "frac_of_hbm_peak": bytes_scan / (scan_ms * 1e-3) / 1e9 / pk["hbm"],
                "algorithmic_tflops": flops_scan / (scan_ms * 1e-3) / 1e12,
                "issued_tf32_tflops": (3.0 if prec == "tf32x3" else 1.0) * (256.0 / N) * flops_scan / (scan_ms * 1e-3) / 1e12,
                "note": "tensor-bound: 3xTF32 and 190 -> 256 prototype lanes; ncu sm__pipe_tc_cycles_active in profiles/"}}


if __name__ == "__main__":
    main()
