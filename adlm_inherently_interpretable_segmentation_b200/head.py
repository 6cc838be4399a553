"""Host side of the prototype head: thin torch wrappers over the C ABI (include/adlm_head.h).

  head_forward      one launch: distances (optional), similarities (optional), logits, argmax
  HeadFunction      autograd drop-in: backward recomputes distances on chip (adlm_head_bwd)
  fused_step        forward + CE + KLD + L1 + backward of segmentation/module.py:137-222,236 in
                    one C-ABI call; the M x N distance map never touches HBM

torch is used for device memory, streams and autograd plumbing only; all arithmetic of the path runs
in libadlm_head.so.  Non-CUDA / non-fp32 inputs raise — there is no fallback.
"""
import ctypes
from ctypes import c_float, c_int, c_longlong, c_size_t

import torch

from . import _lib

DEFAULT_PRECISION = "fp32"


def _prec(p):
    if isinstance(p, int):
        return p
    try:
        return _lib.PREC[p]
    except KeyError:
        raise ValueError(f"unknown precision {p!r}; expected one of {sorted(_lib.PREC)}")


def _act(a):
    if a == "log":
        return _lib.ACT_LOG
    if a == "linear":
        return _lib.ACT_LINEAR
    raise NotImplementedError(
        "the fused prototype head implements prototype_activation_function 'log' and 'linear' "
        f"(model.py:231-237); got {a!r}")


def _check_cuda_f32(name, t):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a tensor")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor: the prototype head has no CPU path")
    if t.dtype != torch.float32:
        raise ValueError(f"{name} must be float32, got {t.dtype}")


def _proto_2d(prototype_vectors):
    """(N,K,1,1) parameter -> (N,K) view; other prototype kernel sizes are not implemented."""
    if prototype_vectors.dim() == 4:
        if prototype_vectors.shape[2] != 1 or prototype_vectors.shape[3] != 1:
            raise NotImplementedError("prototype kernel must be 1x1 (every segmentation config uses 1x1)")
        return prototype_vectors.reshape(prototype_vectors.shape[0], prototype_vectors.shape[1])
    if prototype_vectors.dim() == 2:
        return prototype_vectors
    raise ValueError("prototype_vectors must be (N,K,1,1) or (N,K)")


def uses_big_head_kernel(K, N, C):
    """shapes of the 'tf32' forward served by adlm_head_fwd_big (N-tiled, prototypes streamed by TMA bulk copies)
    instead of the warp-specialised kernel whose operands must fit shared memory (K <= 64, N <= 224, C <= 32)"""
    return (K > 64 or N > 224 or C > 32) and K <= 256 and N <= 1024 and C <= 112


def head_forward(z, prototypes, weight, eps=1e-4, activation="log", precision=None,
                 want_dist=True, want_sim=False, want_argmax=False, pre_sigmoid=False):
    """z (B,K,h,w) -> dict(logits (B,h,w,C), dist (B,N,h,w)|None, sim (B*h*w,N)|None, argmax (B,h,w)|None).
    Mirrors model.py:259-283 (patch-classification branch).
    pre_sigmoid: `z` holds the features BEFORE the add-on Sigmoid (model.py:124-128,174); the kernel applies it while
    staging the tiles and the result carries z = sigmoid(input) under key "z" (adlm_head_fwd_pre; tf32 only)."""
    lib = _lib.load()
    _check_cuda_f32("conv_features", z)
    _check_cuda_f32("prototype_vectors", prototypes)
    _check_cuda_f32("last_layer.weight", weight)
    if z.dim() != 4:
        raise ValueError("conv_features must be (B,K,h,w)")
    z = z.contiguous()
    P = _proto_2d(prototypes).contiguous()
    W = weight.contiguous()
    B, K, h, w = z.shape
    N, C = P.shape[0], W.shape[0]
    if P.shape[1] != K or W.shape[1] != N:
        raise ValueError(f"shape mismatch: z K={K}, prototypes {tuple(P.shape)}, weight {tuple(W.shape)}")
    HW = h * w
    dev = z.device
    with torch.cuda.device(dev):
        logits = torch.empty((B, h, w, C), device=dev, dtype=torch.float32)
        dist = torch.empty((B, N, h, w), device=dev, dtype=torch.float32) if want_dist else None
        sim = torch.empty((B * HW, N), device=dev, dtype=torch.float32) if want_sim else None
        amax = torch.empty((B, h, w), device=dev, dtype=torch.int32) if want_argmax else None
        prec_name = precision or DEFAULT_PRECISION
        if prec_name == "tf32" and not pre_sigmoid and uses_big_head_kernel(K, N, C):
            # beyond the operand budget of the warp-specialised kernel: the N-tiled kernel with streamed prototypes
            if want_sim:
                raise NotImplementedError("the large-head tensor-core forward does not return similarities; "
                                          "use precision 'fp32'")
            need = int(lib.adlm_head_fwd_big_workspace_bytes(K, N, C))
            ws = torch.empty(need, device=dev, dtype=torch.uint8)
            rc = lib.adlm_head_fwd_big(_lib.ptr(z), B, K, HW, _lib.ptr(P), _lib.ptr(W), N, C, c_float(eps),
                                       _act(activation), _lib.ptr(logits), _lib.ptr(dist), _lib.ptr(amax),
                                       _lib.ptr(ws), c_size_t(need), _lib.stream_ptr())
            _lib.check(rc, "adlm_head_fwd_big")
            return dict(logits=logits, dist=dist, sim=sim, argmax=amax)
        if pre_sigmoid:
            z_out = torch.empty_like(z)
            rc = lib.adlm_head_fwd_pre(_lib.ptr(z), B, K, HW, _lib.ptr(P), _lib.ptr(W), N, C,
                                       c_float(eps), _act(activation), _prec(precision or DEFAULT_PRECISION),
                                       _lib.ptr(z_out), _lib.ptr(logits), _lib.ptr(dist), _lib.ptr(sim),
                                       _lib.ptr(amax), _lib.stream_ptr())
            _lib.check(rc, "adlm_head_fwd_pre")
            return dict(logits=logits, dist=dist, sim=sim, argmax=amax, z=z_out)
        rc = lib.adlm_head_fwd(_lib.ptr(z), B, K, HW, _lib.ptr(P), _lib.ptr(W), N, C,
                               c_float(eps), _act(activation), _prec(precision or DEFAULT_PRECISION),
                               _lib.ptr(logits), _lib.ptr(dist), _lib.ptr(sim), _lib.ptr(amax),
                               _lib.stream_ptr())
    _lib.check(rc, "adlm_head_fwd")
    return dict(logits=logits, dist=dist, sim=sim, argmax=amax)


def head_backward(z, prototypes, weight, dlogits=None, ddist=None, dsim=None, eps=1e-4, activation="log",
                  precision=None, need_dz=True, need_dP=True, need_dW=True):
    """-> (dz (B,K,h,w)|None, dP like prototypes|None, dW (C,N)|None)"""
    lib = _lib.load()
    z = z.contiguous()
    P = _proto_2d(prototypes).contiguous()
    W = weight.contiguous()
    B, K, h, w = z.shape
    N, C = P.shape[0], W.shape[0]
    dev = z.device
    for name, g in (("dlogits", dlogits), ("ddist", ddist), ("dsim", dsim)):
        if g is not None:
            _check_cuda_f32(name, g)
    dlogits = dlogits.contiguous() if dlogits is not None else None
    ddist = ddist.contiguous() if ddist is not None else None
    dsim = dsim.contiguous() if dsim is not None else None
    with torch.cuda.device(dev):
        dz = torch.empty_like(z) if need_dz else None
        dP = torch.empty((N, K), device=dev, dtype=torch.float32) if need_dP else None
        dW = torch.empty((C, N), device=dev, dtype=torch.float32) if need_dW else None
        prec = _prec(precision or DEFAULT_PRECISION)
        need = int(lib.adlm_head_bwd_workspace_bytes(K, N, prec))
        ws = torch.empty(need, device=dev, dtype=torch.uint8) if need else None
        rc = lib.adlm_head_bwd(_lib.ptr(z), B, K, h * w, _lib.ptr(P), _lib.ptr(W), N, C,
                               c_float(eps), _act(activation), prec,
                               _lib.ptr(dlogits), _lib.ptr(ddist), _lib.ptr(dsim),
                               _lib.ptr(dz), _lib.ptr(dP), _lib.ptr(dW), _lib.ptr(ws), c_size_t(need),
                               _lib.stream_ptr())
    _lib.check(rc, "adlm_head_bwd")
    if dP is not None:
        dP = dP.reshape(prototypes.shape)
    return dz, dP, dW


class HeadFunction(torch.autograd.Function):
    """(z, prototype_vectors, last_layer.weight) -> (logits, distances, similarities).

    Saves only the inputs: the backward recomputes the distances on chip, so neither the M x N
    distance map nor any of the reference's five saved M x N activations is kept for autograd."""

    @staticmethod
    def forward(ctx, z, prototypes, weight, eps, activation, precision, want_dist, want_sim):
        out = head_forward(z, prototypes, weight, eps, activation, precision, want_dist, want_sim)
        ctx.save_for_backward(z, prototypes, weight)
        ctx.cfg = (eps, activation, precision)
        ctx.set_materialize_grads(False)
        empty = z.new_empty(0)
        dist = out["dist"] if want_dist else empty
        sim = out["sim"] if want_sim else empty
        non_diff = [t for t, want in ((dist, want_dist), (sim, want_sim)) if not want]
        if non_diff:
            ctx.mark_non_differentiable(*non_diff)
        return out["logits"], dist, sim

    @staticmethod
    def backward(ctx, dlogits, ddist, dsim):
        z, prototypes, weight = ctx.saved_tensors
        eps, activation, precision = ctx.cfg
        need = ctx.needs_input_grad
        if dlogits is None and ddist is None and dsim is None:
            return (None,) * 8
        if dlogits is not None:
            dlogits = dlogits.reshape(z.shape[0], z.shape[2] * z.shape[3], -1)
        dz, dP, dW = head_backward(z, prototypes, weight, dlogits, ddist, dsim, eps, activation, precision,
                                   need_dz=need[0], need_dP=need[1], need_dW=need[2])
        return dz, dP, dW, None, None, None, None, None


def proto_class_from_identity(prototype_class_identity):
    """(N,C) one-hot float -> (N,) int32 class ids (model.py:66-71).  Rows must be one-hot."""
    ident = prototype_class_identity
    if ident.dim() != 2:
        raise ValueError("prototype_class_identity must be (N,C)")
    if not bool(((ident == 0) | (ident == 1)).all()) or not bool((ident.sum(dim=1) == 1).all()):
        raise NotImplementedError("prototype_class_identity rows must be one-hot")
    return ident.argmax(dim=1).to(torch.int32)


class _StepResult(dict):
    __getattr__ = dict.__getitem__


_PER_CLASS_OK = {}


def _check_protos_per_class(pc):
    """The KLD side buffers hold at most ADLM_MAX_PROTOS_PER_CLASS (16) own-class distances per pixel; the reference
    has no such limit (module.py:170-190), so a head with a larger class is rejected loudly rather than computed wrong.
    One bincount + sync per distinct proto_class tensor (cached on storage pointer, version, length, device)."""
    key = (pc.data_ptr(), pc._version, pc.numel(), str(pc.device))
    if key in _PER_CLASS_OK:
        return
    if pc.numel():
        if int(pc.min()) < 0:
            raise ValueError("proto_class entries must be >= 0")
        worst = int(torch.bincount(pc.to(torch.int64)).max())
        if worst > _lib.MAX_PROTOS_PER_CLASS:
            raise NotImplementedError(
                f"fused_step supports at most {_lib.MAX_PROTOS_PER_CLASS} prototypes per class "
                f"(ADLM_MAX_PROTOS_PER_CLASS); this head has a class with {worst}")
    if len(_PER_CLASS_OK) > 64:
        _PER_CLASS_OK.clear()
    _PER_CLASS_OK[key] = True


def fused_step(z, prototypes, weight, target, proto_class, ignore_void=True,
               weights=(1.0, 0.25, 1e-4), grad_scale=1.0, eps=1e-4, precision=None,
               need_grads=True, want_logits=False, want_argmax=False, workspace=None, global_counts=None,
               grads_out=None, dz_out=None, dp_comm=None, pre_sigmoid=False):
    """One C-ABI call for module.py:137-222 + :236 on given features.

    z (B,K,h,w); target (B,h,w) integer labels already resized to the feature map (dataset.py:23-31);
    proto_class (N,) int32.  Returns a dict with device tensors:
      scalars (8,) float64 [loss, cross_entropy, kld_loss, l1, n_correct, n_patches, n_kld_terms, 0],
      dz, dP (shaped like prototypes), dW (None when need_grads is False), logits, argmax (optional).
    Gradients are those of (loss * grad_scale).  global_counts: (2,) int32 CUDA tensor overriding the
    CE / KLD normalisers (data-parallel shards, see parallel.py), or None.  grads_out: optional flat fp32
    CUDA buffer with at least N*K + C*N elements; dP and dW are then views into it ([dP | dW], one collective).
    dz_out: optional preallocated contiguous fp32 CUDA tensor shaped like z that receives dz (pipelined callers).
    dp_comm: a parallel.DpExchange — the step then runs adlm_step_fwd_bwd_dp: normalisers, dP, dW and the loss
    sums are exchanged with the peer GPUs inside the step's own kernels (no collective call afterwards).
    need_grads: True (dz, dP, dW), False (loss values only) or "last_layer": only dW — the reference's last-layer phase
    after push (train_and_test.last_only, module.py:106-108,384-390: features, add-on layers and prototypes are frozen).
    In the fp32-class modes that phase runs in (`fp32`, `tf32x3`; distances of pushed prototypes are ~0 there) the
    backward then skips the dz and dP contractions, two thirds of its work; dz and dP are None in the result.
    pre_sigmoid: `z` holds the features BEFORE the add-on Sigmoid; the step applies it in the forward kernel,
    returns z = sigmoid(input) under "z" and dz is the gradient wrt the pre-sigmoid input
    (adlm_step_fwd_bwd_pre; tf32 only, not combined with dp_comm)."""
    lib = _lib.load()
    _check_cuda_f32("conv_features", z)
    _check_cuda_f32("prototype_vectors", prototypes)
    _check_cuda_f32("last_layer.weight", weight)
    z = z.contiguous()
    P = _proto_2d(prototypes).contiguous()
    W = weight.contiguous()
    B, K, h, w = z.shape
    N, C = P.shape[0], W.shape[0]
    HW = h * w
    dev = z.device
    if tuple(target.shape) != (B, h, w) and tuple(target.shape) != (B, HW):
        raise ValueError(f"target must be (B,h,w) = {(B, h, w)}, got {tuple(target.shape)}")
    tgt = target.to(device=dev, dtype=torch.int32).contiguous()
    pc = proto_class.to(device=dev, dtype=torch.int32).contiguous()
    if pc.numel() != N:
        raise ValueError("proto_class must have N entries")
    _check_protos_per_class(pc)
    w_ce, w_kld, w_l1 = (float(x) for x in weights)
    with torch.cuda.device(dev):
        need = int(lib.adlm_step_workspace_bytes(B, K, HW, N, C))
        if workspace is None or workspace.numel() < need:
            workspace = torch.empty(need, device=dev, dtype=torch.uint8)
        scalars = torch.empty(_lib.STEP_NSCALARS, device=dev, dtype=torch.float64)
        last_only = isinstance(need_grads, str)
        if last_only and need_grads != "last_layer":
            raise ValueError('need_grads must be True, False or "last_layer"')
        # the tensor-core backward (tf32) always forms all three gradients; the CUDA-core one skips what is not asked for
        skip_zp = last_only and _prec(precision or DEFAULT_PRECISION) != _lib.PREC_TF32 and dp_comm is None \
            and not pre_sigmoid
        if need_grads and dz_out is not None:
            if dz_out.shape != z.shape or dz_out.dtype != torch.float32 or not dz_out.is_cuda or not dz_out.is_contiguous():
                raise ValueError("dz_out must be a contiguous CUDA float32 tensor shaped like z")
            dz = dz_out
        else:
            dz = torch.empty_like(z) if (need_grads and not skip_zp) else None
        if need_grads and grads_out is not None:
            if grads_out.dtype != torch.float32 or not grads_out.is_cuda or grads_out.numel() < N * K + C * N:
                raise ValueError("grads_out must be a CUDA float32 buffer with >= N*K + C*N elements")
            dP = grads_out[:N * K].view(N, K)
            dW = grads_out[N * K:N * K + C * N].view(C, N)
        else:
            dP = torch.empty((N, K), device=dev, dtype=torch.float32) if (need_grads and not skip_zp) else None
            dW = torch.empty((C, N), device=dev, dtype=torch.float32) if need_grads else None
        logits = torch.empty((B, h, w, C), device=dev, dtype=torch.float32) if want_logits else None
        amax = torch.empty((B, h, w), device=dev, dtype=torch.int32) if want_argmax else None
        z_feat = None
        if pre_sigmoid:
            if dp_comm is not None:
                raise NotImplementedError("pre_sigmoid and dp_comm are not combined (apply torch.sigmoid first)")
            z_feat = torch.empty_like(z)
            rc = lib.adlm_step_fwd_bwd_pre(_lib.ptr(z), B, K, HW, _lib.ptr(P), _lib.ptr(W), N, C, c_float(eps),
                                           _prec(precision or DEFAULT_PRECISION), _lib.ptr(tgt), _lib.ptr(pc),
                                           c_int(1 if ignore_void else 0), c_float(w_ce), c_float(w_kld),
                                           c_float(w_l1), c_float(grad_scale), _lib.ptr(global_counts),
                                           _lib.ptr(z_feat), _lib.ptr(logits), _lib.ptr(amax),
                                           _lib.ptr(dz), _lib.ptr(dP), _lib.ptr(dW), _lib.ptr(scalars),
                                           _lib.ptr(workspace), c_size_t(workspace.numel()), _lib.stream_ptr())
            _lib.check(rc, "adlm_step_fwd_bwd_pre")
        elif dp_comm is not None:
            if global_counts is not None:
                raise ValueError("dp_comm and global_counts are mutually exclusive")
            comm = dp_comm.next_comm(K, N, C)
            rc = lib.adlm_step_fwd_bwd_dp(_lib.ptr(z), B, K, HW, _lib.ptr(P), _lib.ptr(W), N, C, c_float(eps),
                                          _prec(precision or DEFAULT_PRECISION), _lib.ptr(tgt), _lib.ptr(pc),
                                          c_int(1 if ignore_void else 0), c_float(w_ce), c_float(w_kld),
                                          c_float(w_l1), c_float(grad_scale),
                                          _lib.ptr(logits), _lib.ptr(amax),
                                          _lib.ptr(dz), _lib.ptr(dP), _lib.ptr(dW), _lib.ptr(scalars),
                                          _lib.ptr(workspace), c_size_t(workspace.numel()), ctypes.byref(comm),
                                          _lib.stream_ptr())
            _lib.check(rc, "adlm_step_fwd_bwd_dp")
        else:
            rc = lib.adlm_step_fwd_bwd(_lib.ptr(z), B, K, HW, _lib.ptr(P), _lib.ptr(W), N, C, c_float(eps),
                                       _prec(precision or DEFAULT_PRECISION), _lib.ptr(tgt), _lib.ptr(pc),
                                       c_int(1 if ignore_void else 0), c_float(w_ce), c_float(w_kld),
                                       c_float(w_l1), c_float(grad_scale), _lib.ptr(global_counts),
                                       _lib.ptr(logits), _lib.ptr(amax),
                                       _lib.ptr(dz), _lib.ptr(dP), _lib.ptr(dW), _lib.ptr(scalars),
                                       _lib.ptr(workspace), c_size_t(workspace.numel()), _lib.stream_ptr())
    _lib.check(rc, "adlm_step_fwd_bwd")
    if last_only:
        dz = dP = None
    if dP is not None:
        dP = dP.reshape(prototypes.shape)
    return _StepResult(scalars=scalars, dz=dz, dP=dP, dW=dW, logits=logits, argmax=amax, workspace=workspace,
                       z=z_feat)


def step_counts(target, proto_class, num_classes, ignore_void=True, workspace=None, K=1):
    """(2,) int32 CUDA tensor [n CE patches, n KLD terms] of a batch shard (adlm_step_counts)."""
    lib = _lib.load()
    if not target.is_cuda:
        raise ValueError("target must be a CUDA tensor")
    B = target.shape[0]
    HW = target[0].numel()
    dev = target.device
    tgt = target.to(dtype=torch.int32).contiguous()
    pc = proto_class.to(device=dev, dtype=torch.int32).contiguous()
    N, C = pc.numel(), int(num_classes)
    with torch.cuda.device(dev):
        need = int(lib.adlm_step_workspace_bytes(B, K, HW, N, C))
        if workspace is None or workspace.numel() < need:
            workspace = torch.empty(need, device=dev, dtype=torch.uint8)
        counts = torch.empty(2, device=dev, dtype=torch.int32)
        rc = lib.adlm_step_counts(_lib.ptr(tgt), B, HW, _lib.ptr(pc), N, C, c_int(1 if ignore_void else 0),
                                  _lib.ptr(counts), _lib.ptr(workspace), c_size_t(workspace.numel()),
                                  _lib.stream_ptr())
    _lib.check(rc, "adlm_step_counts")
    return counts


class FusedStepFunction(torch.autograd.Function):
    """loss = FusedStepFunction.apply(z, prototype_vectors, last_layer.weight, target, proto_class, cfg)

    Forward runs the whole fused step (gradients included); backward hands the stored gradients
    to autograd scaled by the incoming grad, so `loss.backward()` continues into the backbone with dz."""

    @staticmethod
    def forward(ctx, z, prototypes, weight, target, proto_class, cfg):
        need = any(ctx.needs_input_grad[:3])
        res = fused_step(z, prototypes, weight, target, proto_class, need_grads=need, **cfg)
        ctx.save_for_backward(*(t for t in (res.dz, res.dP, res.dW) if t is not None))
        ctx.have = need
        ctx.scalars = res.scalars
        loss = res.scalars[0].to(torch.float32)
        ctx.mark_non_differentiable(res.scalars)
        return loss, res.scalars

    @staticmethod
    def backward(ctx, dloss, _dscalars):
        if not ctx.have:
            return (None,) * 6
        dz, dP, dW = ctx.saved_tensors
        need = ctx.needs_input_grad
        g = dloss.to(torch.float32)
        return (dz * g if need[0] else None, dP * g if need[1] else None, dW * g if need[2] else None,
                None, None, None)


# ---- legacy min-pool classification branch (model.py:284-299) -------------------------------------------------

#: arithmetic of the min-pool scan: the arg-min decides where the gradient goes, so fp32-class distances by default
MIN_POOL_DEFAULT_PRECISION = "tf32x3"


def min_pool_forward(z, prototypes, precision=None):
    """z (B,K,h,w), prototypes (N,K[,1,1]) -> (min_distances (B,N) f32, argmin (B,N) int32 patch index, packed (B,N)
    int64 workspace for the backward).  The (B,N,h,w) distance map is never written (adlm_min_pool_fwd)."""
    lib = _lib.load()
    _check_cuda_f32("conv_features", z)
    _check_cuda_f32("prototype_vectors", prototypes)
    if z.dim() != 4:
        raise ValueError("conv_features must be (B,K,h,w)")
    z = z.contiguous()
    P = _proto_2d(prototypes).contiguous()
    B, K, h, w = z.shape
    N = P.shape[0]
    if P.shape[1] != K:
        raise ValueError(f"shape mismatch: z K={K}, prototypes {tuple(P.shape)}")
    dev = z.device
    with torch.cuda.device(dev):
        packed = torch.full((B, N), _lib.PUSH_NONE, device=dev, dtype=torch.int64)
        mind = torch.empty((B, N), device=dev, dtype=torch.float32)
        amin = torch.empty((B, N), device=dev, dtype=torch.int32)
        rc = lib.adlm_min_pool_fwd(_lib.ptr(z), B, K, h * w, _lib.ptr(P), N,
                                   _prec(precision or (MIN_POOL_DEFAULT_PRECISION if (K <= 64 and N <= 224) else "fp32")),
                                   _lib.ptr(packed), _lib.ptr(mind), _lib.ptr(amin), _lib.stream_ptr())
    _lib.check(rc, "adlm_min_pool_fwd")
    return mind, amin, packed


class MinPoolFunction(torch.autograd.Function):
    """min_distances = -max_pool2d(-_l2_convolution(z)) (model.py:284-288) with autograd edges to z and the
    prototypes; the backward scatters onto the arg-min patches (adlm_min_pool_bwd)."""

    @staticmethod
    def forward(ctx, z, prototypes, precision):
        mind, amin, packed = min_pool_forward(z.detach(), prototypes.detach(), precision)
        ctx.save_for_backward(z, prototypes, packed)
        ctx.mark_non_differentiable(amin)
        return mind, amin

    @staticmethod
    def backward(ctx, gmin, _gamin):
        z, prototypes, packed = ctx.saved_tensors
        lib = _lib.load()
        zc = z.detach().contiguous()
        P = _proto_2d(prototypes.detach()).contiguous()
        B, K, h, w = zc.shape
        N = P.shape[0]
        need_dz, need_dP = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        g = gmin.contiguous().to(torch.float32)
        with torch.cuda.device(zc.device):
            dz = torch.zeros_like(zc) if need_dz else None
            dP = torch.zeros_like(P) if need_dP else None
            rc = lib.adlm_min_pool_bwd(_lib.ptr(zc), B, K, h * w, _lib.ptr(P), N, _lib.ptr(packed), _lib.ptr(g),
                                       _lib.ptr(dz), _lib.ptr(dP), _lib.stream_ptr())
        _lib.check(rc, "adlm_min_pool_bwd")
        return dz, (dP.reshape(prototypes.shape) if need_dP else None), None
