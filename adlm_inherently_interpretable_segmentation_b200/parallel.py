"""Multi-GPU plumbing for the prototype head: one process per GPU, torch.distributed (NCCL over NVLink on
the B200 box, gloo in the CPU tests).  SURVEY.md §8(e):

  * forward / eval / cfg5 stress: batch-sharded, no communication;
  * data-parallel training step: the batch is sharded by images.  CE is a mean over the non-void patches
    of the WHOLE batch and KLD a mean over all (image,label,pair) terms (module.py:162-165,202-206), so
    the two normalisers are summed over ranks first (one 8-byte all-reduce) and handed to the kernels as
    `global_counts`; the masked-L1 term is split evenly.  Then ONE fp32 all-reduce of the flat buffer
    [dP | dW | scalars] (63 KB at 190x64, C=19) makes every rank hold the single-GPU gradients;
  * push: contiguous image ranges per rank, one all_reduce(MIN) over N packed uint64 words, one sum
    all-reduce of the (N,K) winner vectors (each row is owned by exactly one rank).

Everything here is device-agnostic tensor logic (the kernels are reached through head.py), so the same
functions run under gloo on CPU tensors in tests/test_parallel_gloo.py.
"""
import torch
import torch.distributed as dist


def world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def image_range(n_images, rank, size):
    """contiguous shard [lo, hi) of rank: keeps 'earliest image wins' = lowest global patch id"""
    per = (n_images + size - 1) // size
    return min(n_images, rank * per), min(n_images, (rank + 1) * per)


def allreduce_counts(counts, group=None):
    """(2,) int32 [CE patches, KLD terms] of this shard -> sums over ranks (in place)"""
    if world(group)[1] > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts


def pack_grads(dP, dW, scalars):
    """one flat fp32 buffer [dP | dW | scalars] for a single collective"""
    return torch.cat([dP.reshape(-1), dW.reshape(-1), scalars.to(torch.float32).reshape(-1)])


def unpack_grads(flat, dP, dW, scalars):
    nP, nW = dP.numel(), dW.numel()
    dP.copy_(flat[:nP].view_as(dP))
    dW.copy_(flat[nP:nP + nW].view_as(dW))
    scalars.copy_(flat[nP + nW:nP + nW + scalars.numel()].to(scalars.dtype))


def allreduce_head_grads(dP, dW, scalars, group=None, replicated=(3, 5, 6)):
    """sum-reduce prototype / last-layer gradients and the additive loss scalars over ranks.
    `replicated` scalars (raw L1 norm, the two global normalisers) are identical on every rank already
    and keep their value."""
    rank, size = world(group)
    if size == 1:
        return
    keep = scalars[list(replicated)].clone()
    flat = pack_grads(dP, dW, scalars)
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    unpack_grads(flat, dP, dW, scalars)
    scalars[list(replicated)] = keep


def merge_packed_min(packed, group=None):
    """push: element-wise MIN of the packed (distance bits << 32 | global patch id) words (int64 view of the
    uint64 words: distances are >= 0, so the sign bit is clear and signed order == unsigned order)"""
    if world(group)[1] > 1:
        dist.all_reduce(packed, op=dist.ReduceOp.MIN, group=group)
    return packed


def merge_owned_rows(rows, group=None):
    """push: (N,K) winner vectors, each row non-zero on exactly one rank -> sum over ranks"""
    if world(group)[1] > 1:
        dist.all_reduce(rows, op=dist.ReduceOp.SUM, group=group)
    return rows


_ADDITIVE = {}


def _additive_mask(device, rank, replicated=(3, 5, 6)):
    """(8,) float64: 1 for the additive loss scalars, and for the replicated ones (raw L1 norm, the two global
    normalisers, identical on every rank) 1 on rank 0 only, so that a SUM over ranks keeps their value"""
    key = (device, rank, tuple(replicated))
    if key not in _ADDITIVE:
        m = torch.ones(8, dtype=torch.float64)
        if rank != 0:
            m[list(replicated)] = 0.0
        _ADDITIVE[key] = m.to(device)
    return _ADDITIVE[key]


def dp_fused_step(ppnet_or_params, z, target, proto_class, ignore_void=True, weights=(1.0, 0.25, 1e-4),
                  grad_scale=1.0, precision=None, group=None, workspace=None, dz_out=None):
    """Data-parallel fused step on this rank's image shard.  Returns the head.fused_step result with dP, dW
    and scalars already reduced (identical on every rank, equal to the single-GPU step on the whole batch);
    dz is the gradient for this rank's features only (it feeds the local backbone backward).

    Two collectives per step: the 8-byte sum of the loss normalisers before the kernels run, and one fp32
    sum of the flat buffer [dP | dW | 8 loss scalars] after them (the kernels write dP and dW straight into it)."""
    from . import head
    prototypes, weight = ppnet_or_params
    rank, size = world(group)
    C = weight.shape[0]
    if size == 1:
        return head.fused_step(z, prototypes, weight, target, proto_class, ignore_void=ignore_void, weights=weights,
                               grad_scale=grad_scale, precision=precision, workspace=workspace, dz_out=dz_out)
    counts = head.step_counts(target, proto_class, C, ignore_void, workspace=workspace, K=z.shape[1])
    allreduce_counts(counts, group)
    w_ce, w_kld, w_l1 = weights
    n_grad = prototypes.numel() + weight.numel()
    flat = torch.empty(n_grad + 8, device=z.device, dtype=torch.float32)
    res = head.fused_step(z, prototypes, weight, target, proto_class, ignore_void=ignore_void,
                          weights=(w_ce, w_kld, w_l1 / size), grad_scale=grad_scale, precision=precision,
                          global_counts=counts, workspace=workspace, grads_out=flat, dz_out=dz_out)
    flat[n_grad:].copy_(res.scalars * _additive_mask(z.device, rank))
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    res["scalars"] = flat[n_grad:].to(torch.float64)
    return res
