"""Multi-GPU plumbing for the prototype head: one process per GPU, torch.distributed (NCCL over NVLink on
the B200 box, gloo in the CPU tests).  SURVEY.md §8(e):

  * forward / eval / cfg5 stress: batch-sharded, no communication;
  * data-parallel training step: the batch is sharded by images.  CE is a mean over the non-void patches
    of the WHOLE batch and KLD a mean over all (image,label,pair) terms (module.py:162-165,202-206), so
    the two normalisers must be summed over ranks before the loss kernels run, and dP, dW and the loss sums
    after them.  Two transports:
      - DpExchange (default on CUDA): both exchanges happen INSIDE the step's own kernels over peer-mapped
        memory (torch symmetric memory -> NVLink stores + sequence flags, adlm_step_fwd_bwd_dp): the DP step
        is the same six launches as the single-GPU step and makes no collective call;
      - NCCL (fallback when symmetric memory is unavailable, and the gloo CPU tests): one 8-byte all-reduce
        of the normalisers handed to the kernels as `global_counts`, the masked-L1 term split evenly, then
        ONE fp32 all-reduce of the flat buffer [dP | dW | scalars] (63 KB at 190x64, C=19);
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


class DpExchange:
    """Peer-mapped exchange buffers for adlm_step_fwd_bwd_dp: one buffer per rank, allocated with torch
    symmetric memory (CUDA VMM handles exchanged through the process group's store) so that every rank has
    every peer's buffer mapped over NVLink.  Construction is a collective (all ranks of `group`).

    DpExchange.local(device, world) builds `world` plain buffers on ONE device instead — the single-process
    form used by the tests and by N "ranks" sharing a GPU on separate streams."""

    #: wall-clock bound of every in-kernel wait for a peer (a stalled data loader on one rank is normal; a dead
    #: rank should not hang the others for ever): results are NaN and scalars[7] > 0 after it expires
    timeout_ms = 30000

    def __init__(self, K, N, C, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        lib = _lib_mod().load()
        self.rank, self.world = world(group)
        if self.world > _lib_mod().DP_MAX_WORLD:
            raise NotImplementedError(f"DpExchange: world {self.world} > {_lib_mod().DP_MAX_WORLD}")
        self.shape = (K, N, C)
        nbytes = int(lib.adlm_dp_exchange_bytes(K, N, C, self.world))
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        grp = group if group is not None else dist.group.WORLD
        self.handle = symm_mem.rendezvous(self.buf, group=grp)
        self._set_ptrs([int(p) for p in self.handle.buffer_ptrs])
        dist.barrier(group=group)            # every peer's buffer is zeroed and mapped before the first step
        self.seq = 0

    @classmethod
    def local(cls, K, N, C, device, n_ranks):
        """n_ranks exchange views over plain device buffers of one GPU (no process group needed)"""
        lib = _lib_mod().load()
        nbytes = int(lib.adlm_dp_exchange_bytes(K, N, C, n_ranks))
        bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=device) for _ in range(n_ranks)]
        out = []
        for r in range(n_ranks):
            ex = cls.__new__(cls)
            ex.rank, ex.world, ex.shape, ex.buf, ex.handle, ex.seq = r, n_ranks, (K, N, C), bufs, None, 0
            ex._set_ptrs([b.data_ptr() for b in bufs])
            out.append(ex)
        return out

    def layout(self):
        """dict of byte offsets inside one exchange buffer (adlm_dp_exchange_layout)"""
        import ctypes
        out = (ctypes.c_size_t * 8)()
        K, N, C = self.shape
        _lib_mod().check(_lib_mod().load().adlm_dp_exchange_layout(K, N, C, self.world, ctypes.cast(out, ctypes.c_void_p)),
                         "adlm_dp_exchange_layout")
        keys = ("cnt", "cflag", "acc", "gflag", "grad", "total", "n_grad_pad", "n_blocks")
        return dict(zip(keys, (int(x) for x in out)))

    def _set_ptrs(self, ptrs):
        import ctypes
        self._ptrs = (ctypes.c_void_p * len(ptrs))(*ptrs)

    def next_comm(self, K, N, C):
        """the adlm_dp_comm of the next step on this exchange (advances the sequence number)"""
        import ctypes
        if (K, N, C) != self.shape:
            raise ValueError(f"DpExchange was built for (K,N,C)={self.shape}, step has {(K, N, C)}")
        self.seq += 1
        L = _lib_mod()
        return L.DpComm(self.rank, self.world, self.seq, ctypes.cast(self._ptrs, ctypes.POINTER(ctypes.c_void_p)),
                        int(self.timeout_ms))


def _lib_mod():
    from . import _lib
    return _lib


_EXCHANGES = {}
_SYMM_BROKEN = []


def default_exchange(K, N, C, device, group=None):
    """cached DpExchange of (group, shape, device); None when symmetric memory cannot be set up here
    (reported once on stderr; the caller then takes the NCCL transport).  Collective on first use."""
    import os
    import sys
    if os.environ.get("ADLM_DP_TRANSPORT", "").lower() == "nccl" or _SYMM_BROKEN:
        return None
    key = (id(group) if group is not None else 0, K, N, C, str(device))
    if key not in _EXCHANGES:
        try:
            _EXCHANGES[key] = DpExchange(K, N, C, device, group)
        except Exception as e:  # noqa: BLE001 - any failure of the symmetric-memory stack selects NCCL
            _SYMM_BROKEN.append(repr(e))
            print(f"[adlm] symmetric memory unavailable ({e!r}); data-parallel step uses NCCL all-reduces",
                  file=sys.stderr, flush=True)
            return None
    return _EXCHANGES[key]


def dp_supported_fused(K, N, C, precision):
    """shapes / modes adlm_step_fwd_bwd_dp implements (include/adlm_head.h)"""
    return precision == "tf32" and K <= 64 and N <= 192 and C <= 32


def dp_fused_step(ppnet_or_params, z, target, proto_class, ignore_void=True, weights=(1.0, 0.25, 1e-4),
                  grad_scale=1.0, precision=None, group=None, workspace=None, dz_out=None, exchange="auto"):
    """Data-parallel fused step on this rank's image shard.  Returns the head.fused_step result with dP, dW
    and scalars already reduced (identical on every rank, equal to the single-GPU step on the whole batch);
    dz is the gradient for this rank's features only (it feeds the local backbone backward).

    exchange: a DpExchange, "auto" (cached default_exchange when the shape/mode supports it) or None (NCCL).
    Peer-memory transport: no collective call at all (see DpExchange).  NCCL transport: two collectives per
    step, the 8-byte sum of the loss normalisers before the kernels run and one fp32 sum of the flat buffer
    [dP | dW | 8 loss scalars] after them (the kernels write dP and dW straight into it)."""
    from . import head
    prototypes, weight = ppnet_or_params
    rank, size = world(group)
    C = weight.shape[0]
    if size == 1:
        return head.fused_step(z, prototypes, weight, target, proto_class, ignore_void=ignore_void, weights=weights,
                               grad_scale=grad_scale, precision=precision, workspace=workspace, dz_out=dz_out)
    N, K = prototypes.shape[0], prototypes.shape[1]
    prec_name = precision or head.DEFAULT_PRECISION
    if isinstance(exchange, str) and exchange == "auto":
        exchange = default_exchange(K, N, C, z.device, group) if (z.is_cuda and dp_supported_fused(K, N, C, prec_name)) else None
    if exchange is not None:
        return head.fused_step(z, prototypes, weight, target, proto_class, ignore_void=ignore_void, weights=weights,
                               grad_scale=grad_scale, precision=precision, workspace=workspace, dz_out=dz_out,
                               dp_comm=exchange)
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
