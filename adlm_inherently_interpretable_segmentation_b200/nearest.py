"""k nearest patches of every prototype and the pruning rule built on them.

find_nearest.find_k_nearest_patches_to_prototypes (find_nearest.py:66-342) pushes every image through the network,
copies the (N,h,w) distance map to the host, masks void pixels (labels resized with PIL NEAREST), takes the
per-prototype minimum in numpy and keeps a heap of the k best images per prototype; prune.prune_prototypes
(prune.py:11-42) prunes a prototype when fewer than `prune_threshold` of those k patches carry its own class.

Here the per-image search is the push scan kernel in its per-image mode (adlm_push_scan_images: ONE launch for
all images, distances on chip, one packed (distance bits, patch index) minimum per (image, prototype)) run with a
void mask instead of the class mask; the top-k over
images is one stable sort of the (n_img, N) minima on the device, and only the N*k winning patches have their label
regions looked at.  Deviation from the reference: an image without any non-void patch yields no candidate (the
reference keeps it with distance + 1e7, find_nearest.py:132).
"""
from ctypes import c_int

import numpy as np
import torch

from . import _lib
from .head import _check_cuda_f32, _prec, _proto_2d
from .push import PUSH_DEFAULT_PRECISION, unpack


def pil_nearest_lut(in_size, out_size):
    """source index of every output pixel of PIL's Image.resize(..., NEAREST), which resize_label uses
    (segmentation/dataset.py:23-31): Pillow accumulates the source coordinate in double (xo = a/2; xo += a)"""
    a = in_size / out_size
    xo = a * 0.5
    lut = np.empty(out_size, np.int64)
    for x in range(out_size):
        lut[x] = min(int(xo), in_size - 1)
        xo += a
    return lut


def void_mask(labels, h, w):
    """labels (n,H,W) uint8 CUDA (0 = void) -> (n,h*w) int32: all class bits set where the NEAREST-resized label is
    not void (find_nearest.py:128-132), 0 otherwise — the push scan then admits every prototype on those patches"""
    n, H, W = labels.shape
    rl = torch.from_numpy(pil_nearest_lut(H, h)).to(labels.device)
    cl = torch.from_numpy(pil_nearest_lut(W, w)).to(labels.device)
    small = labels[:, rl][:, :, cl]
    return torch.where(small != 0, torch.full_like(small, -1, dtype=torch.int32),
                       torch.zeros_like(small, dtype=torch.int32)).reshape(n, h * w).contiguous()


def patch_label(label_img, pi, pj, h, w, target_class):
    """class of the image region behind patch (pi, pj) (find_nearest.py:163-168,196-205): the prototype's class if
    any pixel carries it, else the most frequent class (-1 = void); None for an empty region"""
    H, W = label_img.shape
    ph, pw = H / h, W / w
    lab = label_img[int(pi * ph):int((pi + 1) * ph), int(pj * pw):int((pj + 1) * pw)].astype(np.int64) - 1
    if lab.size == 0:
        return None
    if np.any(lab == target_class):
        return int(target_class)
    vals, cnts = np.unique(lab, return_counts=True)
    return int(vals[np.argmax(cnts)])


def image_minima(z, labels, prototypes, proto_class, precision=None):
    """per-(image, prototype) nearest non-void patch of a batch in ONE scan launch (adlm_push_scan_images):
    -> (n,N) int64 packed words (float bits of the distance << 32 | patch index inside the image)"""
    lib = _lib.load()
    _check_cuda_f32("conv_features", z)
    z = z.contiguous()
    P = _proto_2d(prototypes.detach()).contiguous()
    _check_cuda_f32("prototype_vectors", P)
    n, K, h, w = z.shape
    N = P.shape[0]
    dev = z.device
    if labels.dtype != torch.uint8 or not labels.is_cuda or labels.shape[0] != n:
        raise ValueError("labels must be a CUDA uint8 tensor (n,H,W)")
    if int(proto_class.max()) >= _lib.MAX_PUSH_CLASSES:
        raise NotImplementedError(f"the scan supports up to {_lib.MAX_PUSH_CLASSES} classes (uint32 patch mask)")
    pc = proto_class.to(device=dev, dtype=torch.int32).contiguous()
    mask = void_mask(labels, h, w)
    packed = torch.full((n, N), _lib.PUSH_NONE, device=dev, dtype=torch.int64)
    prec = _prec(precision or PUSH_DEFAULT_PRECISION)
    with torch.cuda.device(dev):                             # one launch: per-(image, prototype) minima
        rc = lib.adlm_push_scan_images(_lib.ptr(z), n, K, h * w, _lib.ptr(mask), _lib.ptr(P), N, _lib.ptr(pc),
                                       c_int(prec), _lib.ptr(packed), _lib.stream_ptr())
    _lib.check(rc, "adlm_push_scan_images")
    return packed


def top_k_and_labels(packed, k, label_of, pc_np, h, w, image_base=0, n_total=None, group=None):
    """top-k images per prototype and the label rule on the winners (find_nearest.py:163-226).
    packed (n_local,N) int64 minima of this rank's images [image_base, image_base + n_local) out of n_total;
    label_of(i) -> (H,W) label image of LOCAL image i.  With a process group the minima are merged by one MIN
    all-reduce of the (n_total,N) table (4.5 MB for Cityscapes train) and the labels of the N*k winners, each known to
    the rank that owns the image, by one MAX all-reduce; every rank returns the same dict."""
    import torch.distributed as dist
    n_local, N = packed.shape
    sharded = group is not None or (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
                                    and n_total is not None)
    if n_total is None:
        n_total = n_local
    if sharded:
        full = torch.full((n_total, N), _lib.PUSH_NONE, device=packed.device, dtype=torch.int64)
        full[image_base:image_base + n_local] = packed
        dist.all_reduce(full, op=dist.ReduceOp.MIN, group=group)
    else:
        full = packed
    # packed words order like (distance, patch index); a stable sort over the image axis keeps the earlier image on
    # ties, as the reference's heap does (find_nearest.py:217-223)
    order = torch.sort(full, dim=0, stable=True)
    top_words = order.values[:k].t().contiguous().cpu().numpy()          # (N,kk)
    top_imgs = order.indices[:k].t().contiguous().cpu().numpy()
    dist_np, patch = unpack(top_words.reshape(-1))
    dist_np, patch = dist_np.reshape(top_words.shape), patch.reshape(top_words.shape)
    kk = top_words.shape[1]
    NOT_MINE, EMPTY = -10, -3
    lab = np.full((N, kk), NOT_MINE, np.int64)
    for j in range(N):
        for t in range(kk):
            i = int(top_imgs[j, t])
            if patch[j, t] < 0 or not (image_base <= i < image_base + n_local):
                continue
            pi, pj = divmod(int(patch[j, t]), w)
            got = patch_label(label_of(i - image_base), pi, pj, h, w, int(pc_np[j]))
            lab[j, t] = EMPTY if got is None else got
    if sharded:
        lt = torch.from_numpy(lab)
        if dist.get_backend(group) == "nccl":
            lt = lt.to(packed.device)
        dist.all_reduce(lt, op=dist.ReduceOp.MAX, group=group)
        lab = lt.cpu().numpy()
    out = dict(labels=np.full((N, k), -2, np.int64), distances=np.full((N, k), np.inf),
               images=np.full((N, k), -1, np.int64), patch_i=np.full((N, k), -1, np.int64),
               patch_j=np.full((N, k), -1, np.int64))
    for j in range(N):
        t_out = 0
        for t in range(kk):
            if patch[j, t] < 0 or lab[j, t] == EMPTY or lab[j, t] == NOT_MINE:
                continue
            pi, pj = divmod(int(patch[j, t]), w)
            out["labels"][j, t_out], out["distances"][j, t_out] = lab[j, t], dist_np[j, t]
            out["images"][j, t_out], out["patch_i"][j, t_out], out["patch_j"][j, t_out] = int(top_imgs[j, t]), pi, pj
            t_out += 1
    return out


def find_k_nearest_tensors(z, labels, prototypes, proto_class, k=5, precision=None, labels_host=None,
                           image_base=0, n_total=None, group=None):
    """z (n,K,h,w) CUDA features of the n images (PPNet.push_forward's first output), labels (n,H,W) CUDA uint8
    (0 = void, class + 1 otherwise).  -> dict of numpy arrays, each (N,k), ordered by ascending distance:
      labels     class id of the patch (find_nearest's labels_all_prototype; -1 void, -2 = fewer than k candidates)
      distances, images, patch_i, patch_j
    Multi-GPU: every rank passes its contiguous shard (parallel.image_range) with image_base / n_total; the result is
    identical on all ranks (see top_k_and_labels)."""
    packed = image_minima(z, labels, prototypes, proto_class, precision)
    _, _, h, w = z.shape
    pc_np = proto_class.detach().cpu().numpy().astype(np.int64)
    host = {}

    def label_of(i):
        if i not in host:
            host[i] = labels_host[i] if labels_host is not None else labels[i].cpu().numpy()
        return host[i]
    return top_k_and_labels(packed, k, label_of, pc_np, h, w, image_base, n_total, group)


def prototypes_to_prune(nearest_labels, proto_class, prune_threshold):
    """prune.py:31-42: indices of the prototypes with fewer than prune_threshold own-class patches among their k
    nearest; feed to PPNet.prune_prototypes"""
    pc = np.asarray(proto_class.detach().cpu() if torch.is_tensor(proto_class) else proto_class).astype(np.int64)
    lab = np.asarray(nearest_labels)
    return [int(j) for j in range(len(pc)) if int(np.sum(lab[j] == pc[j])) < prune_threshold]
