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
from .push import default_search_precision, unpack


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
    prec = _prec(precision or default_search_precision(K, N))
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


# ------------------------------------------------------------------------------ the reference entry points
def find_k_nearest_patches_to_prototypes(dataset,
                                         prototype_network_parallel,
                                         k=5,
                                         preprocess_input_function=None,
                                         full_save=False,
                                         root_dir_for_saving_images='./nearest',
                                         log=print,
                                         prototype_activation_function_in_numpy=None,
                                         chunk=16):
    """Drop-in for find_nearest.find_k_nearest_patches_to_prototypes (find_nearest.py:66-342): same arguments (plus
    `chunk`, images per scan launch), same return value — the (N,k) array of patch class ids, nearest first (-1 void;
    -2 where the dataset offers fewer than k candidates, which the reference cannot produce because it keeps void-only
    images at distance + 1e7).  Images are read exactly as the reference does (RGB, margin crop, ImageNet
    normalisation — `preprocess_input_function` is unused there too), pushed through `conv_features`, and scanned on
    the device; only the N*k winners have their label regions read back.  With `full_save` the numeric artefacts are
    written (`<root>/<j>/class_id.npy`, `<root>/full_class_id.npy`); the PNG dumps of find_nearest.py:228-325 are
    visualisation and are not produced.  Under torch.distributed every rank scans a contiguous slice of the images
    and all ranks return the same array."""
    import os
    import time
    import torch.distributed as dist
    from .push import _load_sample, _to_normalized_tensor
    ppnet = prototype_network_parallel.module if hasattr(prototype_network_parallel, 'module') \
        else prototype_network_parallel
    prototype_network_parallel.eval()
    log('find nearest patches')
    start = time.time()
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0
    img_ids = list(dataset.img_ids)
    n_total = len(img_ids)
    per = (n_total + world - 1) // world
    lo, hi = min(n_total, rank * per), min(n_total, (rank + 1) * per)
    proto_class = ppnet.proto_class()
    dev = ppnet.prototype_vectors.device
    N = ppnet.num_prototypes
    prec = ppnet.head_precision if ppnet.head_precision in ("fp32", "tf32x3") else None
    rows, hw = [], None
    with torch.no_grad():
        for c0 in range(lo, hi, chunk):
            zs, labs = [], []
            for idx in range(c0, min(hi, c0 + chunk)):
                img, ann = _load_sample(dataset, img_ids[idx])
                zs.append(ppnet.conv_features(_to_normalized_tensor(img).unsqueeze(0).to(dev)).float())
                labs.append(torch.from_numpy(np.ascontiguousarray(ann).astype(np.uint8)))
            if len({tuple(t.shape) for t in zs}) != 1 or len({tuple(t.shape) for t in labs}) != 1:
                raise NotImplementedError("all images of one search must share the image and feature-map size")
            z = torch.cat(zs)
            if hw is None:
                hw = (z.shape[2], z.shape[3])
            rows.append(image_minima(z, torch.stack(labs).to(dev), ppnet.prototype_vectors, proto_class, prec))
    packed = torch.cat(rows) if rows else torch.full((0, N), _lib.PUSH_NONE, device=dev, dtype=torch.int64)
    if world > 1:
        t = torch.tensor(list(hw or (0, 0)), device=dev, dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        hw = (int(t[0]), int(t[1]))
    cache = {}

    def label_of(i):
        if i not in cache:
            cache[i] = np.asarray(_load_sample(dataset, img_ids[lo + i])[1]).astype(np.uint8)
        return cache[i]
    pc_np = proto_class.detach().cpu().numpy().astype(np.int64)
    res = top_k_and_labels(packed, k, label_of, pc_np, hw[0], hw[1], image_base=lo,
                           n_total=n_total if world > 1 else None)
    labels_all_prototype = res["labels"]
    if full_save and rank == 0:
        for j in range(N):
            d = os.path.join(root_dir_for_saving_images, str(j))
            os.makedirs(d, exist_ok=True)
            np.save(os.path.join(d, 'class_id.npy'), labels_all_prototype[j])
        np.save(os.path.join(root_dir_for_saving_images, 'full_class_id.npy'), labels_all_prototype)
    log('\tfind nearest patches time: \t{0}'.format(time.time() - start))
    return labels_all_prototype


def prune_prototypes(dataset,
                     prototype_network_parallel,
                     k,
                     prune_threshold,
                     preprocess_input_function,
                     original_model_dir,
                     epoch_number,
                     log=print,
                     copy_prototype_imgs=True):
    """Drop-in for prune.prune_prototypes (prune.py:11-110), as segmentation/run_pruning.py:60-69 calls it: k nearest
    patches of every prototype, prune those with fewer than `prune_threshold` own-class patches among them, write
    `pruned_prototypes_epoch{e}_k{k}_pt{t}/prune_info.npy` and `prototypes_to_keep.json`, shrink the network in
    place; returns prune_info (pruned index, its class).  The copies of the prototype images (prune.py:73-108) are made
    for the files that exist."""
    import json
    import os
    import shutil
    ppnet = prototype_network_parallel.module if hasattr(prototype_network_parallel, 'module') \
        else prototype_network_parallel
    nearest = find_k_nearest_patches_to_prototypes(
        dataset=dataset, prototype_network_parallel=prototype_network_parallel, k=k,
        preprocess_input_function=preprocess_input_function, full_save=True,
        root_dir_for_saving_images=os.path.join(original_model_dir, 'img'), log=log)
    original_num_prototypes = ppnet.num_prototypes
    pc = ppnet.proto_class().cpu().numpy().astype(np.int64)
    to_prune = prototypes_to_prune(nearest, pc, prune_threshold)
    log('k = {}, prune_threshold = {}'.format(k, prune_threshold))
    log('{} prototypes will be pruned'.format(len(to_prune)))
    prune_info = np.hstack((np.array(to_prune, dtype=np.int64).reshape(-1, 1), pc[to_prune].reshape(-1, 1)))
    out_dir = os.path.join(original_model_dir, 'pruned_prototypes_epoch{}_k{}_pt{}'.format(epoch_number, k, prune_threshold))
    os.makedirs(out_dir, exist_ok=True)
    np.save(os.path.join(out_dir, 'prune_info.npy'), prune_info)
    ppnet.prune_prototypes(to_prune)
    keep = list(sorted(set(range(original_num_prototypes)) - set(to_prune)))
    log(f'Prototypes to keep: {keep}')
    with open(os.path.join(original_model_dir, 'prototypes_to_keep.json'), 'w') as fp:
        json.dump(keep, fp)
    if copy_prototype_imgs:
        src_dir = os.path.join(original_model_dir, 'img', 'epoch-%d' % epoch_number)
        dst_dir = os.path.join(out_dir, 'img', 'epoch-%d' % epoch_number)
        os.makedirs(dst_dir, exist_ok=True)
        for new, old in enumerate(keep):
            for pat in ('prototype-img%d.png', 'prototype-img-original%d.png',
                        'prototype-img-original_with_self_act%d.png', 'prototype-self-act%d.npy'):
                src = os.path.join(src_dir, pat % old)
                if os.path.exists(src):
                    shutil.copyfile(src, os.path.join(dst_dir, pat % new))
        for name in ('bb%d.npy' % epoch_number, 'bb-receptive_field%d.npy' % epoch_number):
            src = os.path.join(src_dir, name)
            if os.path.exists(src):
                np.save(os.path.join(dst_dir, name), np.load(src)[keep])
    return prune_info
