"""Prototype push (segmentation/push.py) on libadlm_head.so.

`push_prototypes` keeps the reference's 12-argument signature and side effects (push.py:24-158):
it mutates the network in place (prototype_vectors <- nearest same-class patches, de-dup, prune),
writes `unique_prototypes.json` and, when asked, the `bb*.npy` / `bb-receptive_field*.npy` boxes.

What changes is the search.  The reference pulls features and the (h,w,N) distance map of every image
to the host, builds class->patch sets in a 2M-iteration Python loop and argmins in numpy
(push.py:209-280; 12-19 s per 1024x2048 image).  Here each image costs two kernels:
  adlm_patch_class_mask   labels -> per-patch class bitmask                       (push.py:214-230)
  adlm_push_scan          distances on chip -> per-prototype packed (dist,id) min (push.py:249-280)
and the running minimum lives in ONE device array of N uint64 for the whole dataset.  Multi-GPU: every
rank scans a contiguous range of images, then a single all_reduce(MIN) of the N packed words picks
the winners; ties go to the lowest global patch id = earliest image, as the
reference's strict '<' (push.py:265).

The receptive-field boxes (push.py:288-307) and the activation-crop boxes (push.py:309-350) depend
only on the final winner of each prototype, so they are computed after the reduce, for <= N patches.
"""
import json
import os
import time
from ctypes import c_int, c_longlong

import numpy as np
import torch

from . import _lib
from .head import _check_cuda_f32, _prec, _proto_2d

#: the search must reproduce the fp32 argmin: tcgen05 3xTF32 (fp32-class distances) unless told otherwise
PUSH_DEFAULT_PRECISION = "tf32x3"
#: operand budget of the tcgen05 scan (csrc/tc_push.cu); beyond it the library refuses the tensor-core modes
TC_SCAN_MAX_K, TC_SCAN_MAX_N = 64, 224


def default_search_precision(K, N):
    """precision used when the caller does not name one: 3xTF32 on the tensor cores where that kernel exists, else the
    fp32 CUDA-core scan — both fp32-class, chosen HERE and visibly rather than rerouted inside the library"""
    return PUSH_DEFAULT_PRECISION if (K <= TC_SCAN_MAX_K and N <= TC_SCAN_MAX_N) else "fp32"

NORM_MEAN = (0.485, 0.456, 0.406)   # segmentation/constants.py:6-7 (find_nearest.to_normalized_tensor)
NORM_STD = (0.229, 0.224, 0.225)


# ------------------------------------------------------------------------------ host-side integer logic
def patch_luts(H, W, h, w):
    """pixel -> patch index with the reference's float expression int(pixel / (H / h))
    (push.py:201-202,217,219).  Computed in Python floats so the boundaries match bit for bit."""
    ph, pw = H / h, W / w
    rows = np.fromiter((int(i / ph) for i in range(H)), dtype=np.int32, count=H)
    cols = np.fromiter((int(j / pw) for j in range(W)), dtype=np.int32, count=W)
    # int() can land on h when H/h rounds down (never for the reference's sizes); the reference would then
    # index past the distance map and raise
    if rows.max() >= h or cols.max() >= w:
        raise ValueError("pixel->patch map leaves the feature map (push.py:217-219 would raise IndexError)")
    return rows, cols


def rf_box(patch_i, patch_j, H, W, h, w):
    """receptive-field box of a patch (push.py:288-292), same float arithmetic"""
    ph, pw = H / h, W / w
    return (int(patch_i * ph), int(patch_i * ph + ph) + 1, int(patch_j * pw), int(patch_j * pw + pw) + 1)


def unpack(packed):
    """(N,) int64 packed words -> (dist float32 (inf where none), global patch id int64 (-1 where none))"""
    p = np.asarray(packed, dtype=np.int64)
    none = p == _lib.PUSH_NONE
    bits = (p >> 32).astype(np.uint32)
    dist = bits.view(np.float32).astype(np.float64)
    ids = (p & 0xFFFFFFFF).astype(np.int64)
    dist[none] = np.inf
    ids[none] = -1
    return dist, ids


# ------------------------------------------------------------------------------ device search
_LUT_CACHE = {}


class PushSearch:
    """Running per-prototype nearest-patch state for one push pass on one device.

    scan(z, labels, first_image_index) may be called any number of times (any batch size);
    global patch id = image_index * h*w + patch, so shards on different GPUs just use their own
    image indices."""

    def __init__(self, prototypes, proto_class, num_classes=None, precision=None, device=None):
        self.lib = _lib.load()
        P = _proto_2d(prototypes.detach())
        _check_cuda_f32("prototype_vectors", P)
        self.P = P.contiguous().clone()          # the search is against the pre-push prototypes
        self.device = self.P.device
        self.N, self.K = self.P.shape
        self.proto_class = proto_class.to(device=self.device, dtype=torch.int32).contiguous()
        if int(self.proto_class.max()) >= _lib.MAX_PUSH_CLASSES:
            raise NotImplementedError(f"push supports up to {_lib.MAX_PUSH_CLASSES} classes (uint32 patch mask)")
        self.prec = _prec(precision or default_search_precision(self.K, self.N))
        self.packed = torch.full((self.N,), _lib.PUSH_NONE, device=self.device, dtype=torch.int64)
        self.hw = None

    def _luts(self, H, W, h, w):
        # the tables depend on the four sizes only: one copy per device for the life of the process (a push pass builds
        # a new PushSearch; recomputing and re-uploading them was host time inside every pass)
        key = (H, W, h, w, str(self.device))
        if key not in _LUT_CACHE:
            r, c = patch_luts(H, W, h, w)
            _LUT_CACHE[key] = (torch.from_numpy(r).to(self.device), torch.from_numpy(c).to(self.device))
        return _LUT_CACHE[key]

    def class_mask(self, labels, h, w):
        """labels (n,H,W) uint8 on the device (0 = void, class = label-1) -> (n,h*w) int32 bitmask"""
        if labels.dtype != torch.uint8 or not labels.is_cuda:
            raise ValueError("labels must be a CUDA uint8 tensor (n,H,W)")
        labels = labels.contiguous()
        n, H, W = labels.shape
        rows, cols = self._luts(H, W, h, w)
        mask = torch.empty((n, h * w), device=self.device, dtype=torch.int32)
        with torch.cuda.device(self.device):
            rc = self.lib.adlm_patch_class_mask(_lib.ptr(labels), n, H, W, _lib.ptr(rows), _lib.ptr(cols),
                                                h, w, _lib.ptr(mask), _lib.stream_ptr())
        _lib.check(rc, "adlm_patch_class_mask")
        return mask

    def scan(self, z, labels, first_image_index, mask=None):
        """z (n,K,h,w) features of n consecutive images, labels (n,H,W) uint8 (or a precomputed mask)"""
        _check_cuda_f32("conv_features", z)
        z = z.contiguous()
        n, K, h, w = z.shape
        if K != self.K:
            raise ValueError("feature depth != prototype depth")
        if self.hw is None:
            self.hw = (h, w)
        elif self.hw != (h, w):
            raise NotImplementedError("all images of one push pass must share the feature-map size "
                                      "(global patch id = image * h*w + patch)")
        if mask is None:
            mask = self.class_mask(labels, h, w)
        base = int(first_image_index) * h * w
        with torch.cuda.device(self.device):
            rc = self.lib.adlm_push_scan(_lib.ptr(z), n, K, h * w, _lib.ptr(mask), _lib.ptr(self.P), self.N,
                                         _lib.ptr(self.proto_class), c_longlong(base), c_int(self.prec),
                                         _lib.ptr(self.packed), _lib.stream_ptr())
        _lib.check(rc, "adlm_push_scan")

    def gather(self, z, first_image_index, out):
        """fill out[j] (N,K) with the winner vector of every prototype whose winner lies in z's images"""
        z = z.contiguous()
        n, K, h, w = z.shape
        base = int(first_image_index) * h * w
        with torch.cuda.device(self.device):
            rc = self.lib.adlm_push_gather(_lib.ptr(z), n, K, h * w, c_longlong(base), _lib.ptr(self.packed),
                                           self.N, _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "adlm_push_gather")

    def reduce(self, group=None):
        """multi-GPU merge: one all_reduce(MIN) over the N packed words (NCCL over NVLink)"""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.packed, op=dist.ReduceOp.MIN, group=group)

    def sync_map_size(self, group=None):
        """multi-GPU: a rank whose image range is empty (more ranks than images) never scanned and does not know the
        feature-map size the global patch ids are built on; take it from a rank that did (one MAX all-reduce)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            t = torch.tensor(list(self.hw or (0, 0)), device=self.device, dtype=torch.int64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            hw = (int(t[0]), int(t[1]))
            if hw != (0, 0):
                self.hw = hw

    def winners(self):
        """-> (min distance float64 (N,), image index (N,), patch_i (N,), patch_j (N,)); -1 where none"""
        d, ids = unpack(self.packed.cpu().numpy())
        if self.hw is None:                      # nothing was scanned anywhere: no winners (push.py:57-61)
            none = np.full(self.N, -1, dtype=np.int64)
            return d, none, none.copy(), none.copy()
        h, w = self.hw
        img = np.where(ids >= 0, ids // (h * w), -1)
        patch = np.where(ids >= 0, ids % (h * w), -1)
        return d, img, np.where(ids >= 0, patch // w, -1), np.where(ids >= 0, patch % w, -1)


def push_search_tensors(z, labels, prototypes, proto_class, precision=None, chunk=64, group=None,
                        image_offset=0):
    """Search a resident shard: z (n,K,h,w) CUDA features, labels (n,H,W) CUDA uint8.
    Returns (search, new_prototypes (N,K) float32 CUDA) after the cross-rank reduce; rows of prototypes
    with no candidate stay zero (push.py:57-61)."""
    s = PushSearch(prototypes, proto_class, precision=precision)
    n = z.shape[0]
    for i in range(0, n, chunk):
        s.scan(z[i:i + chunk], labels[i:i + chunk], image_offset + i)
    s.reduce(group)
    out = torch.zeros((s.N, s.K), device=s.device, dtype=torch.float32)
    for i in range(0, n, chunk):
        s.gather(z[i:i + chunk], image_offset + i, out)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)   # each row is owned by exactly one rank
    return s, out


# ------------------------------------------------------------------------------ boxes of the winners
def grow_high_activation_box(act, box, threshold, margin=5):
    """Greedy window growth around `box` over pixels with act >= threshold, then a fixed margin
    (helpers.py:48-82 behaviour): each side advances one pixel per round while the strip just outside
    it holds a high-activation pixel."""
    hi = act >= threshold
    top, bottom, left, right = box
    Hh, Ww = act.shape
    live = [True] * 4
    while any(live):
        if live[0] and top > 0 and hi[top - 1, left:right + 1].any():
            top -= 1
        else:
            live[0] = False
        if live[1] and bottom < Hh - 1 and hi[bottom + 1, left:right + 1].any():
            bottom += 1
        else:
            live[1] = False
        if live[2] and left > 0 and hi[top:bottom + 1, left - 1].any():
            left -= 1
        else:
            live[2] = False
        if live[3] and right < Ww - 1 and hi[top:bottom + 1, right + 1].any():
            right += 1
        else:
            live[3] = False
    return (max(top - margin, 0), min(bottom + margin, Hh - 1) + 1,
            max(left - margin, 0), min(right + margin, Ww - 1) + 1)


def activation_bound_box(dist_map, labels, target_class, box, activation='log', eps=1e-4, max_dist=None):
    """push.py:310-330: similarity map of the winning prototype, cubic upsample to the image, 95th
    percentile threshold, growth restricted to ground-truth pixels of the prototype's class."""
    import cv2
    if activation == 'log':
        act = np.log((dist_map + 1) / (dist_map + eps))
    elif activation == 'linear':
        act = max_dist - dist_map
    else:
        act = activation(dist_map)
    H, W = labels.shape
    up = cv2.resize(act.astype(np.float32), dsize=(W, H), interpolation=cv2.INTER_CUBIC)
    threshold = np.percentile(up, 95)
    gt = labels == (target_class + 1)
    return grow_high_activation_box(up * gt, box, threshold)


# ------------------------------------------------------------------------------ the reference entry point
def _to_normalized_tensor(img):
    arr = np.asarray(img, dtype=np.float32) / 255.0          # transforms.ToTensor
    t = torch.from_numpy(arr).permute(2, 0, 1)
    mean = torch.tensor(NORM_MEAN).view(3, 1, 1)
    std = torch.tensor(NORM_STD).view(3, 1, 1)
    return (t - mean) / std


def _load_sample(dataset, img_id):
    from PIL import Image
    with open(dataset.get_img_path(img_id), 'rb') as f:
        img = Image.open(f).convert('RGB')
    m = dataset.image_margin_size                             # push.py:110-113
    img = img.crop((m, m, img.width - m, img.height - m))
    ann = np.load(os.path.join(dataset.annotations_dir, img_id + '.npy'))
    if getattr(dataset, 'convert_targets', None) is not None:
        ann = dataset.convert_targets(ann)
    return img, np.asarray(ann)


def push_prototypes(dataset,
                    prototype_network_parallel,
                    prototype_layer_stride=1,
                    root_dir_for_saving_prototypes=None,
                    epoch_number=None,
                    prototype_img_filename_prefix=None,
                    prototype_self_act_filename_prefix=None,
                    proto_bound_boxes_filename_prefix=None,
                    save_prototype_class_identity=True,
                    log=print,
                    pascal=False,
                    prototype_activation_function_in_numpy=None):
    """Drop-in for segmentation.push.push_prototypes (push.py:24-158).  Under torch.distributed every rank
    scans its contiguous slice of `dataset.img_ids` and the winners are merged with one MIN
    all-reduce; files are written by rank 0."""
    import torch.distributed as dist
    ppnet = prototype_network_parallel.module if hasattr(prototype_network_parallel, 'module') \
        else prototype_network_parallel
    ppnet.eval()
    log('\tpush')
    start = time.time()
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0

    N = ppnet.num_prototypes
    n_cols = 6 if save_prototype_class_identity else 5
    proto_rf_boxes = np.full([N, n_cols], -1)
    proto_bound_boxes = np.full([N, n_cols], -1)
    if root_dir_for_saving_prototypes is not None and epoch_number is not None:
        proto_epoch_dir = os.path.join(root_dir_for_saving_prototypes, 'epoch-' + str(epoch_number))
        os.makedirs(proto_epoch_dir, exist_ok=True)
    else:
        proto_epoch_dir = root_dir_for_saving_prototypes

    proto_class = ppnet.proto_class()
    # The winners must be the fp32 arg-min (push.py:259-265): a model trained with head_precision 'tf32' still
    # searches with fp32-class distances, exactly as model._forward_min_pool does.
    prec = ppnet.head_precision if ppnet.head_precision in ("fp32", "tf32x3") else None
    search = PushSearch(ppnet.prototype_vectors, proto_class, precision=prec)
    dev = search.device
    img_ids = list(dataset.img_ids)
    per = (len(img_ids) + world - 1) // world
    lo, hi = rank * per, min(len(img_ids), (rank + 1) * per)   # contiguous ranges keep "earliest image wins"

    log('Updating prototypes...')
    img_hw = {}
    with torch.no_grad():
        for idx in range(lo, hi):
            img, ann = _load_sample(dataset, img_ids[idx])
            z = ppnet.conv_features(_to_normalized_tensor(img).unsqueeze(0).to(dev))
            labels = torch.from_numpy(np.ascontiguousarray(ann).astype(np.uint8)).unsqueeze(0).to(dev)
            search.scan(z.float(), labels, idx)
            img_hw[idx] = ann.shape
        search.reduce()
        search.sync_map_size()
        dmin, w_img, w_pi, w_pj = search.winners()
        new_protos = torch.zeros((N, search.K), device=dev, dtype=torch.float32)
        pc = proto_class.cpu().numpy()
        # second visit, winners only: feature vector, receptive-field box, activation-crop box
        for idx in sorted(set(int(i) for i in w_img if lo <= i < hi)):
            img, ann = _load_sample(dataset, img_ids[idx])
            z = ppnet.conv_features(_to_normalized_tensor(img).unsqueeze(0).to(dev)).float()
            search.gather(z, idx, new_protos)
            h, w = z.shape[2], z.shape[3]
            H, W = ann.shape
            dist_map = None
            for j in np.nonzero(w_img == idx)[0]:
                box = rf_box(int(w_pi[j]), int(w_pj[j]), H, W, h, w)
                proto_rf_boxes[j, :5] = (idx,) + box
                if n_cols == 6:
                    proto_rf_boxes[j, 5] = pc[j]
                try:
                    if dist_map is None:
                        dist_map = ppnet._l2_convolution(z)[0].cpu().numpy()
                    act_fn = ppnet.prototype_activation_function
                    if not isinstance(act_fn, str):
                        act_fn = prototype_activation_function_in_numpy
                    bb = activation_bound_box(dist_map[j], ann, int(pc[j]), box, act_fn, ppnet.epsilon,
                                              max_dist=float(np.prod(ppnet.prototype_shape[1:])))
                    proto_bound_boxes[j, :5] = (idx,) + bb
                    if n_cols == 6:
                        proto_bound_boxes[j, 5] = pc[j]
                except ImportError:      # cv2 missing: boxes stay -1, the search result is unaffected
                    pass
        if world > 1:
            dist.all_reduce(new_protos, op=dist.ReduceOp.SUM)
            for arr in (proto_rf_boxes, proto_bound_boxes):
                t = torch.from_numpy(arr.astype(np.int64)).to(dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)      # unowned rows are -1 everywhere else
                arr[...] = t.cpu().numpy()

    if proto_epoch_dir is not None and proto_bound_boxes_filename_prefix is not None and rank == 0:
        np.save(os.path.join(proto_epoch_dir, proto_bound_boxes_filename_prefix + '-receptive_field'
                             + str(epoch_number) + '.npy'), proto_rf_boxes)
        np.save(os.path.join(proto_epoch_dir, proto_bound_boxes_filename_prefix + str(epoch_number) + '.npy'),
                proto_bound_boxes)

    log('\tExecuting push ...')
    ppnet.prototype_vectors.data.copy_(new_protos.reshape(tuple(ppnet.prototype_shape)))
    unique_index, duplicate_idx = dedup_prototypes(new_protos.cpu().numpy().astype(np.float64))
    log(f'Removing {len(duplicate_idx)} duplicate prototypes.')
    ppnet.prune_prototypes(duplicate_idx)
    if rank == 0 and root_dir_for_saving_prototypes is not None:
        os.makedirs(root_dir_for_saving_prototypes, exist_ok=True)
        with open(os.path.join(root_dir_for_saving_prototypes, 'unique_prototypes.json'), 'w') as fp:
            json.dump([int(i) for i in sorted(unique_index)], fp)
    log('\tpush time: \t{0}'.format(time.time() - start))


def dedup_prototypes(prototype_update):
    """np.unique(axis=0) de-dup of push.py:147-149 -> (unique_index, duplicate_idx list)"""
    flat = np.reshape(prototype_update, (prototype_update.shape[0], -1))
    _, unique_index = np.unique(flat, axis=0, return_index=True)
    keep = set(int(i) for i in unique_index)
    return unique_index, [i for i in range(flat.shape[0]) if i not in keep]
